// spmv.cuh -- SpMV and Mat helpers shared with the KSP.
#pragma once
#include "mesh.cuh"
#include "comm.cuh"

namespace pyn {
// Optional fusions for the Krylov loop: `skip` (device flag, non-zero = solver
// already converged -> return at once) and DOT (also reduce sum_i y_i*x_i, the
// p.Ap of CG, into dot_out with the deterministic grid reduction).
struct SpmvFuse {
  const double* skip;
  double* partials;
  unsigned int* ticket;
  double* dot_out;
  P2PRed red;          // nranks > 1: push the dot partial to every rank's window instead
  // single-reduction CG: with dot_r the reduction has three values, {y.x, dot_r.x, x.x}
  const double* dot_r;
  // hw.win != NULL: the ghost entries of x were pushed into this rank's halo window by the
  // kernel that produced x; wait for them at CTA start and read ghost columns from the window
  HaloWait hw;
  // with hw and n_int > 0: the first n_int rows read no ghost column; a warp waits for the ghosts only
  // when it reaches its first row >= n_int, so the arrival of the halo overlaps the interior rows
  int64_t n_int;
};

// y = alpha*A*x + beta*z (z may be NULL when beta == 0; z may alias y)
int spmv(pyn_mat* m, double alpha, const double* x, double beta, const double* z, double* y);
// red != NULL with nranks > 1: the dot partial is pushed to every rank's window (kind 0) instead of dot_out
int spmv_dot(pyn_mat* m, const double* x, double* y, double* dot_out, const double* skip,
             const P2PRed* red = nullptr, const HaloWait* hw = nullptr);
// y = A x and out3 = {x.y, r.x, x.x} in one pass (single-reduction CG)
int spmv_dot3(pyn_mat* m, const double* x, double* y, const double* r, double* out3, const double* skip,
              const P2PRed* red = nullptr);
// row-partitioned matrix: count the leading owned rows without a ghost column (n_interior) and the
// blocks of the owned rows
int spmv_build_row_order(pyn_mat* m);
int diag_blocks(pyn_mat* m, double* dblk);   // [nbrows][b][b]
int scale_rows(pyn_mat* m, const double* l);
int mat_add_new(pyn_mat* a, pyn_mat* b, pyn_mat** out);

}  // namespace pyn
