// matfree.cuh -- matrix-free product with the KLE operator K.
//
// y = K x without the assembled matrix: per element, gather x, evaluate the reference
// gradients at the quadrature points by sum factorisation (1-D tables, three passes), apply
// the point-wise bilinear form of Spectral.getElemKLEMatrices (spectral.py:116-159), apply
// the transposed passes, and add the element results per node in ascending element order.
// Bytes per product: vectors + 80 B of geometry per quadrature point instead of 8 B per
// stored non-zero (0.03 GB instead of 0.90 GB at 16^3 / ngl = 4); same result up to the
// order of the floating-point sums.
#pragma once
#include "spmv.cuh"

struct pyn_matfree {
  pyn_mesh* mesh = nullptr;
  int npf = 0, npr = 0;                  // points per direction: full / reduced rule
  int64_t npts = 0;                      // npf^3 + npr^3
  pyn::DevBuf<double> geo;               // [nel][10][npts]: J^-1 (k*3+d), w|J|; points in tensor order
  pyn::DevBuf<int32_t> nodes_t;          // [nel][nn] element nodes in tensor order (i fastest)
  pyn::DevBuf<int32_t> loc_of_t;         // [nn] reference local index of tensor position t
  pyn::DevBuf<double> ye;                // [nel][nn][3] element results, reference local order
};

namespace pyn {
int matfree_create(pyn_mesh* mesh, pyn_matfree** out);
// y = alpha*K*x + beta*z on rows [0, nrows) (nodes); dot: also fuse.dot_out = sum y_i x_i
int matfree_apply(pyn_mat* m, int64_t nrows, double alpha, const double* x, double beta,
                  const double* z, double* y, bool dot, const SpmvFuse& fuse);
}  // namespace pyn
