// asm_pc.cuh -- element-block additive Schwarz preconditioner for the KLE operator (-pc_type asm)
#pragma once
#include <vector>
#include "mesh.cuh"

namespace pyn {

// z = sum_e R_e^T W_e inv(R_e K R_e^T) W_e R_e r, W = 1/sqrt(elements sharing the node).
// Elements whose assembled blocks R_e K R_e^T are bitwise equal share one inverse (27 on a
// uniform box with one boundary-condition kind per side).
struct AsmPc {
  pyn_mesh* mesh = nullptr;
  int B = 0, ld = 0;               // dofs per node, dofs per element (nn * B)
  int64_t nclass = 0, ntiles = 0;
  DevBuf<double> inv;              // [nclass][ld][ld], symmetric (or the transposed inverse of an unsymmetric block)
  DevBuf<int32_t> tile_elems;      // [ntiles][kAsmTile (asm_kernels.cuh)] elements of one class, -1 padded
  DevBuf<int32_t> tile_class;      // [ntiles]
  DevBuf<double> wnode;            // [nnodes]
  // row-partitioned operator: the apply refreshes the ghost entries of r, multiplies over all local
  // elements (block + halo layer) and sums into the owned nodes only -- no reverse exchange
  pyn_halo* halo = nullptr;
  int64_t n_own = 0;               // nodes summed into (all of them on one GPU)
  int64_t ncomplete = 0;           // elements whose block could be read here (all rows local)
  double hash_quantum = 0.0;       // values were grouped as multiples of this (0: bitwise equal blocks)
  DevBuf<double> ybuf;             // [nel][ld] element-local results of the last apply
  size_t smem = 0;
  // second level (coarse_steps > 0): Q1 space on the element corners, Galerkin operator
  // Ac = P^T K P probed through K's own product, coarse_steps Chebyshev steps with point Jacobi
  int coarse_steps = 0;
  int64_t nv = 0;
  int ncorner = 0, ncolors = 0;
  double lmax = 0.0;
  DevBuf<int32_t> elem_verts;      // [nel][ncorner]
  DevBuf<double> phi, keep;        // [nn][ncorner], [nnodes*B]
  DevBuf<int64_t> v2e_ptr;
  DevBuf<int2> v2e;
  DevBuf<int64_t> ac_ptr;          // block CSR of Ac over the vertex adjacency
  DevBuf<int32_t> ac_col;
  DevBuf<double> ac_val, ac_dinv;
  DevBuf<double> rc, xc, d0, d1;
  // restriction as explicit lists (fused coarse kernel): vertex v gathers w * keep * r over
  // rl_node / rl_w [rl_ptr[v], rl_ptr[v+1]) -- the entries asm_restrict_kernel visits, in its order
  DevBuf<int64_t> rl_ptr;
  DevBuf<int32_t> rl_node;
  DevBuf<double> rl_w;
  DevBuf<double> cheb_c;           // [2][steps+1] Chebyshev coefficients c1, c2 on the device
  int coarse_grid = 0;             // co-resident grid of the fused kernel (cooperative launch)
  std::vector<double> c1, c2;      // Chebyshev coefficients (asm_host.cuh)
};

// more distinct element blocks than this (bytes of inverses) is refused: the apply would read
// more than the SpMV it is meant to save
constexpr int64_t kAsmMaxInverseBytes = (int64_t)1 << 30;
constexpr double kAsmHashQuantum = 1e-7;   // second grouping attempt: values as multiples of this x max |K|

// coarse_steps: 0 = one level; k > 0 = plus the coarse correction with k Chebyshev steps
int asm_setup(pyn_ctx* ctx, pyn_mat* A, int coarse_steps, AsmPc** out);
// z = M^-1 r.  dots != NULL: dots[0] = r.z, dots[1] = z.z (device).  done != NULL: a device flag;
// when non-zero the apply is skipped (iterations launched after convergence).
// Row-partitioned: r's ghost entries are refreshed in place (they are copies of the owners' values
// by definition), dots are this rank's partial sums.
int asm_apply(pyn_ctx* ctx, AsmPc* pc, double* r, double* z, double* dots, const double* done);

}  // namespace pyn
