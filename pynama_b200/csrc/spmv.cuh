// spmv.cuh -- SpMV and Mat helpers shared with the KSP.
#pragma once
#include "mesh.cuh"

namespace pyn {
// y = alpha*A*x + beta*z (z may be NULL when beta == 0; z may alias y)
int spmv(pyn_mat* m, double alpha, const double* x, double beta, const double* z, double* y);
int spmv_dot(pyn_mat* m, const double* x, double* y, double* dot_out, const double* skip);
int diag_blocks(pyn_mat* m, double* dblk);   // [nbrows][b][b]
int scale_rows(pyn_mat* m, const double* l);
int mat_add_new(pyn_mat* a, pyn_mat* b, pyn_mat** out);

}  // namespace pyn
