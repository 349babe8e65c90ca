// ibm.cu -- 2-D immersed boundary: the interpolation matrix H, the spreading matrix S and
// A = H S of ImmersedBoundaryStatic.buildIBMMatrix (src/cases/immersed_boundary.py:170-193),
// built on the device from the Lagrangian markers and the Euler nodes of the affected cells.
//
//   collectNodes / computeClose   immersed_boundary.py:210-229, :263-274 (static), :346-358 (dynamic)
//   computeDirac / getDiracs      immersed_boundary.py:255-262, src/domain/immersed_body.py:261-266
//   fourGrid / threeGrid / linear src/domain/immersed_body.py:398-424
//
// Unscaled D[l][n] = prod_axis phi(|x_n - X_l| / h) / h over the kept nodes n of marker l
// (explicit zeros included, as MatSetValues stores them); H = h^2 D, S = (dl h) D^T, A = H S
// with the pattern of PETSc's MatMatMult (markers sharing at least one kept node).  The three
// matrices are node-block matrices with 2x2 blocks w*I (row l*2+dof, column n*2+dof), so the
// SpMV and the Krylov solver of the KLE path apply to them unchanged.
//
// The arithmetic uses the _rn intrinsics in the reference's order of operations: no FMA
// contraction, so the weights equal the Python floats of the reference.
#include "mesh.cuh"

namespace pyn {

__device__ __forceinline__ double phi_four(double r) {
  if (r <= 1.0) {
    const double t = __dsub_rn(__dadd_rn(1.0, __dmul_rn(4.0, r)), __dmul_rn(4.0, __dmul_rn(r, r)));
    return __ddiv_rn(__dadd_rn(__dsub_rn(3.0, __dmul_rn(2.0, r)), __dsqrt_rn(t)), 8.0);
  }
  if (r <= 2.0) {
    const double t = __dsub_rn(__dadd_rn(-7.0, __dmul_rn(12.0, r)), __dmul_rn(4.0, __dmul_rn(r, r)));
    return __ddiv_rn(__dsub_rn(__dsub_rn(5.0, __dmul_rn(2.0, r)), __dsqrt_rn(t)), 8.0);
  }
  return 0.0;
}
__device__ __forceinline__ double phi_three(double r) {
  if (r <= 0.5) {
    const double t = __dadd_rn(__dmul_rn(-3.0, __dmul_rn(r, r)), 1.0);
    return __ddiv_rn(__dadd_rn(1.0, __dsqrt_rn(t)), 3.0);
  }
  if (r <= 1.5) {
    const double u = __dsub_rn(1.0, r);
    const double t = __dadd_rn(__dmul_rn(-3.0, __dmul_rn(u, u)), 1.0);
    return __ddiv_rn(__dsub_rn(__dsub_rn(5.0, __dmul_rn(3.0, r)), __dsqrt_rn(t)), 6.0);
  }
  return 0.0;
}
__device__ __forceinline__ double phi_linear(double r) { return r < 1.0 ? __dsub_rn(1.0, r) : 0.0; }

__device__ __forceinline__ double phi(int kind, double r) {
  return kind == 0 ? phi_four(r) : (kind == 1 ? phi_three(r) : phi_linear(r));
}

// ImmersedBody.getDiracs: d = 1; for r in dist: d *= phi(|r|/h); d /= h
__device__ __forceinline__ double dirac_weight(int kind, double dx, double dy, double h) {
  double d = phi(kind, __ddiv_rn(fabs(dx), h));
  d = __ddiv_rn(d, h);
  d = __dmul_rn(d, phi(kind, __ddiv_rn(fabs(dy), h)));
  return __ddiv_rn(d, h);
}

// computeClose: is Euler node (ex, ey) kept for the marker (lx, ly)?
__device__ __forceinline__ bool is_close(bool dynamic, double ex, double ey, double lx, double ly, double dl) {
  const double d0 = __ddiv_rn(__dsub_rn(ex, lx), dl), d1 = __ddiv_rn(__dsub_rn(ey, ly), dl);
  if (dynamic) return fabs(d0) < 2.0 && fabs(__dmul_rn(d1, dl)) < __dadd_rn(1.0, __dmul_rn(2.0, dl));
  return fabs(d0) < 3.0 && fabs(d1) < 3.0;
}

// 2x2 block w*I of a row-plane matrix
__device__ __forceinline__ void store_block(double* val, int64_t p0, int64_t nb, int64_t u, double v) {
  val[p0 * 4 + 0 * nb + u] = v;
  val[p0 * 4 + 1 * nb + u] = 0.0;
  val[p0 * 4 + 2 * nb + u] = 0.0;
  val[p0 * 4 + 3 * nb + u] = v;
}

// One warp per "row" point scanning all "column" points in their given order (ascending node id /
// marker id), kept ones compacted in order with ballots.  ROWS_ARE_MARKERS: rows of D (markers x
// candidate nodes) else rows of D^T (candidate nodes x markers; row index = the node id).
// PASS 0 counts, PASS 1 writes column ids, the unscaled weights (D only) and the scaled blocks.
template <int PASS, bool ROWS_ARE_MARKERS>
__global__ void ibm_scan_kernel(int64_t nrow, const double* __restrict__ row_xy,
                                const int32_t* __restrict__ row_id, int64_t ncol,
                                const double* __restrict__ col_xy, const int32_t* __restrict__ col_id,
                                double h, double dl, int kind, int dynamic, double scale,
                                int32_t* __restrict__ rowlen, const int64_t* __restrict__ rowptr,
                                int32_t* __restrict__ colidx, double* __restrict__ w,
                                double* __restrict__ val) {
  const int lane = threadIdx.x & 31;
  const int64_t warp = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
  for (int64_t i = warp; i < nrow; i += nwarps) {
    const double rx = row_xy[2 * i], ry = row_xy[2 * i + 1];
    const int64_t r = row_id ? (int64_t)row_id[i] : i;
    const int64_t p0 = PASS ? rowptr[r] : 0;
    const int64_t nb = PASS ? rowptr[r + 1] - p0 : 0;
    int64_t out = 0;
    for (int64_t j0 = 0; j0 < ncol; j0 += 32) {
      const int64_t j = j0 + lane;
      bool keep = false;
      double cx = 0.0, cy = 0.0;
      if (j < ncol) {
        cx = col_xy[2 * j];
        cy = col_xy[2 * j + 1];
        // the test is the reference's, marker = "lag", node = "euler", whichever is the row
        keep = ROWS_ARE_MARKERS ? is_close(dynamic != 0, cx, cy, rx, ry, dl)
                                : is_close(dynamic != 0, rx, ry, cx, cy, dl);
      }
      const unsigned bal = __ballot_sync(0xffffffffu, keep);
      if (PASS && keep) {
        const int64_t u = out + __popc(bal & ((1u << lane) - 1u));
        const double d = ROWS_ARE_MARKERS ? dirac_weight(kind, __dsub_rn(cx, rx), __dsub_rn(cy, ry), h)
                                          : dirac_weight(kind, __dsub_rn(rx, cx), __dsub_rn(ry, cy), h);
        colidx[p0 + u] = col_id ? col_id[j] : (int32_t)j;
        if (w) w[p0 + u] = d;
        store_block(val, p0, nb, u, __dmul_rn(d, scale));
      }
      out += __popc(bal);
    }
    if (!PASS && lane == 0) rowlen[r] = (int32_t)out;
  }
}

// A = H S: block (l, m) = sum over the common kept nodes n, ascending, of H[l][n] S[n][m].
// One warp per marker l, lanes over markers m; both rows of D are sorted by node id.
template <int PASS>
__global__ void ibm_hs_kernel(int64_t nlag, const double* __restrict__ lag_xy, double dl, int dynamic,
                              const int64_t* __restrict__ d_rowptr, const int32_t* __restrict__ d_col,
                              const double* __restrict__ d_w, double scale_h, double scale_s,
                              int32_t* __restrict__ rowlen, const int64_t* __restrict__ rowptr,
                              int32_t* __restrict__ colidx, double* __restrict__ val) {
  const int lane = threadIdx.x & 31;
  const int64_t warp = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
  // two markers further apart than this along an axis cannot share a kept node
  const double reach = dynamic ? fmax(4.0 * dl, 2.0 * (1.0 + 2.0 * dl)) : 6.0 * dl;
  for (int64_t l = warp; l < nlag; l += nwarps) {
    const double lx = lag_xy[2 * l], ly = lag_xy[2 * l + 1];
    const int64_t a0 = d_rowptr[l], a1 = d_rowptr[l + 1];
    const int64_t p0 = PASS ? rowptr[l] : 0;
    const int64_t nb = PASS ? rowptr[l + 1] - p0 : 0;
    int64_t out = 0;
    for (int64_t m0 = 0; m0 < nlag; m0 += 32) {
      const int64_t m = m0 + lane;
      bool common = false;
      double acc = 0.0;
      if (m < nlag && fabs(lag_xy[2 * m] - lx) <= reach * 1.0000001 && fabs(lag_xy[2 * m + 1] - ly) <= reach * 1.0000001) {
        int64_t a = a0, b = d_rowptr[m];
        const int64_t b1 = d_rowptr[m + 1];
        while (a < a1 && b < b1) {
          const int32_t ca = d_col[a], cb = d_col[b];
          if (ca == cb) {
            common = true;
            if (PASS) acc = __dadd_rn(acc, __dmul_rn(__dmul_rn(d_w[a], scale_h), __dmul_rn(d_w[b], scale_s)));
            ++a;
            ++b;
          } else if (ca < cb) {
            ++a;
          } else {
            ++b;
          }
        }
      }
      const unsigned bal = __ballot_sync(0xffffffffu, common);
      if (PASS && common) {
        const int64_t u = out + __popc(bal & ((1u << lane) - 1u));
        colidx[p0 + u] = (int32_t)m;
        store_block(val, p0, nb, u, acc);
      }
      out += __popc(bal);
    }
    if (!PASS && lane == 0) rowlen[l] = (int32_t)out;
  }
}

__global__ void ibm_affected_kernel(int64_t n, const int32_t* __restrict__ col, const double* __restrict__ w,
                                    uint8_t* __restrict__ mark) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
    if (w[i] > 0.0) mark[col[i]] = 1;
}

static pyn_mat* new_free_mat(pyn_ctx* ctx, int64_t nbrows, int64_t nbcols) {
  pyn_mat* m = new pyn_mat();
  m->ctx = ctx;
  m->mesh = nullptr;
  m->graph = nullptr;
  m->br = m->bc = 2;
  m->nbrows = nbrows;
  m->nbcols = nbcols;
  m->free_pattern = true;
  m->diag_only = true;
  m->ref_dense = false;
  return m;
}

static int alloc_blocks(pyn_mat* m, const int32_t* rowlen) {
  PYN_CHECK(m->rowptr.alloc(m->ctx, m->nbrows + 1));
  PYN_CHECK(exclusive_scan_i32_to_i64(m->ctx, rowlen, m->rowptr.p, m->nbrows, &m->nblocks));
  PYN_CHECK(m->colidx.alloc(m->ctx, m->nblocks));
  PYN_CHECK(m->val.alloc(m->ctx, m->nblocks * 4));
  return PYN_OK;
}

static int ibm_build(pyn_ctx* ctx, int64_t n_euler, int64_t nlag, const double* lag_xy_h, int64_t ncand,
                     const int32_t* cand_nodes_h, const double* cand_xy_h, double h, double dl,
                     double scale_h, double scale_s, int kind, int dynamic, pyn_mat* H, pyn_mat* S,
                     pyn_mat* A, int64_t* max_per_lag, uint8_t* affected_h) {
  DevBuf<double> lag_xy, cand_xy, w;
  DevBuf<int32_t> cand_nodes, rowlen;
  PYN_CHECK(lag_xy.alloc(ctx, 2 * nlag));
  PYN_CHECK(lag_xy.upload(lag_xy_h, 2 * nlag));
  PYN_CHECK(cand_xy.alloc(ctx, 2 * ncand));
  PYN_CHECK(cand_xy.upload(cand_xy_h, 2 * ncand));
  PYN_CHECK(cand_nodes.alloc(ctx, ncand));
  PYN_CHECK(cand_nodes.upload(cand_nodes_h, ncand));
  cudaStream_t st = ctx->stream;

  // H = scale_h * D: rows = markers, columns = candidate nodes in ascending node id
  PYN_CHECK(rowlen.alloc(ctx, nlag));
  const int gl = grid_for(ctx, nlag * 32, 128);
  ibm_scan_kernel<0, true><<<gl, 128, 0, st>>>(nlag, lag_xy.p, nullptr, ncand, cand_xy.p, cand_nodes.p, h, dl,
                                              kind, dynamic, scale_h, rowlen.p, nullptr, nullptr, nullptr,
                                              nullptr);
  PYN_LAUNCHED(ctx);
  std::vector<int32_t> len_h((size_t)nlag);
  PYN_CHECK(rowlen.download(len_h.data(), nlag));
  int64_t mx = 0;
  for (int64_t l = 0; l < nlag; ++l) {
    // collectNodes raises "Lag Node without Euler" (immersed_boundary.py:222-223)
    PYN_REQUIRE(len_h[(size_t)l] > 0, "Lag Node without Euler (marker %lld)", (long long)l);
    mx = std::max<int64_t>(mx, len_h[(size_t)l]);
  }
  if (max_per_lag) *max_per_lag = mx;
  PYN_CHECK(alloc_blocks(H, rowlen.p));
  PYN_CHECK(w.alloc(ctx, H->nblocks));
  ibm_scan_kernel<1, true><<<gl, 128, 0, st>>>(nlag, lag_xy.p, nullptr, ncand, cand_xy.p, cand_nodes.p, h, dl,
                                              kind, dynamic, scale_h, nullptr, H->rowptr.p, H->colidx.p, w.p,
                                              H->val.p);
  PYN_LAUNCHED(ctx);

  // S = scale_s * D^T: rows = all Euler nodes (most of them empty), columns = markers
  DevBuf<int32_t> lenS;
  PYN_CHECK(lenS.alloc(ctx, n_euler));
  PYN_CHECK(lenS.zero());
  const int gc = grid_for(ctx, ncand * 32, 128);
  ibm_scan_kernel<0, false><<<gc, 128, 0, st>>>(ncand, cand_xy.p, cand_nodes.p, nlag, lag_xy.p, nullptr, h, dl,
                                               kind, dynamic, scale_s, lenS.p, nullptr, nullptr, nullptr,
                                               nullptr);
  PYN_LAUNCHED(ctx);
  PYN_CHECK(alloc_blocks(S, lenS.p));
  PYN_REQUIRE(S->nblocks == H->nblocks, "IBM transpose: %lld blocks against %lld", (long long)S->nblocks,
              (long long)H->nblocks);
  ibm_scan_kernel<1, false><<<gc, 128, 0, st>>>(ncand, cand_xy.p, cand_nodes.p, nlag, lag_xy.p, nullptr, h, dl,
                                               kind, dynamic, scale_s, nullptr, S->rowptr.p, S->colidx.p,
                                               nullptr, S->val.p);
  PYN_LAUNCHED(ctx);

  // A = H S (H.matMult(S), immersed_boundary.py:188)
  ibm_hs_kernel<0><<<gl, 128, 0, st>>>(nlag, lag_xy.p, dl, dynamic, H->rowptr.p, H->colidx.p, w.p, scale_h,
                                       scale_s, rowlen.p, nullptr, nullptr, nullptr);
  PYN_LAUNCHED(ctx);
  PYN_CHECK(alloc_blocks(A, rowlen.p));
  ibm_hs_kernel<1><<<gl, 128, 0, st>>>(nlag, lag_xy.p, dl, dynamic, H->rowptr.p, H->colidx.p, w.p, scale_h,
                                       scale_s, nullptr, A->rowptr.p, A->colidx.p, A->val.p);
  PYN_LAUNCHED(ctx);

  if (affected_h) {   // nodes with a positive weight (buildIBMMatrix's return value, :181,:193)
    DevBuf<uint8_t> mark;
    PYN_CHECK(mark.alloc(ctx, n_euler));
    PYN_CHECK(mark.zero());
    ibm_affected_kernel<<<grid_for(ctx, H->nblocks, 256), 256, 0, st>>>(H->nblocks, H->colidx.p, w.p, mark.p);
    PYN_LAUNCHED(ctx);
    PYN_CHECK(mark.download(affected_h, n_euler));
  }
  PYN_CUDA(cudaStreamSynchronize(st));
  return PYN_OK;
}

}  // namespace pyn

using namespace pyn;

extern "C" int pyn_ibm_build(pyn_ctx* ctx, int64_t n_euler_nodes, int64_t nlag, const double* lag_xy,
                             int64_t ncand, const int32_t* cand_nodes, const double* cand_xy, double h,
                             double dl, double scale_h, double scale_s, int kind, int dynamic,
                             pyn_mat** H, pyn_mat** S, pyn_mat** A, int64_t* max_per_lag,
                             uint8_t* affected) {
  PYN_REQUIRE(ctx && lag_xy && cand_nodes && cand_xy && H && S && A, "pyn_ibm_build: null argument");
  PYN_REQUIRE(nlag > 0 && ncand > 0 && n_euler_nodes > 0, "pyn_ibm_build: empty body or node set");
  PYN_REQUIRE(h > 0.0 && dl > 0.0, "pyn_ibm_build: h and dl must be positive");
  PYN_REQUIRE(kind >= 0 && kind <= 2, "pyn_ibm_build: delta function %d (0 four, 1 three, 2 linear)", kind);
  for (int64_t j = 0; j < ncand; ++j) {
    PYN_REQUIRE(cand_nodes[j] >= 0 && cand_nodes[j] < n_euler_nodes, "pyn_ibm_build: node %d outside the mesh",
                cand_nodes[j]);
    PYN_REQUIRE(j == 0 || cand_nodes[j] > cand_nodes[j - 1], "pyn_ibm_build: candidate nodes must ascend");
  }
  *H = *S = *A = nullptr;
  pyn_mat* h_ = new_free_mat(ctx, nlag, n_euler_nodes);
  pyn_mat* s_ = new_free_mat(ctx, n_euler_nodes, nlag);
  pyn_mat* a_ = new_free_mat(ctx, nlag, nlag);
  const int rc = ibm_build(ctx, n_euler_nodes, nlag, lag_xy, ncand, cand_nodes, cand_xy, h, dl, scale_h,
                           scale_s, kind, dynamic, h_, s_, a_, max_per_lag, affected);
  if (rc != PYN_OK) {
    delete h_;
    delete s_;
    delete a_;
    return rc;
  }
  *H = h_;
  *S = s_;
  *A = a_;
  return PYN_OK;
}
