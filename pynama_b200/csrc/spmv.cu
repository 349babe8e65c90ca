// spmv.cu -- node-block SpMV (k6) and the small Mat operations around it.
//
// y = alpha*A*x + beta*z on the row-plane BSR layout: a warp (or a sub-warp
// group of G lanes for short rows) owns one block row; lane l takes blocks
// l, l+G, ... and reads the same position of each of the br*bc planes, so
// every plane load of the group is one contiguous run of 8-byte words.
// Replaces PETSc MatMult on AIJ (kle_solver.py:35-41, base_problem.py:123-136).
// HBM-bound: per block 8*br*bc bytes of values + 4 bytes of column index; x
// (a few MB) lives in L2.
#include "elem.cuh"
#include "spmv.cuh"
#include "vec.cuh"
#include "comm.cuh"
#include "matfree.cuh"
#include <stdlib.h>

namespace pyn {

// Column ids of the next pass (and the extent + first column ids of the group's next row)
// are fetched before the current pass is consumed, which takes the colidx -> x dependent
// load off the critical path; the grid is exactly the resident set (4 CTAs of 256 threads
// per SM at <= 64 registers: one wave, rows interleaved across warps).  Measured on K at
// 16^3/ngl=4: 150.8 us = 5.95 TB/s (first version without the prefetch, 2 waves: 204 us).
// Tighter register caps (6-8 CTAs/SM) made ptxas serialise the nine plane loads and lost
// 40-90 %; unrolling the pass loop lost 3-20 %; a TMA-bulk shared-memory ring fed by a
// producer warp (cp.async.bulk + mbarrier, 128 KB ring) was correct but 6x slower -- 16
// consumer warps each holding a whole row (<= 24.7 KB) span more than the ring, so the
// warps serialise (profiles/README.md, round 1).
// ORD: row-partitioned matrix with the fused halo whose first fuse.n_int rows read no ghost column
// (pynama_b200/parallel.py numbers the owned nodes that way): a warp waits for the ghosts only when it
// reaches its first row >= n_int.  A separate instantiation, so that the one-GPU kernel keeps the exact
// code it was tuned with (r2d: extra state in the common kernel cost 11 % at the 64-register cap; r2f:
// walking the rows through a permutation array instead cost 19 % -- one more dependent load per row).
template <int BR, int BC, int G, int NDOT, bool ORD = false>
__global__ void __launch_bounds__(256, 4)
spmv_kernel(int64_t nbrows, const int64_t* __restrict__ rowptr, const int32_t* __restrict__ colidx,
             const double* __restrict__ val, const double* __restrict__ x, double alpha, double beta,
             const double* __restrict__ z, double* __restrict__ y, SpmvFuse fuse) {
  constexpr int BB = BR * BC;
  constexpr int NV = NDOT > 0 ? NDOT : 1;
  __shared__ double red_sm[NV * 32];
  if (fuse.skip && *fuse.skip != 0.0) return;
  const double* xg = nullptr;          // ghost section of x in the halo window (fused exchange)
  int64_t n_own = 0;
  bool waited = !ORD;                  // deferred, per warp, to the first boundary row
  if (fuse.hw.win) {
    if (!ORD && !halo_wait_cta(fuse.hw)) return;
    xg = halo_buf(fuse.hw.win, (int)(fuse.hw.seq & 1ull), fuse.hw.nghost, fuse.hw.max_bs);
    n_own = fuse.hw.n_own;
  }
  double dotacc[NV];
#pragma unroll
  for (int k = 0; k < NV; ++k) dotacc[k] = 0.0;
  const int lane = threadIdx.x & (G - 1);
  const int64_t grp = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) / G;
  const int64_t ngrp = ((int64_t)gridDim.x * blockDim.x) / G;
  const int64_t iters = (nbrows + ngrp - 1) / ngrp;
  int64_t r = grp;
  bool live = r < nbrows;
  int64_t p0 = live ? rowptr[r] : 0;
  int nb = live ? (int)(rowptr[r + 1] - p0) : 0;
  int32_t cnext = lane < nb ? colidx[p0 + lane] : 0;
  for (int64_t it = 0; it < iters; ++it) {
    // next row's extent, fetched early
    const int64_t rn = r + ngrp;
    const bool liven = rn < nbrows;
    const int64_t p0n = liven ? rowptr[rn] : 0;
    const int nbn = liven ? (int)(rowptr[rn + 1] - p0n) : 0;
    if (ORD) {
      if (!waited && __any_sync(0xffffffffu, live && r >= fuse.n_int)) {
        halo_wait_warp(fuse.hw);
        waited = true;
      }
    }
    const double* __restrict__ v = val + p0 * BB;
    double acc[BR];
#pragma unroll
    for (int i = 0; i < BR; ++i) acc[i] = 0.0;
    for (int u = lane; u < nb; u += G) {
      const int64_t c = cnext;
      if (u + G < nb) cnext = colidx[p0 + u + G];
      double xv[BC];
      if (xg && c >= n_own) {            // written by a peer GPU: read through L2, not L1
#pragma unroll
        for (int j = 0; j < BC; ++j) xv[j] = __ldcg(xg + (c - n_own) * BC + j);
      } else {
#pragma unroll
        for (int j = 0; j < BC; ++j) xv[j] = x[c * BC + j];
      }
#pragma unroll
      for (int i = 0; i < BR; ++i)
#pragma unroll
        for (int j = 0; j < BC; ++j) acc[i] += __ldcs(v + (int64_t)(i * BC + j) * nb + u) * xv[j];
    }
    if (lane < nbn) cnext = colidx[p0n + lane];
#pragma unroll
    for (int i = 0; i < BR; ++i)
#pragma unroll
      for (int o = G / 2; o > 0; o >>= 1) acc[i] += __shfl_xor_sync(0xffffffffu, acc[i], o, G);
#pragma unroll
    for (int i = 0; i < BR; ++i) {
      if (live && lane == i % G) {
        double out = alpha * acc[i];
        if (beta != 0.0) out += beta * z[r * BR + i];
        y[r * BR + i] = out;
        if (NDOT > 0) {
          const double xr = x[r * BC + i];
          dotacc[0] += out * xr;
          if (NDOT == 3) {
            dotacc[NV > 1 ? 1 : 0] += fuse.dot_r[r * BR + i] * xr;
            dotacc[NV > 2 ? 2 : 0] += xr * xr;
          }
        }
      }
    }
    r = rn;
    live = liven;
    p0 = p0n;
    nb = nbn;
  }
  if (NDOT > 0) {
    block_reduce<NV>(dotacc, red_sm);
    if (fuse.red.nranks > 1)
      grid_reduce_push<NV>(dotacc, fuse.partials, fuse.ticket, fuse.red, 0);
    else
      grid_reduce_finish<NV>(dotacc, fuse.partials, fuse.ticket, fuse.dot_out);
  }
}

static int spmv_launch(pyn_mat* m, double alpha, const double* x, double beta, const double* z,
                       double* y, bool dot, SpmvFuse fuse) {
  pyn_ctx* ctx = m->ctx;
  if (m->nbrows == 0) return PYN_OK;
  int64_t nrows = m->nbrows;
  if (m->halo) {                       // distributed: refresh the ghosts of x, own rows only
    if (!fuse.hw.win) PYN_CHECK(halo_exchange(m->halo, const_cast<double*>(x), m->bc));
    nrows = m->halo->n_own;
    if (nrows == 0) return PYN_OK;
  }
  if (dot && fuse.hw.win && m->n_interior > 0 && !g_halo_wait_at_start) fuse.n_int = m->n_interior;
  if (m->mf) return matfree_apply(m, nrows, alpha, x, beta, z, y, dot, fuse);
  const double avg = m->nbrows ? (double)m->nblocks / (double)m->nbrows : 0.0;
  const int G = avg > 24.0 ? 32 : (avg > 10.0 ? 16 : 8);
  // one wave: 4 resident CTAs per SM (__launch_bounds__(256, 4))
  const int grid = grid_for(ctx, nrows * G, 256, 4);
#define LAUNCH(BR, BC, GG, DD)                                                                        \
  do {                                                                                                \
    if (DD > 0 && fuse.n_int > 0)                                                                     \
      spmv_kernel<BR, BC, GG, DD, (DD > 0)><<<grid, 256, 0, ctx->stream>>>(                           \
          nrows, m->rowptr.p, m->colidx.p, m->val.p, x, alpha, beta, z, y, fuse);                     \
    else                                                                                              \
      spmv_kernel<BR, BC, GG, DD, false><<<grid, 256, 0, ctx->stream>>>(                              \
          nrows, m->rowptr.p, m->colidx.p, m->val.p, x, alpha, beta, z, y, fuse);                     \
  } while (0)
#define CALL(BR, BC)                                     \
  do {                                                   \
    if (dot) {                                           \
      if (BR != BC) {                                    \
        set_error("spmv+dot needs square blocks");       \
        return PYN_E_INVALID;                            \
      }                                                  \
      if (fuse.dot_r) {                                  \
        if (G == 32) LAUNCH(BR, BC, 32, 3);              \
        else if (G == 16) LAUNCH(BR, BC, 16, 3);         \
        else LAUNCH(BR, BC, 8, 3);                       \
      } else {                                           \
        if (G == 32) LAUNCH(BR, BC, 32, 1);              \
        else if (G == 16) LAUNCH(BR, BC, 16, 1);         \
        else LAUNCH(BR, BC, 8, 1);                       \
      }                                                  \
    } else {                                             \
      if (G == 32) LAUNCH(BR, BC, 32, 0);                \
      else if (G == 16) LAUNCH(BR, BC, 16, 0);           \
      else LAUNCH(BR, BC, 8, 0);                         \
    }                                                    \
  } while (0)
  prof_begin(ctx, dot ? PROF_SPMV_KSP : PROF_SPMV);
  PYN_DISPATCH_BLOCK(m->br, m->bc, CALL);
#undef CALL
#undef LAUNCH
  prof_end(ctx);
  PYN_LAUNCHED(ctx);
  return PYN_OK;
}

// flag[r] = 1 when row r holds a column >= n_own (one warp per row)
// *nreal += blocks of the rows that are not padding (padding repeats the row's last column)
__global__ void ghost_rows_kernel(int64_t nrows, int64_t n_own, const int64_t* __restrict__ rowptr,
                                  const int32_t* __restrict__ colidx, uint8_t* __restrict__ flag,
                                  unsigned long long* __restrict__ nreal) {
  const int lane = threadIdx.x & 31;
  const int64_t warp = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
  for (int64_t r = warp; r < nrows; r += nwarps) {
    bool g = false;
    unsigned int real = 0;
    for (int64_t k = rowptr[r] + lane; k < rowptr[r + 1]; k += 32) {
      g |= colidx[k] >= n_own;
      real += (k == rowptr[r] || colidx[k] != colidx[k - 1]) ? 1u : 0u;
    }
    g = __any_sync(0xffffffffu, g);
    for (int o = 16; o > 0; o >>= 1) real += __shfl_xor_sync(0xffffffffu, real, o);
    if (lane == 0) {
      flag[r] = g ? 1 : 0;
      atomicAdd(nreal, (unsigned long long)real);      // integer sum: order does not matter
    }
  }
}

int spmv_build_row_order(pyn_mat* m) {
  pyn_ctx* ctx = m->ctx;
  m->n_interior = -1;
  if (!m->halo || m->halo->n_own <= 0 || m->nblocks == 0) return PYN_OK;
  const int64_t n_own = m->halo->n_own;
  DevBuf<uint8_t> flag;
  DevBuf<unsigned long long> nreal;
  PYN_CHECK(flag.alloc(ctx, n_own));
  PYN_CHECK(nreal.alloc(ctx, 1));
  PYN_CHECK(nreal.zero());
  ghost_rows_kernel<<<grid_for(ctx, n_own * 32, 256), 256, 0, ctx->stream>>>(n_own, n_own, m->rowptr.p,
                                                                              m->colidx.p, flag.p, nreal.p);
  PYN_LAUNCHED(ctx);
  unsigned long long nr = 0;
  PYN_CHECK(nreal.download(&nr, 1));
  m->own_blocks_real = (int64_t)nr;
  std::vector<uint8_t> f((size_t)n_own);
  PYN_CHECK(flag.download(f.data(), n_own));
  // the deferred wait needs the rows that read ghosts at the END of the owned rows (the numbering of
  // pynama_b200/parallel.py); any other numbering keeps the wait at CTA start
  int64_t n_int = 0;
  while (n_int < n_own && !f[(size_t)n_int]) ++n_int;
  // (rows after n_int need not all hold a ghost column: only the rows BEFORE n_int must be ghost-free)
  m->n_interior = n_int;
  return PYN_OK;
}

int spmv(pyn_mat* m, double alpha, const double* x, double beta, const double* z, double* y) {
  SpmvFuse f;
  memset(&f, 0, sizeof(f));
  return spmv_launch(m, alpha, x, beta, z, y, false, f);
}

// y = A x and dot_out = x . y in one pass; no-op when *skip != 0
int spmv_dot(pyn_mat* m, const double* x, double* y, double* dot_out, const double* skip, const P2PRed* red,
             const HaloWait* hw) {
  pyn_ctx* ctx = m->ctx;
  SpmvFuse f;
  memset(&f, 0, sizeof(f));
  f.skip = skip;
  f.partials = ctx->red_partials;
  f.ticket = ctx->red_ticket;
  f.dot_out = dot_out;
  if (red) f.red = *red;
  if (hw) f.hw = *hw;
  return spmv_launch(m, 1.0, x, 0.0, nullptr, y, true, f);
}

int spmv_dot3(pyn_mat* m, const double* x, double* y, const double* r, double* out3, const double* skip,
              const P2PRed* red) {
  pyn_ctx* ctx = m->ctx;
  SpmvFuse f;
  memset(&f, 0, sizeof(f));
  f.skip = skip;
  f.partials = ctx->red_partials;
  f.ticket = ctx->red_ticket;
  f.dot_out = out3;
  f.dot_r = r;
  if (red) f.red = *red;
  return spmv_launch(m, 1.0, x, 0.0, nullptr, y, true, f);
}

// ---- diagonal blocks / row scaling ------------------------------------------------
// dblk[r][i][j] = diagonal node block of a square-block matrix (zeros if absent)
template <int B>
__global__ void diag_blocks_kernel(int64_t nbrows, const int64_t* __restrict__ rowptr,
                                   const int32_t* __restrict__ colidx, const double* __restrict__ val,
                                   double* __restrict__ dblk) {
  for (int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; r < nbrows;
       r += (int64_t)gridDim.x * blockDim.x) {
    const int64_t p0 = rowptr[r];
    const int nb = (int)(rowptr[r + 1] - p0);
    int lo = 0, hi = nb - 1, pos = -1;
    while (lo <= hi) {
      const int mid = (lo + hi) >> 1;
      const int32_t c = colidx[p0 + mid];
      if (c == r) { pos = mid; break; }
      if (c < r) lo = mid + 1; else hi = mid - 1;
    }
    while (pos > 0 && colidx[p0 + pos - 1] == (int32_t)r) --pos;   // not a padding repeat of the column
    for (int q = 0; q < B * B; ++q)
      dblk[r * B * B + q] = pos >= 0 ? val[p0 * B * B + (int64_t)q * nb + pos] : 0.0;
  }
}

int diag_blocks(pyn_mat* m, double* dblk) {
  pyn_ctx* ctx = m->ctx;
  PYN_REQUIRE(m->br == m->bc, "diagonal of a non-square-block matrix");
  const int grid = grid_for(ctx, m->nbrows, 128);
  if (m->br == 3)
    diag_blocks_kernel<3><<<grid, 128, 0, ctx->stream>>>(m->nbrows, m->rowptr.p, m->colidx.p, m->val.p, dblk);
  else if (m->br == 2)
    diag_blocks_kernel<2><<<grid, 128, 0, ctx->stream>>>(m->nbrows, m->rowptr.p, m->colidx.p, m->val.p, dblk);
  else if (m->br == 1)
    diag_blocks_kernel<1><<<grid, 128, 0, ctx->stream>>>(m->nbrows, m->rowptr.p, m->colidx.p, m->val.p, dblk);
  else {
    set_error("diag_blocks: unsupported block size %d", m->br);
    return PYN_E_UNSUPPORTED;
  }
  PYN_LAUNCHED(ctx);
  return PYN_OK;
}

__global__ void scale_rows_kernel(int br, int bc, int64_t nbrows, const int64_t* __restrict__ rowptr,
                                  const double* __restrict__ l, double* __restrict__ val) {
  const int lane = threadIdx.x & 31;
  const int64_t warp = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
  const int bb = br * bc;
  for (int64_t r = warp; r < nbrows; r += nwarps) {
    const int64_t p0 = rowptr[r];
    const int64_t nb = rowptr[r + 1] - p0;
    for (int64_t t = lane; t < nb * bb; t += 32) {
      const int i = (int)(t / nb) / bc;
      val[p0 * bb + t] *= l[r * br + i];
    }
  }
}

int scale_rows(pyn_mat* m, const double* l) {
  pyn_ctx* ctx = m->ctx;
  scale_rows_kernel<<<grid_for(ctx, m->nbrows * 32, 256), 256, 0, ctx->stream>>>(
      m->br, m->bc, m->nbrows, m->rowptr.p, l, m->val.p);
  PYN_LAUNCHED(ctx);
  return PYN_OK;
}

// ---- C = A + B on the union pattern (both filtered from the same graph) -------------
__global__ void mark_slots_kernel(const int64_t* __restrict__ base, int64_t n, int64_t* __restrict__ pos) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n;
       i += (int64_t)gridDim.x * blockDim.x)
    if (base[i] >= 0) pos[base[i]] = i;            // padding blocks have no slot
}

// one warp per row of C: for every block find the source positions in A and B
__global__ void add_values_kernel(int bb, int64_t nbrows, const int64_t* __restrict__ c_rowptr,
                                  const int64_t* __restrict__ c_base,
                                  const int64_t* __restrict__ a_rowptr, const int64_t* __restrict__ a_pos,
                                  const double* __restrict__ a_val, const int64_t* __restrict__ b_rowptr,
                                  const int64_t* __restrict__ b_pos, const double* __restrict__ b_val,
                                  double* __restrict__ c_val) {
  const int lane = threadIdx.x & 31;
  const int64_t warp = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
  for (int64_t r = warp; r < nbrows; r += nwarps) {
    const int64_t p0 = c_rowptr[r], nb = c_rowptr[r + 1] - p0;
    const int64_t ap0 = a_rowptr[r], anb = a_rowptr[r + 1] - ap0;
    const int64_t bp0 = b_rowptr[r], bnb = b_rowptr[r + 1] - bp0;
    for (int64_t u = lane; u < nb; u += 32) {
      const int64_t s = c_base[p0 + u];
      const int64_t pa = s >= 0 ? a_pos[s] : -1, pb = s >= 0 ? b_pos[s] : -1;
      for (int q = 0; q < bb; ++q) {
        double v = 0.0;
        if (pa >= 0) v += a_val[ap0 * bb + (int64_t)q * anb + (pa - ap0)];
        if (pb >= 0) v += b_val[bp0 * bb + (int64_t)q * bnb + (pb - bp0)];
        c_val[p0 * bb + (int64_t)q * nb + u] = v;
      }
    }
  }
}

__global__ void fill_i64_kernel(int64_t* p, int64_t n, int64_t v) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n;
       i += (int64_t)gridDim.x * blockDim.x)
    p[i] = v;
}

int mat_add_new(pyn_mat* a, pyn_mat* b, pyn_mat** out) {
  PYN_REQUIRE(a->graph == b->graph && a->br == b->br && a->bc == b->bc,
              "mat_add: operands must come from the same node graph with equal block shape");
  PYN_REQUIRE(a->use_rowcls == b->use_rowcls && a->use_colcls == b->use_colcls,
              "mat_add: operands use different dof-class masks");
  pyn_ctx* ctx = a->ctx;
  pyn_mat* c = new pyn_mat();
  int rc = make_pattern(a->mesh, a->graph, a->br, a->bc, a->use_rowcls, a->use_colcls,
                        a->pairmask | b->pairmask, a->diagmask | b->diagmask, c);
  if (rc != PYN_OK) {
    delete c;
    return rc;
  }
  c->ref_dense = a->ref_dense;
  c->halo = a->halo;
  DevBuf<int64_t> apos, bpos;
  const int64_t ns = a->graph->nslots;
  rc = apos.alloc(ctx, ns);
  if (rc == PYN_OK) rc = bpos.alloc(ctx, ns);
  if (rc != PYN_OK) {
    delete c;
    return rc;
  }
  const int g1 = grid_for(ctx, ns, 256);
  fill_i64_kernel<<<g1, 256, 0, ctx->stream>>>(apos.p, ns, -1);
  fill_i64_kernel<<<g1, 256, 0, ctx->stream>>>(bpos.p, ns, -1);
  ctx->launches += 2;
  if (a->nblocks) {
    mark_slots_kernel<<<grid_for(ctx, a->nblocks, 256), 256, 0, ctx->stream>>>(a->base_slot.p, a->nblocks, apos.p);
    ctx->launches++;
  }
  if (b->nblocks) {
    mark_slots_kernel<<<grid_for(ctx, b->nblocks, 256), 256, 0, ctx->stream>>>(b->base_slot.p, b->nblocks, bpos.p);
    ctx->launches++;
  }
  if (c->nblocks) {
    add_values_kernel<<<grid_for(ctx, c->nbrows * 32, 256), 256, 0, ctx->stream>>>(
        c->br * c->bc, c->nbrows, c->rowptr.p, c->base_slot.p, a->rowptr.p, apos.p, a->val.p,
        b->rowptr.p, bpos.p, b->val.p, c->val.p);
    ctx->launches++;
  }
  cudaError_t e = cudaStreamSynchronize(ctx->stream);
  if (e != cudaSuccess) {
    set_error("mat_add: %s", cudaGetErrorString(e));
    delete c;
    return PYN_E_CUDA;
  }
  *out = c;
  return PYN_OK;
}

}  // namespace pyn
