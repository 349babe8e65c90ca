// assemble.cu -- numeric phase of the assembly (k4, k5): segmented reduction of
// the sorted COO into node-block matrices, with the reference's Dirichlet /
// no-slip masking, diagonal fix-ups and lumped-mass row scaling fused in.
//
// Follows MatFS.buildFS (mat_fs.py:150-192), MatNS.buildNS (mat_ns.py:47-145),
// Operators.setValues/assembleAll (mat_fs.py:238-271).  The dof split of the
// reference (free / Dirichlet / no-slip normal / no-slip tangential) is a
// per-dof class; which (row class, column class) combinations a matrix takes
// from the element block is a 16-bit mask.
#include "elem.cuh"

namespace pyn {

struct DiagAdd {
  double v[4];
};

__device__ __forceinline__ bool block_kept(int br, int bc, int64_t r, int32_t c,
                                           const uint8_t* __restrict__ cls, bool use_rowcls,
                                           bool use_colcls, uint32_t pairmask, uint32_t diagmask) {
  bool keep = false;
  for (int i = 0; i < br; ++i) {
    const int rc = use_rowcls ? cls[r * br + i] : 0;
    if (r == c && ((diagmask >> rc) & 1u)) keep = true;
    for (int j = 0; j < bc; ++j) {
      const int cc = use_colcls ? cls[(int64_t)c * bc + j] : 0;
      if ((pairmask >> (rc * 4 + cc)) & 1u) keep = true;
    }
  }
  return keep;
}

// pass 0: count kept slots per row; pass 1: write colidx/base_slot (ordered)
template <int PASS>
__global__ void filter_kernel(int br, int bc, int64_t nrows, const int64_t* __restrict__ g_rowptr,
                              const int32_t* __restrict__ g_colidx, const uint8_t* __restrict__ cls,
                              bool use_rowcls, bool use_colcls, uint32_t pairmask, uint32_t diagmask,
                              int32_t* __restrict__ rowlen, const int64_t* __restrict__ m_rowptr,
                              int32_t* __restrict__ m_colidx, int64_t* __restrict__ m_base) {
  const int lane = threadIdx.x & 31;
  const int64_t warp = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
  for (int64_t r = warp; r < nrows; r += nwarps) {
    const int64_t p0 = g_rowptr[r], len = g_rowptr[r + 1] - p0;
    int64_t out = PASS ? m_rowptr[r] : 0;
    int total = 0;
    for (int64_t u0 = 0; u0 < len; u0 += 32) {
      const int64_t u = u0 + lane;
      bool keep = false;
      int32_t c = 0;
      if (u < len) {
        c = g_colidx[p0 + u];
        keep = block_kept(br, bc, r, c, cls, use_rowcls, use_colcls, pairmask, diagmask);
      }
      const unsigned bal = __ballot_sync(0xffffffffu, keep);
      if (PASS && keep) {
        const int64_t pos = out + __popc(bal & ((1u << lane) - 1u));
        m_colidx[pos] = c;
        m_base[pos] = p0 + u;
      }
      out += __popc(bal);
      total += __popc(bal);
    }
    if (!PASS && lane == 0) rowlen[r] = total;
  }
}

template <int BR, int BC>
__global__ void __launch_bounds__(256)
numeric_kernel(int64_t nrows, const int64_t* __restrict__ m_rowptr,
               const int32_t* __restrict__ m_colidx, const int64_t* __restrict__ m_base,
               const int64_t* __restrict__ cptr, const int2* __restrict__ csrc,
               const double* __restrict__ ebs, int64_t ebs_pairs, int ebs_ppa, int64_t e_base,
               const int32_t* __restrict__ pair_map, double sign,
               const uint8_t* __restrict__ cls, bool use_rowcls, bool use_colcls,
               uint32_t pairmask, uint32_t diagmask, DiagAdd dadd,
               const double* __restrict__ row_scale, double* __restrict__ val) {
  constexpr int BB = BR * BC;
  const int lane = threadIdx.x & 31;
  const int64_t warp = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
  for (int64_t r = warp; r < nrows; r += nwarps) {
    const int64_t p0 = m_rowptr[r];
    const int64_t nb = m_rowptr[r + 1] - p0;
    int rc[BR];
#pragma unroll
    for (int i = 0; i < BR; ++i) rc[i] = use_rowcls ? cls[r * BR + i] : 0;
    const double rs = row_scale ? row_scale[r] : 1.0;
    for (int64_t u = lane; u < nb; u += 32) {
      const int64_t s = m_base[p0 + u];
      const int32_t c = m_colidx[p0 + u];
      double acc[BB];
#pragma unroll
      for (int q = 0; q < BB; ++q) acc[q] = 0.0;
      const int64_t k1 = cptr[s + 1];
      for (int64_t k = cptr[s]; k < k1; ++k) {
        const int2 eq = csrc[k];
        const int32_t qq = pair_map ? pair_map[eq.y] : eq.y;
        if (qq >= 0) {
          if (ebs_ppa > 0) {       // plane-major store
            const int a = qq / ebs_ppa, m = qq - a * ebs_ppa;
            const double* src = ebs + (((int64_t)eq.x - e_base) * ebs_pairs + (int64_t)a * ebs_ppa) * BB + m;
#pragma unroll
            for (int q = 0; q < BB; ++q) acc[q] += __ldcs(src + (int64_t)q * ebs_ppa);
          } else {
            const double* src = ebs + (((int64_t)eq.x - e_base) * ebs_pairs + qq) * BB;
#pragma unroll
            for (int q = 0; q < BB; ++q) acc[q] += src[q];
          }
        }
      }
#pragma unroll
      for (int i = 0; i < BR; ++i) {
#pragma unroll
        for (int j = 0; j < BC; ++j) {
          const int cc = use_colcls ? cls[(int64_t)c * BC + j] : 0;
          const bool ok = (pairmask >> (rc[i] * 4 + cc)) & 1u;
          double v = ok ? sign * acc[i * BC + j] : 0.0;
          if (BR == BC && i == j && c == r && ((diagmask >> rc[i]) & 1u)) v += dadd.v[rc[i]];
          if (row_scale) v = v * rs;
          val[p0 * BB + (int64_t)(i * BC + j) * nb + u] = v;
        }
      }
    }
  }
}

__global__ void node_weights_kernel(const int64_t* __restrict__ n2e_ptr, const int2* __restrict__ n2e,
                                    const double* __restrict__ ebsW, int nn, int64_t nnodes,
                                    double* __restrict__ W, double* __restrict__ Winv) {
  for (int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; r < nnodes;
       r += (int64_t)gridDim.x * blockDim.x) {
    double s = 0.0;
    for (int64_t k = n2e_ptr[r]; k < n2e_ptr[r + 1]; ++k) s += ebsW[(int64_t)n2e[k].x * nn + n2e[k].y];
    W[r] = s;
    if (Winv) Winv[r] = 1.0 / s;     // Vec.reciprocal (mat_fs.py:256,261,266)
  }
}

int assemble_node_weights(pyn_mesh* mesh, const double* ebsW, double* W, double* Winv) {
  pyn_ctx* ctx = mesh->ctx;
  node_weights_kernel<<<grid_for(ctx, mesh->nnodes, 128), 128, 0, ctx->stream>>>(
      mesh->n2e_ptr.p, mesh->n2e.p, ebsW, mesh->nn, mesh->nnodes, W, Winv);
  PYN_LAUNCHED(ctx);
  return PYN_OK;
}

// Filter `graph` into the pattern of a new matrix (no values yet).
int make_pattern(pyn_mesh* mesh, const NodeGraph* graph, int br, int bc, bool use_rowcls,
                 bool use_colcls, uint32_t pairmask, uint32_t diagmask, pyn_mat* m) {
  pyn_ctx* ctx = mesh->ctx;
  m->ctx = ctx;
  m->mesh = mesh;
  m->br = br;
  m->bc = bc;
  m->nbrows = graph->nrows;
  m->nbcols = graph->nrows;
  m->graph = graph;
  m->use_rowcls = use_rowcls;
  m->use_colcls = use_colcls;
  m->pairmask = pairmask;
  m->diagmask = diagmask;
  if (use_rowcls || use_colcls)
    PYN_REQUIRE(mesh->has_cls, "dof classes not set (pyn_mesh_set_dof_class)");
  DevBuf<int32_t> rowlen;
  PYN_CHECK(rowlen.alloc(ctx, graph->nrows));
  PYN_CHECK(m->rowptr.alloc(ctx, graph->nrows + 1));
  const int grid = grid_for(ctx, graph->nrows * 32, 256);
  filter_kernel<0><<<grid, 256, 0, ctx->stream>>>(br, bc, graph->nrows, graph->rowptr.p,
                                                  graph->colidx.p, mesh->dofcls.p, use_rowcls,
                                                  use_colcls, pairmask, diagmask, rowlen.p, nullptr,
                                                  nullptr, nullptr);
  PYN_LAUNCHED(ctx);
  PYN_CHECK(exclusive_scan_i32_to_i64(ctx, rowlen.p, m->rowptr.p, graph->nrows, &m->nblocks));
  PYN_CHECK(m->colidx.alloc(ctx, m->nblocks));
  PYN_CHECK(m->base_slot.alloc(ctx, m->nblocks));
  filter_kernel<1><<<grid, 256, 0, ctx->stream>>>(br, bc, graph->nrows, graph->rowptr.p,
                                                  graph->colidx.p, mesh->dofcls.p, use_rowcls,
                                                  use_colcls, pairmask, diagmask, nullptr,
                                                  m->rowptr.p, m->colidx.p, m->base_slot.p);
  PYN_LAUNCHED(ctx);
  PYN_CHECK(m->val.alloc(ctx, m->nblocks * br * bc));
  PYN_CUDA(cudaStreamSynchronize(ctx->stream));
  return PYN_OK;
}

int assemble_matrix(pyn_mesh* mesh, const MatSpec& sp, pyn_mat** out) {
  pyn_ctx* ctx = mesh->ctx;
  pyn_mat* m = new pyn_mat();
  int rc = make_pattern(mesh, sp.graph, sp.br, sp.bc, sp.use_rowcls, sp.use_colcls, sp.pairmask,
                        sp.diagmask, m);
  if (rc != PYN_OK) {
    delete m;
    return rc;
  }
  m->ref_dense = sp.ref_dense;
  DiagAdd da;
  for (int i = 0; i < 4; ++i) da.v[i] = sp.diag_add[i];
  if (m->nblocks > 0) {
    const int grid = grid_for(ctx, m->nbrows * 32, 256);
#define CALL(BR, BC)                                                                              \
  numeric_kernel<BR, BC><<<grid, 256, 0, ctx->stream>>>(                                          \
      m->nbrows, m->rowptr.p, m->colidx.p, m->base_slot.p, sp.graph->cptr.p, sp.graph->csrc.p,    \
      sp.ebs, sp.ebs_pairs, sp.ebs_ppa, 0, sp.pair_map, sp.sign, mesh->dofcls.p, sp.use_rowcls, sp.use_colcls, \
      sp.pairmask, sp.diagmask, da, sp.row_scale, m->val.p)
    prof_begin(ctx, PROF_NUMERIC);
    PYN_DISPATCH_BLOCK(sp.br, sp.bc, CALL);
#undef CALL
    prof_end(ctx);
    PYN_LAUNCHED(ctx);
  }
  *out = m;
  return PYN_OK;
}

}  // namespace pyn
