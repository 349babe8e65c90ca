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

// rows hold a multiple of 4 blocks: padding blocks after the kept ones
__host__ __device__ __forceinline__ int pad_blocks(int kept) { return (4 - (kept & 3)) & 3; }

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
                              int32_t* __restrict__ rowlen, unsigned long long* __restrict__ nreal,
                              const int64_t* __restrict__ m_rowptr,
                              int32_t* __restrict__ m_colidx, int64_t* __restrict__ m_base) {
  const int lane = threadIdx.x & 31;
  const int64_t warp = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
  for (int64_t r = warp; r < nrows; r += nwarps) {
    const int64_t p0 = g_rowptr[r], len = g_rowptr[r + 1] - p0;
    int64_t out = PASS ? m_rowptr[r] : 0;
    int total = 0;
    int32_t lastc = 0;
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
      if (bal) lastc = __shfl_sync(0xffffffffu, c, 31 - __clz(bal));
      out += __popc(bal);
      total += __popc(bal);
    }
    const int npad = pad_blocks(total);
    if (PASS && lane < npad) {                       // zero blocks repeating the last column
      m_colidx[out + lane] = lastc;
      m_base[out + lane] = -1;
    }
    if (!PASS && lane == 0) {
      rowlen[r] = total + npad;
      if (total) atomicAdd(nreal, (unsigned long long)total);
    }
  }
}

// The same two passes for up to four matrices filtered from one graph with row classes (the
// K, Krhs, Rw, Rd of pyn_assemble_kle): column ids and dof classes are read once.
struct FilterTarget {
  int bc;
  int use_colcls;
  uint32_t pairmask, diagmask;
  int32_t* rowlen;           // pass 0
  unsigned long long* nreal; // pass 0: kept blocks of the whole matrix
  const int64_t* rowptr;     // pass 1
  int32_t* colidx;
  int64_t* base;
};
struct FilterTargets {
  FilterTarget t[4];
};

// Packed dof classes of a node (pyn_mesh::nodecls): 2 bits per dof.
__device__ __forceinline__ int dof_class(unsigned pack, int i) { return (pack >> (2 * i)) & 3u; }
// set of the classes among the first D dofs (bit a = class a present)
template <int D>
__device__ __forceinline__ unsigned class_set(unsigned pack) {
  unsigned s = 0;
#pragma unroll
  for (int i = 0; i < D; ++i) s |= 1u << dof_class(pack, i);
  return s;
}
// bits (rowcls*4 + colcls) of a pair mask reachable from a set of row classes / column classes
__device__ __forceinline__ unsigned row_class_bits(unsigned rset) {
  return ((rset & 1u) ? 0x000Fu : 0u) | ((rset & 2u) ? 0x00F0u : 0u) | ((rset & 4u) ? 0x0F00u : 0u) |
         ((rset & 8u) ? 0xF000u : 0u);
}
__device__ __forceinline__ unsigned col_class_bits(unsigned cset) { return cset * 0x1111u; }

template <int PASS, int D>
__global__ void filter4_kernel(int64_t nrows, const int64_t* __restrict__ g_rowptr,
                               const int32_t* __restrict__ g_colidx,
                               const uint8_t* __restrict__ nodecls, FilterTargets T) {
  const int lane = threadIdx.x & 31;
  const int64_t warp = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
  for (int64_t r = warp; r < nrows; r += nwarps) {
    const int64_t p0 = g_rowptr[r];
    const int len = (int)(g_rowptr[r + 1] - p0);
    const unsigned rset = class_set<D>(nodecls[r]);
    const unsigned rbits = row_class_bits(rset);
    int64_t out[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) out[k] = PASS ? T.t[k].rowptr[r] : 0;
    int total[4] = {0, 0, 0, 0};
    int32_t lastc[4] = {0, 0, 0, 0};
    for (int u0 = 0; u0 < len; u0 += 32) {
      const int u = u0 + lane;
      int32_t c = 0;
      unsigned cbits = 0;
      if (u < len) {
        c = g_colidx[p0 + u];
        cbits = col_class_bits(class_set<D>(nodecls[c]));
      }
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        // block_kept: some (row class, column class) pair allowed, or a diagonal fix-up
        const unsigned cb = T.t[k].use_colcls ? cbits : 0x1111u;
        const bool keep = u < len && ((T.t[k].pairmask & rbits & cb) != 0u ||
                                      ((int64_t)c == r && (T.t[k].diagmask & rset) != 0u));
        const unsigned bal = __ballot_sync(0xffffffffu, keep);
        if (PASS && keep) {
          const int64_t pos = out[k] + __popc(bal & ((1u << lane) - 1u));
          T.t[k].colidx[pos] = c;
          T.t[k].base[pos] = p0 + u;
        }
        if (bal) lastc[k] = __shfl_sync(0xffffffffu, c, 31 - __clz(bal));
        out[k] += __popc(bal);
        total[k] += __popc(bal);
      }
    }
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const int npad = pad_blocks(total[k]);
      if (PASS && lane < npad) {                     // zero blocks repeating the last column
        T.t[k].colidx[out[k] + lane] = lastc[k];
        T.t[k].base[out[k] + lane] = -1;
      }
      if (!PASS && lane == 0) {
        T.t[k].rowlen[r] = total[k] + npad;
        if (total[k]) atomicAdd(T.t[k].nreal, (unsigned long long)total[k]);
      }
    }
  }
}

// G lanes per block row (32, or 16 / 8 for the short rows of the stencil graph)
template <int BR, int BC, int G>
__global__ void __launch_bounds__(256)
numeric_kernel(int64_t nrows, const int64_t* __restrict__ m_rowptr,
               const int32_t* __restrict__ m_colidx, const int64_t* __restrict__ m_base,
               const int64_t* __restrict__ cptr, const int2* __restrict__ csrc,
               const double* __restrict__ ebs,
               int64_t ebs_pairs, int ebs_ppa, int ebs_stride, int ebs_off,
               const int32_t* __restrict__ pair_map, double sign,
               const uint8_t* __restrict__ cls, bool use_rowcls, bool use_colcls,
               uint32_t pairmask, uint32_t diagmask, DiagAdd dadd,
               const double* __restrict__ row_scale, double* __restrict__ val) {
  constexpr int BB = BR * BC;
  const int lane = threadIdx.x & (G - 1);
  const int64_t warp = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) / G;
  const int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) / G;
  for (int64_t r = warp; r < nrows; r += nwarps) {
    const int64_t p0 = m_rowptr[r];
    const int64_t nb = m_rowptr[r + 1] - p0;
    int rc[BR];
#pragma unroll
    for (int i = 0; i < BR; ++i) rc[i] = use_rowcls ? cls[r * BR + i] : 0;
    const double rs = row_scale ? row_scale[r] : 1.0;
    for (int64_t u = lane; u < nb; u += G) {
      const int64_t s = m_base[p0 + u];
      const int32_t c = m_colidx[p0 + u];
      if (s < 0) {                                   // padding block
#pragma unroll
        for (int q = 0; q < BB; ++q) val[p0 * BB + (int64_t)q * nb + u] = 0.0;
        continue;
      }
      double acc[BB];
#pragma unroll
      for (int q = 0; q < BB; ++q) acc[q] = 0.0;
      const int64_t k1 = cptr[s + 1];
      for (int64_t k = cptr[s]; k < k1; ++k) {
        const int2 eq = csrc[k];
        const int32_t qq = pair_map ? pair_map[eq.y] : eq.y;
        const int64_t es = eq.x;                                         // index in the store
        if (qq >= 0) {
          if (ebs_ppa > 0) {       // plane-major store
            const int a = qq / ebs_ppa, m = qq - a * ebs_ppa;
            const double* src = ebs + (es * ebs_pairs + (int64_t)a * ebs_ppa) * BB + m;
#pragma unroll
            for (int q = 0; q < BB; ++q) acc[q] += __ldcs(src + (int64_t)q * ebs_ppa);
          } else {
            const double* src = ebs + (es * ebs_pairs + qq) * ebs_stride + ebs_off;
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
                                    const double* __restrict__ ebsW,
                                    const int32_t* __restrict__ eclass, int nn, int64_t nnodes,
                                    double* __restrict__ W, double* __restrict__ Winv) {
  for (int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; r < nnodes;
       r += (int64_t)gridDim.x * blockDim.x) {
    double s = 0.0;
    for (int64_t k = n2e_ptr[r]; k < n2e_ptr[r + 1]; ++k) {
      const int2 ea = n2e[k];
      s += ebsW[(eclass ? (int64_t)eclass[ea.x] : (int64_t)ea.x) * nn + ea.y];
    }
    W[r] = s;
    if (Winv) Winv[r] = 1.0 / s;     // Vec.reciprocal (mat_fs.py:256,261,266)
  }
}

int assemble_node_weights(pyn_mesh* mesh, const double* ebsW, double* W, double* Winv) {
  pyn_ctx* ctx = mesh->ctx;
  node_weights_kernel<<<grid_for(ctx, mesh->nnodes, 128), 128, 0, ctx->stream>>>(
      mesh->n2e_ptr.p, mesh->n2e.p, ebsW, mesh->elem_class.p, mesh->nn, mesh->nnodes, W, Winv);
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
  DevBuf<unsigned long long> nreal;
  PYN_CHECK(rowlen.alloc(ctx, graph->nrows));
  PYN_CHECK(nreal.alloc(ctx, 1));
  PYN_CHECK(nreal.zero());
  PYN_CHECK(m->rowptr.alloc(ctx, graph->nrows + 1));
  const int grid = grid_for(ctx, graph->nrows * 32, 256);
  filter_kernel<0><<<grid, 256, 0, ctx->stream>>>(br, bc, graph->nrows, graph->rowptr.p,
                                                  graph->colidx.p, mesh->dofcls.p, use_rowcls,
                                                  use_colcls, pairmask, diagmask, rowlen.p, nreal.p,
                                                  nullptr, nullptr, nullptr);
  PYN_LAUNCHED(ctx);
  {
    unsigned long long h = 0;
    PYN_CHECK(nreal.download(&h, 1));
    m->nblocks_real = (int64_t)h;
  }
  PYN_CHECK(exclusive_scan_i32_to_i64(ctx, rowlen.p, m->rowptr.p, graph->nrows, &m->nblocks));
  PYN_CHECK(m->colidx.alloc(ctx, m->nblocks));
  PYN_CHECK(m->base_slot.alloc(ctx, m->nblocks));
  filter_kernel<1><<<grid, 256, 0, ctx->stream>>>(br, bc, graph->nrows, graph->rowptr.p,
                                                  graph->colidx.p, mesh->dofcls.p, use_rowcls,
                                                  use_colcls, pairmask, diagmask, nullptr, nullptr,
                                                  m->rowptr.p, m->colidx.p, m->base_slot.p);
  PYN_LAUNCHED(ctx);
  PYN_CHECK(m->val.alloc(ctx, m->nblocks * br * bc));
  PYN_CUDA(cudaStreamSynchronize(ctx->stream));
  return PYN_OK;
}

__global__ void resolve_class_kernel(const int2* __restrict__ csrc, const int32_t* __restrict__ eclass,
                                     int64_t n, int2* __restrict__ out) {
  for (int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; k < n; k += (int64_t)gridDim.x * blockDim.x) {
    const int2 eq = csrc[k];
    out[k] = make_int2(eclass[eq.x], eq.y);
  }
}

int class_sources(pyn_mesh* mesh, const NodeGraph* g, const int2** out) {
  if (mesh->nclass == mesh->nel) {                 // nothing merged: store index == element
    *out = g->csrc.p;
    return PYN_OK;
  }
  if (g->cls_version != mesh->class_version) {
    pyn_ctx* ctx = mesh->ctx;
    PYN_CHECK(g->csrc_cls.alloc(ctx, g->ncontrib));
    resolve_class_kernel<<<grid_for(ctx, g->ncontrib, 256), 256, 0, ctx->stream>>>(
        g->csrc.p, mesh->elem_class.p, g->ncontrib, g->csrc_cls.p);
    PYN_LAUNCHED(ctx);
    g->cls_version = mesh->class_version;
  }
  *out = g->csrc_cls.p;
  return PYN_OK;
}

// Patterns of four matrices filtered from the same graph (all with row classes) in one pair of
// passes; same result as four make_pattern calls.
static int make_patterns4(pyn_mesh* mesh, const MatSpec* sp, pyn_mat** m) {
  pyn_ctx* ctx = mesh->ctx;
  const NodeGraph* graph = sp[0].graph;
  PYN_REQUIRE(mesh->has_cls, "dof classes not set (pyn_mesh_set_dof_class)");
  DevBuf<int32_t> rowlen[4];
  DevBuf<unsigned long long> nreal;
  PYN_CHECK(nreal.alloc(ctx, 4));
  PYN_CHECK(nreal.zero());
  FilterTargets T;
  for (int k = 0; k < 4; ++k) {
    pyn_mat* a = m[k];
    a->ctx = ctx;
    a->mesh = mesh;
    a->br = sp[k].br;
    a->bc = sp[k].bc;
    a->nbrows = a->nbcols = graph->nrows;
    a->graph = graph;
    a->use_rowcls = true;
    a->use_colcls = sp[k].use_colcls;
    a->pairmask = sp[k].pairmask;
    a->diagmask = sp[k].diagmask;
    a->ref_dense = sp[k].ref_dense;
    PYN_CHECK(rowlen[k].alloc(ctx, graph->nrows));
    PYN_CHECK(a->rowptr.alloc(ctx, graph->nrows + 1));
    T.t[k].bc = sp[k].bc;
    T.t[k].use_colcls = sp[k].use_colcls ? 1 : 0;
    T.t[k].pairmask = sp[k].pairmask;
    T.t[k].diagmask = sp[k].diagmask;
    T.t[k].rowlen = rowlen[k].p;
    T.t[k].nreal = nreal.p + k;
    T.t[k].rowptr = nullptr;
    T.t[k].colidx = nullptr;
    T.t[k].base = nullptr;
  }
  const int grid = grid_for(ctx, graph->nrows * 32, 256);
  if (sp[0].br == 2)
    filter4_kernel<0, 2><<<grid, 256, 0, ctx->stream>>>(graph->nrows, graph->rowptr.p, graph->colidx.p,
                                                        mesh->nodecls.p, T);
  else
    filter4_kernel<0, 3><<<grid, 256, 0, ctx->stream>>>(graph->nrows, graph->rowptr.p, graph->colidx.p,
                                                        mesh->nodecls.p, T);
  PYN_LAUNCHED(ctx);
  unsigned long long real_h[4] = {0, 0, 0, 0};
  PYN_CHECK(nreal.download(real_h, 4));
  for (int k = 0; k < 4; ++k) {
    pyn_mat* a = m[k];
    a->nblocks_real = (int64_t)real_h[k];
    PYN_CHECK(exclusive_scan_i32_to_i64(ctx, rowlen[k].p, a->rowptr.p, graph->nrows, &a->nblocks));
    PYN_CHECK(a->colidx.alloc(ctx, a->nblocks));
    PYN_CHECK(a->base_slot.alloc(ctx, a->nblocks));
    PYN_CHECK(a->val.alloc(ctx, a->nblocks * a->br * a->bc));
    T.t[k].rowptr = a->rowptr.p;
    T.t[k].colidx = a->colidx.p;
    T.t[k].base = a->base_slot.p;
  }
  if (sp[0].br == 2)
    filter4_kernel<1, 2><<<grid, 256, 0, ctx->stream>>>(graph->nrows, graph->rowptr.p, graph->colidx.p,
                                                        mesh->nodecls.p, T);
  else
    filter4_kernel<1, 3><<<grid, 256, 0, ctx->stream>>>(graph->nrows, graph->rowptr.p, graph->colidx.p,
                                                        mesh->nodecls.p, T);
  PYN_LAUNCHED(ctx);
  return PYN_OK;
}

static int new_matrix(pyn_mesh* mesh, const MatSpec& sp, pyn_mat** out) {
  pyn_mat* m = new pyn_mat();
  int rc = make_pattern(mesh, sp.graph, sp.br, sp.bc, sp.use_rowcls, sp.use_colcls, sp.pairmask,
                        sp.diagmask, m);
  if (rc != PYN_OK) {
    delete m;
    return rc;
  }
  m->ref_dense = sp.ref_dense;
  *out = m;
  return PYN_OK;
}

int assemble_matrix(pyn_mesh* mesh, const MatSpec& sp, pyn_mat** out) {
  pyn_ctx* ctx = mesh->ctx;
  pyn_mat* m = nullptr;
  PYN_CHECK(new_matrix(mesh, sp, &m));
  *out = m;
  DiagAdd da;
  for (int i = 0; i < 4; ++i) da.v[i] = sp.diag_add[i];
  const int2* src = nullptr;
  PYN_CHECK(class_sources(mesh, sp.graph, &src));
  if (m->nblocks > 0) {
    // lanes per row: the stencil graph has 3 ngl - 2 .. ~20 blocks per row, the dense graph 27 .. 343+
    const int64_t avg = m->nblocks / std::max<int64_t>(m->nbrows, 1);
    const int G = avg <= 6 ? 8 : (avg <= 20 ? 16 : 32);
    const int grid = grid_for(ctx, m->nbrows * G, 256);
#define CALL(BR, BC)                                                                              \
  if (G == 8) numeric_kernel<BR, BC, 8> CALL_ARGS; else if (G == 16) numeric_kernel<BR, BC, 16> CALL_ARGS; else numeric_kernel<BR, BC, 32> CALL_ARGS
#define CALL_ARGS <<<grid, 256, 0, ctx->stream>>>(                                          \
      m->nbrows, m->rowptr.p, m->colidx.p, m->base_slot.p, sp.graph->cptr.p, src,    \
      sp.ebs, sp.ebs_pairs, sp.ebs_ppa, sp.ebs_stride ? sp.ebs_stride : sp.br * sp.bc, sp.ebs_off,  \
      sp.pair_map, sp.sign, mesh->dofcls.p, sp.use_rowcls, sp.use_colcls, \
      sp.pairmask, sp.diagmask, da, sp.row_scale, m->val.p)
    prof_begin(ctx, PROF_NUMERIC);
    PYN_DISPATCH_BLOCK(sp.br, sp.bc, CALL);
#undef CALL
#undef CALL_ARGS
    prof_end(ctx);
    PYN_LAUNCHED(ctx);
  }
  return PYN_OK;
}

// ---- K, Krhs, Rw, Rd in one pass over the dense node graph -------------------------------
// The four matrices of MatFS.buildFS / the free rows of MatNS.buildNS take their values from
// the same element blocks at the same node pairs, so one walk over the sorted COO feeds all
// of them: the contribution list (8 B per contribution), the segment pointers and the column
// ids are read once instead of once per matrix, and no slot -> block table is needed -- a
// matrix keeps the slots of a row in graph order (filter_kernel), so the position of a kept
// block is a running count over the warp's ballots.
struct KleTarget {
  const int64_t* rowptr;     // of the matrix (NULL: matrix not wanted)
  double* val;
  uint32_t pairmask, diagmask;
  double sign;
  double dadd[4];
};
struct KleTargets {
  KleTarget K, Krhs, Rw, Rd;
};

template <int BR, int BC, bool COLCLS>
__device__ __forceinline__ void kle_emit(const KleTarget& t, bool valid, bool diag, unsigned rpack,
                                         unsigned cpack, const double* acc, int lane, int64_t p0,
                                         int nb, int& out) {
  constexpr int BB = BR * BC;
  unsigned okbits = 0, dbits = 0;                    // entries taken from the block / diagonal fix-ups
#pragma unroll
  for (int i = 0; i < BR; ++i) {
    const int rc = dof_class(rpack, i);
#pragma unroll
    for (int j = 0; j < BC; ++j) {
      const int cc = COLCLS ? dof_class(cpack, j) : 0;
      okbits |= ((t.pairmask >> (rc * 4 + cc)) & 1u) << (i * BC + j);
    }
    if (diag) dbits |= ((t.diagmask >> rc) & 1u) << i;
  }
  const bool keep = valid && (okbits | dbits);
  const unsigned bal = __ballot_sync(0xffffffffu, keep);
  if (keep) {
    const int u = out + __popc(bal & ((1u << lane) - 1u));
    double* dst = t.val + p0 * BB + u;
#pragma unroll
    for (int i = 0; i < BR; ++i)
#pragma unroll
      for (int j = 0; j < BC; ++j) {
        double v = ((okbits >> (i * BC + j)) & 1u) ? t.sign * acc[i * BC + j] : 0.0;
        if (BR == BC && i == j && ((dbits >> i) & 1u)) v += t.dadd[dof_class(rpack, i)];
        __stcs(dst + (int64_t)(i * BC + j) * nb, v);
      }
  }
  out += __popc(bal);
}

template <int BB>
__device__ __forceinline__ void kle_zero_pad(const KleTarget& t, int64_t p0, int nb, int out, int lane) {
  if (out + lane < nb) {
    double* dst = t.val + p0 * BB + (out + lane);
#pragma unroll
    for (int q = 0; q < BB; ++q, dst += nb) __stcs(dst, 0.0);
  }
}

// 32 bytes of a record, one sector, through the non-coherent path
__device__ __forceinline__ void load_sector(const double* p, double* v) {
  asm volatile("ld.global.nc.v4.f64 {%0,%1,%2,%3}, [%4];"
               : "=d"(v[0]), "=d"(v[1]), "=d"(v[2]), "=d"(v[3])
               : "l"(p));
}

// The walk over a row is software-pipelined over its chunks of 32 slots, because a warp
// issues in order and the gather is a chain of dependent loads:
//   stage A (two chunks ahead): column id, segment pointers            -- no dependence
//   stage B (one chunk ahead):  first two contributions, column classes -- needs A
//   stage C (this chunk):       the records, sum, stores                -- needs B
// so that only the record loads of stage C are waited for.
// I: int32_t when the contribution count and the record offsets fit 31 bits (half the index
// arithmetic and registers), else int64_t
template <typename I>
struct SlotA {
  int32_t c;
  I k0, k1;
};
template <typename I>
struct SlotB {
  int32_t c;
  unsigned cpack;
  int cnt;
  I k0;
  int2 e0, e1;        // (store index, pair) of the first two contributions
};
template <typename I>
__device__ __forceinline__ SlotA<I> slot_stage_a(bool valid, int64_t s, const int32_t* __restrict__ g_colidx,
                                                 const int64_t* __restrict__ cptr) {
  SlotA<I> a;
  a.c = 0;
  a.k0 = a.k1 = 0;
  if (valid) {
    a.c = g_colidx[s];
    a.k0 = (I)cptr[s];
    a.k1 = (I)cptr[s + 1];
  }
  return a;
}
template <typename I>
__device__ __forceinline__ SlotB<I> slot_stage_b(const SlotA<I>& a, const int2* __restrict__ csrc,
                                                 const uint8_t* __restrict__ nodecls) {
  SlotB<I> b;
  b.c = a.c;
  b.k0 = a.k0;
  b.cnt = (int)(a.k1 - a.k0);
  b.cpack = 0;
  b.e0 = b.e1 = make_int2(0, 0);
  if (b.cnt > 0) {
    b.e0 = csrc[a.k0];
    if (b.cnt > 1) b.e1 = csrc[a.k0 + 1];
    b.cpack = nodecls[a.c];
  }
  return b;
}

// acc[0..H) = sum over the contributions of the slot of H doubles at offset OFF of their
// records, ascending element (fixed summation order)
template <int RS, int OFF, int H, typename I>
__device__ __forceinline__ void gather_half(const SlotB<I>& b, const int2* __restrict__ csrc,
                                            const double* __restrict__ ebs, I npairs, double* acc) {
  const double* r0 = ebs + ((I)b.e0.x * npairs + b.e0.y) * RS + OFF;
#pragma unroll
  for (int q = 0; q < H; q += 4) load_sector(r0 + q, acc + q);
  if (b.cnt > 1) {
    const double* r1 = ebs + ((I)b.e1.x * npairs + b.e1.y) * RS + OFF;
    double v[H];
#pragma unroll
    for (int q = 0; q < H; q += 4) load_sector(r1 + q, v + q);
#pragma unroll
    for (int q = 0; q < H; ++q) acc[q] += v[q];
    for (int i = 2; i < b.cnt; ++i) {
      const int2 eq = csrc[b.k0 + i];
      const double* rec = ebs + ((I)eq.x * npairs + eq.y) * RS + OFF;
#pragma unroll
      for (int q = 0; q < H; q += 4) load_sector(rec + q, v + q);
#pragma unroll
      for (int q = 0; q < H; ++q) acc[q] += v[q];
    }
  }
}

// csrc: contributions as (index in the element block store, pair) -- class_sources()
template <int D, int DW, int MINB, typename I>
__global__ void __launch_bounds__(256, MINB)
numeric_kle_kernel(int64_t nrows, const int64_t* __restrict__ g_rowptr,
                   const int32_t* __restrict__ g_colidx, const int64_t* __restrict__ cptr,
                   const int2* __restrict__ csrc, const double* __restrict__ ebs, int64_t npairs64,
                   const uint8_t* __restrict__ nodecls, KleTargets T) {
  const I npairs = (I)npairs64;
  using Rec = KleRecord<D>;
  constexpr int RS = Rec::RS, H = Rec::HALF;
  const int lane = threadIdx.x & 31;
  const int64_t warp = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
  for (int64_t r = warp; r < nrows; r += nwarps) {
    const int64_t s0 = g_rowptr[r];
    const int len = (int)(g_rowptr[r + 1] - s0);
    int64_t pK = 0, pKr = 0, pW = 0, pD = 0;
    int nK = 0, nKr = 0, nW = 0, nD = 0;
    if (T.K.rowptr) { pK = T.K.rowptr[r]; nK = (int)(T.K.rowptr[r + 1] - pK); }
    if (T.Krhs.rowptr) { pKr = T.Krhs.rowptr[r]; nKr = (int)(T.Krhs.rowptr[r + 1] - pKr); }
    if (T.Rw.rowptr) { pW = T.Rw.rowptr[r]; nW = (int)(T.Rw.rowptr[r + 1] - pW); }
    if (T.Rd.rowptr) { pD = T.Rd.rowptr[r]; nD = (int)(T.Rd.rowptr[r + 1] - pD); }
    const unsigned rpack = nodecls[r];
    if (nK + nKr + nW + nD == 0) continue;         // e.g. a fully constrained row of Rw
    int oK = 0, oKr = 0, oW = 0, oD = 0;
    const SlotA<I> a1 = slot_stage_a<I>(lane < len, s0 + lane, g_colidx, cptr);
    SlotA<I> a2 = slot_stage_a<I>(32 + lane < len, s0 + 32 + lane, g_colidx, cptr);
    SlotB<I> cur = slot_stage_b<I>(a1, csrc, nodecls);
    for (int u0 = 0; u0 < len; u0 += 32) {
      const SlotA<I> a3 = slot_stage_a<I>(u0 + 64 + lane < len, s0 + u0 + 64 + lane, g_colidx, cptr);
      const SlotB<I> nxt = slot_stage_b<I>(a2, csrc, nodecls);
      const bool valid = cur.cnt > 0;
      const bool diag = valid && (int64_t)cur.c == r;
      double acc[H];
      if (nK + nKr) {                                // first half of the records: K
#pragma unroll
        for (int q = 0; q < H; ++q) acc[q] = 0.0;
        if (valid) gather_half<RS, Rec::OFF_K, H, I>(cur, csrc, ebs, npairs, acc);
        if (nK) kle_emit<D, D, true>(T.K, valid, diag, rpack, cur.cpack, acc, lane, pK, nK, oK);
        if (nKr) kle_emit<D, D, true>(T.Krhs, valid, diag, rpack, cur.cpack, acc, lane, pKr, nKr, oKr);
      }
      if (nW + nD) {                                 // second half: Rw, Rd
#pragma unroll
        for (int q = 0; q < H; ++q) acc[q] = 0.0;
        if (valid) gather_half<RS, Rec::OFF_RW, H, I>(cur, csrc, ebs, npairs, acc);
        if (nW) kle_emit<D, DW, false>(T.Rw, valid, diag, rpack, cur.cpack, acc, lane, pW, nW, oW);
        if (nD)
          kle_emit<D, 1, false>(T.Rd, valid, diag, rpack, cur.cpack, acc + (Rec::OFF_RD - Rec::OFF_RW), lane,
                                pD, nD, oD);
      }
      cur = nxt;
      a2 = a3;
    }
    // padding blocks of the row (rows hold a multiple of 4 blocks)
    kle_zero_pad<D * D>(T.K, pK, nK, oK, lane);
    kle_zero_pad<D * D>(T.Krhs, pKr, nKr, oKr, lane);
    kle_zero_pad<D * DW>(T.Rw, pW, nW, oW, lane);
    kle_zero_pad<D>(T.Rd, pD, nD, oD, lane);
  }
}

static KleTarget kle_target(const MatSpec& sp, const pyn_mat* m) {
  KleTarget t;
  t.rowptr = m->nblocks > 0 ? m->rowptr.p : nullptr;
  t.val = m->val.p;
  t.pairmask = sp.pairmask;
  t.diagmask = sp.diagmask;
  t.sign = sp.sign;
  for (int i = 0; i < 4; ++i) t.dadd[i] = sp.diag_add[i];
  return t;
}

// specs / out in the order K, Krhs, Rw, Rd; all on the dense graph with row classes, K and
// Krhs also with column classes (pyn_assemble_kle)
static int assemble_kle_fused_impl(pyn_mesh* mesh, const MatSpec* sp, pyn_mat** out);

int assemble_kle_fused(pyn_mesh* mesh, const MatSpec* sp, pyn_mat** out) {
  for (int i = 0; i < 4; ++i) out[i] = nullptr;
  const int rc = assemble_kle_fused_impl(mesh, sp, out);
  if (rc != PYN_OK)
    for (int i = 0; i < 4; ++i) {                  // nothing half-built leaves the library
      delete out[i];
      out[i] = nullptr;
    }
  return rc;
}

static int assemble_kle_fused_impl(pyn_mesh* mesh, const MatSpec* sp, pyn_mat** out) {
  pyn_ctx* ctx = mesh->ctx;
  const int d = mesh->dim;
  for (int i = 0; i < 4; ++i)
    PYN_REQUIRE(sp[i].graph == &mesh->dense && sp[i].use_rowcls && sp[i].ebs_ppa == 0 &&
                    !sp[i].pair_map && !sp[i].row_scale && sp[i].br == d &&
                    sp[i].use_colcls == (i < 2) && sp[i].ebs == sp[0].ebs &&
                    sp[i].ebs_stride == kle_record_stride(d),
                "assemble_kle_fused: unexpected matrix spec");
  for (int i = 0; i < 4; ++i) out[i] = new pyn_mat();
  PYN_CHECK(make_patterns4(mesh, sp, out));
  KleTargets T;
  T.K = kle_target(sp[0], out[0]);
  T.Krhs = kle_target(sp[1], out[1]);
  T.Rw = kle_target(sp[2], out[2]);
  T.Rd = kle_target(sp[3], out[3]);
  if (!T.K.rowptr && !T.Krhs.rowptr && !T.Rw.rowptr && !T.Rd.rowptr) return PYN_OK;
  const NodeGraph& g = mesh->dense;
  const int64_t npairs = (int64_t)mesh->nn * mesh->nn;
  const int2* src = nullptr;
  PYN_CHECK(class_sources(mesh, &g, &src));
  // compiled for 2 resident CTAs per SM (128 registers, no spills): measured fastest on B200
  // (3 CTAs at 80 registers spill the pipeline state: 0.93 against 0.78 ms at 16^3, ngl = 4;
  // two half passes K|Krhs and Rw|Rd: 0.82 ms)
  const int grid = grid_for(ctx, g.nrows * 32, 256, 2);   // one resident wave
  // 32-bit contribution / record indices when they fit
  const bool idx32 = !g_force_idx64 && g.ncontrib < (int64_t)0x7fffffff &&
                     mesh->nclass * npairs * kle_record_stride(d) < (int64_t)0x7fffffff;
#define NKLE_ARGS                                                                                   \
  <<<grid, 256, 0, ctx->stream>>>(g.nrows, g.rowptr.p, g.colidx.p, g.cptr.p, src, sp[0].ebs, npairs, \
                                  mesh->nodecls.p, T)
#define NKLE_CALL(D, DW, MB)                                              \
  do {                                                                    \
    if (idx32) numeric_kle_kernel<D, DW, MB, int32_t> NKLE_ARGS;          \
    else numeric_kle_kernel<D, DW, MB, int64_t> NKLE_ARGS;                \
  } while (0)
  prof_begin(ctx, PROF_NUMERIC);
  if (d == 2) NKLE_CALL(2, 1, 2);
  else NKLE_CALL(3, 3, 2);
#undef NKLE_CALL
#undef NKLE_ARGS
  prof_end(ctx);
  PYN_LAUNCHED(ctx);
  return PYN_OK;
}

}  // namespace pyn
