// graph.cu -- symbolic phase of the assembly (k4): node->element adjacency and
// the node graph with its sorted COO ("contribution") lists.
//
// Replaces DMPlexDom.getConnectivityNodes (dmplex.py:352-372) and the
// preallocation loops of mat_fs.py:21-113 / mat_ns.py:7-45.  For every row
// node a CTA generates the COO keys (col node, generation index) of all its
// contributions, sorts them with a shared-memory bitonic network, and emits
// the unique columns plus, per column, the contiguous segment of
// contributions ordered by element id -- i.e. the reference's cell-loop
// insertion order, so the numeric phase can reduce each segment in a fixed
// order without atomics.
#include "elem.cuh"

namespace pyn {

__global__ void n2e_count_kernel(const int32_t* __restrict__ elem_nodes, int64_t total,
                                 int32_t* __restrict__ cnt) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total;
       i += (int64_t)gridDim.x * blockDim.x)
    atomicAdd(&cnt[elem_nodes[i]], 1);
}

__global__ void n2e_fill_kernel(const int32_t* __restrict__ elem_nodes, int64_t total, int nn,
                                const int64_t* __restrict__ ptr, int32_t* __restrict__ cursor,
                                int2* __restrict__ n2e) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total;
       i += (int64_t)gridDim.x * blockDim.x) {
    const int32_t node = elem_nodes[i];
    const int32_t k = atomicAdd(&cursor[node], 1);
    n2e[ptr[node] + k] = make_int2((int)(i / nn), (int)(i % nn));
  }
}

// order every node's list by element id (fixed summation order); lists are short
__global__ void n2e_sort_kernel(const int64_t* __restrict__ ptr, int64_t nnodes,
                                int2* __restrict__ n2e) {
  for (int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; r < nnodes;
       r += (int64_t)gridDim.x * blockDim.x) {
    const int64_t b = ptr[r], e = ptr[r + 1];
    for (int64_t i = b + 1; i < e; ++i) {
      int2 v = n2e[i];
      int64_t j = i - 1;
      while (j >= b && (n2e[j].x > v.x || (n2e[j].x == v.x && n2e[j].y > v.y))) {
        n2e[j + 1] = n2e[j];
        --j;
      }
      n2e[j + 1] = v;
    }
  }
}

__global__ void max_i32_kernel(const int32_t* __restrict__ v, int64_t n, int32_t* __restrict__ out) {
  int32_t m = 0;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n;
       i += (int64_t)gridDim.x * blockDim.x)
    m = max(m, v[i]);
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) m = max(m, __shfl_xor_sync(0xffffffffu, m, o));
  if ((threadIdx.x & 31) == 0) atomicMax(out, m);
}

int build_n2e(pyn_mesh* mesh) {
  pyn_ctx* ctx = mesh->ctx;
  const int64_t total = mesh->nel * mesh->nn;
  DevBuf<int32_t> cnt, cursor, mx;
  PYN_CHECK(cnt.alloc(ctx, mesh->nnodes));
  PYN_CHECK(cursor.alloc(ctx, mesh->nnodes));
  PYN_CHECK(mx.alloc(ctx, 1));
  PYN_CHECK(cnt.zero());
  PYN_CHECK(cursor.zero());
  PYN_CHECK(mx.zero());
  PYN_CHECK(mesh->n2e_ptr.alloc(ctx, mesh->nnodes + 1));
  PYN_CHECK(mesh->n2e.alloc(ctx, total));
  const int grid = grid_for(ctx, total, 256);
  n2e_count_kernel<<<grid, 256, 0, ctx->stream>>>(mesh->elem_nodes.p, total, cnt.p);
  PYN_LAUNCHED(ctx);
  int64_t tot = 0;
  PYN_CHECK(exclusive_scan_i32_to_i64(ctx, cnt.p, mesh->n2e_ptr.p, mesh->nnodes, &tot));
  PYN_REQUIRE(tot == total, "n2e: count mismatch");
  n2e_fill_kernel<<<grid, 256, 0, ctx->stream>>>(mesh->elem_nodes.p, total, mesh->nn,
                                                 mesh->n2e_ptr.p, cursor.p, mesh->n2e.p);
  PYN_LAUNCHED(ctx);
  n2e_sort_kernel<<<grid_for(ctx, mesh->nnodes, 128), 128, 0, ctx->stream>>>(mesh->n2e_ptr.p,
                                                                             mesh->nnodes, mesh->n2e.p);
  PYN_LAUNCHED(ctx);
  max_i32_kernel<<<grid_for(ctx, mesh->nnodes, 256), 256, 0, ctx->stream>>>(cnt.p, mesh->nnodes, mx.p);
  PYN_LAUNCHED(ctx);
  int32_t h = 0;
  PYN_CHECK(mx.download(&h, 1));
  mesh->max_valence = h;
  return PYN_OK;
}

// ---- per-row sort + unique ------------------------------------------------------
__global__ void __launch_bounds__(256)
graph_row_kernel(const int64_t* __restrict__ n2e_ptr, const int2* __restrict__ n2e,
                 const int32_t* __restrict__ elem_nodes, const int32_t* __restrict__ pair_b, int nn,
                 int ppa, int2* __restrict__ csrc, int32_t* __restrict__ tmpcol,
                 int32_t* __restrict__ chead, int32_t* __restrict__ rowlen) {
  extern __shared__ unsigned long long keys[];
  __shared__ int thread_heads[256];
  const int64_t r = blockIdx.x;
  const int tid = threadIdx.x, nth = blockDim.x;
  const int64_t n0 = n2e_ptr[r];
  const int cnt = (int)(n2e_ptr[r + 1] - n0) * ppa;
  const int64_t cand0 = n0 * ppa;
  int np2 = 1;
  while (np2 < cnt) np2 <<= 1;
  for (int idx = tid; idx < np2; idx += nth) {
    unsigned long long key = ~0ULL;
    if (idx < cnt) {
      const int i = idx / ppa, m = idx - i * ppa;
      const int2 ea = n2e[n0 + i];
      const int b = pair_b[ea.y * ppa + m];
      const unsigned int c = (unsigned int)elem_nodes[(int64_t)ea.x * nn + b];
      key = ((unsigned long long)c << 32) | (unsigned int)idx;
    }
    keys[idx] = key;
  }
  __syncthreads();
  for (int k = 2; k <= np2; k <<= 1) {
    for (int j = k >> 1; j > 0; j >>= 1) {
      for (int i = tid; i < np2; i += nth) {
        const int ixj = i ^ j;
        if (ixj > i) {
          const unsigned long long x = keys[i], y = keys[ixj];
          const bool up = (i & k) == 0;
          if ((x > y) == up) {
            keys[i] = y;
            keys[ixj] = x;
          }
        }
      }
      __syncthreads();
    }
  }
  // contiguous chunk per thread: count heads, scan, then emit
  const int chunk = (cnt + nth - 1) / nth;
  const int lo = min(cnt, tid * chunk), hi = min(cnt, lo + chunk);
  int heads = 0;
  for (int p = lo; p < hi; ++p) {
    const unsigned int c = (unsigned int)(keys[p] >> 32);
    if (p == 0 || c != (unsigned int)(keys[p - 1] >> 32)) ++heads;
  }
  thread_heads[tid] = heads;
  __syncthreads();
  if (tid == 0) {
    int run = 0;
    for (int t = 0; t < nth; ++t) {
      const int h = thread_heads[t];
      thread_heads[t] = run;
      run += h;
    }
    rowlen[r] = run;
  }
  __syncthreads();
  int rank = thread_heads[tid];
  for (int p = lo; p < hi; ++p) {
    const unsigned long long key = keys[p];
    const unsigned int c = (unsigned int)(key >> 32);
    const int idx = (int)(key & 0xffffffffu);
    const int i = idx / ppa, m = idx - i * ppa;
    const int2 ea = n2e[n0 + i];
    csrc[cand0 + p] = make_int2(ea.x, ea.y * ppa + m);
    if (p == 0 || c != (unsigned int)(keys[p - 1] >> 32)) {
      tmpcol[cand0 + rank] = (int32_t)c;
      chead[cand0 + rank] = p;
      ++rank;
    }
  }
}

__global__ void graph_compact_kernel(const int64_t* __restrict__ n2e_ptr, int ppa,
                                     const int64_t* __restrict__ rowptr,
                                     const int32_t* __restrict__ tmpcol,
                                     const int32_t* __restrict__ chead, int64_t nrows,
                                     int64_t ncontrib, int32_t* __restrict__ colidx,
                                     int64_t* __restrict__ cptr) {
  const int lane = threadIdx.x & 31;
  const int64_t warp = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
  for (int64_t r = warp; r < nrows; r += nwarps) {
    const int64_t p0 = rowptr[r], len = rowptr[r + 1] - p0, cand0 = n2e_ptr[r] * ppa;
    for (int64_t u = lane; u < len; u += 32) {
      colidx[p0 + u] = tmpcol[cand0 + u];
      cptr[p0 + u] = cand0 + chead[cand0 + u];
    }
    if (r == nrows - 1 && lane == 0) cptr[p0 + len] = ncontrib;
  }
}

int build_graph(pyn_mesh* mesh, NodeGraph* g, const std::vector<int32_t>& pair_b, int ppa) {
  pyn_ctx* ctx = mesh->ctx;
  const int nn = mesh->nn;
  PYN_REQUIRE((int)pair_b.size() == nn * ppa, "build_graph: pair table size");
  g->ppa = ppa;
  g->nrows = mesh->nnodes;
  g->ncontrib = mesh->nel * nn * ppa;
  PYN_CHECK(g->pair_b.alloc(ctx, (int64_t)nn * ppa));
  PYN_CHECK(g->pair_b.upload(pair_b.data(), (int64_t)nn * ppa));
  PYN_CHECK(g->csrc.alloc(ctx, g->ncontrib));
  DevBuf<int32_t> tmpcol, chead, rowlen, mx;
  PYN_CHECK(tmpcol.alloc(ctx, g->ncontrib));
  PYN_CHECK(chead.alloc(ctx, g->ncontrib));
  PYN_CHECK(rowlen.alloc(ctx, g->nrows));
  PYN_CHECK(mx.alloc(ctx, 1));
  PYN_CHECK(mx.zero());
  int64_t maxc = (int64_t)mesh->max_valence * ppa;
  int64_t np2 = 1;
  while (np2 < maxc) np2 <<= 1;
  const size_t smem = (size_t)np2 * sizeof(unsigned long long);
  PYN_REQUIRE(smem <= ctx->smem_optin - 2048,
              "node valence %d x %d pairs needs %zu B of shared memory for the row sort",
              mesh->max_valence, ppa, smem);
  PYN_CUDA(cudaFuncSetAttribute(graph_row_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  PYN_REQUIRE(g->nrows < (int64_t)2147483647, "too many nodes");
  graph_row_kernel<<<(unsigned)g->nrows, 256, smem, ctx->stream>>>(
      mesh->n2e_ptr.p, mesh->n2e.p, mesh->elem_nodes.p, g->pair_b.p, nn, ppa, g->csrc.p, tmpcol.p,
      chead.p, rowlen.p);
  PYN_LAUNCHED(ctx);
  PYN_CHECK(g->rowptr.alloc(ctx, g->nrows + 1));
  PYN_CHECK(exclusive_scan_i32_to_i64(ctx, rowlen.p, g->rowptr.p, g->nrows, &g->nslots));
  max_i32_kernel<<<grid_for(ctx, g->nrows, 256), 256, 0, ctx->stream>>>(rowlen.p, g->nrows, mx.p);
  PYN_LAUNCHED(ctx);
  PYN_CHECK(g->colidx.alloc(ctx, g->nslots));
  PYN_CHECK(g->cptr.alloc(ctx, g->nslots + 1));
  graph_compact_kernel<<<grid_for(ctx, g->nrows * 32, 256), 256, 0, ctx->stream>>>(
      mesh->n2e_ptr.p, ppa, g->rowptr.p, tmpcol.p, chead.p, g->nrows, g->ncontrib, g->colidx.p,
      g->cptr.p);
  PYN_LAUNCHED(ctx);
  int32_t h = 0;
  PYN_CHECK(mx.download(&h, 1));
  g->max_row = h;
  return PYN_OK;
}

// grid-line ("stencil") pairs of the collocated operators: b shares all but at
// most one tensor index with a.
int ensure_stencil(pyn_mesh* mesh) {
  if (mesh->has_stencil) return PYN_OK;
  const int nn = mesh->nn, dim = mesh->dim, ngl = mesh->ngl;
  const int ppa = dim * (ngl - 1) + 1;
  std::vector<int32_t> pb((size_t)nn * ppa), sod((size_t)nn * nn, -1);
  for (int a = 0; a < nn; ++a) {
    int m = 0;
    for (int b = 0; b < nn; ++b) {
      int diff = 0;
      for (int d = 0; d < dim; ++d) diff += mesh->htab.node_t[a][d] != mesh->htab.node_t[b][d];
      if (diff <= 1) {
        PYN_REQUIRE(m < ppa, "stencil: pair count");
        pb[(size_t)a * ppa + m] = b;
        sod[(size_t)a * nn + b] = a * ppa + m;
        ++m;
      }
    }
    PYN_REQUIRE(m == ppa, "stencil: pair count");
  }
  PYN_CHECK(build_graph(mesh, &mesh->stencil, pb, ppa));
  PYN_CHECK(mesh->stencil_of_dense.alloc(mesh->ctx, (int64_t)nn * nn));
  PYN_CHECK(mesh->stencil_of_dense.upload(sod.data(), (int64_t)nn * nn));
  PYN_CUDA(cudaStreamSynchronize(mesh->ctx->stream));
  mesh->has_stencil = true;
  return PYN_OK;
}

}  // namespace pyn
