// asm_pc.cu -- element-block additive Schwarz preconditioner for the Krylov solve of the KLE
// system (PETSc -pc_type asm with one subdomain per spectral element; the reference takes its
// preconditioner from the options database, kle_solver.py:49-64).
//
//   z = sum_e R_e^T W_e inv(K_e) W_e R_e r,   K_e = R_e K R_e^T,   W = diag(1/sqrt(valence))
//
// Set-up (once per operator): every element block K_e is read out of the assembled row-plane
// BSR and hashed on the device; elements with equal hashes share one class; the dense block of
// one representative per class is extracted, inverted on the host (27 blocks of 192^2 on the
// uniform 3-D box at ngl = 4) and uploaded.
// Apply (every Krylov iteration), two launches:
//   asm_mult_kernel  one CTA per tile of 16 elements of a class: gather W r into shared memory,
//                    Y = inv(K_c) X as a register-tiled fp64 product (the inverse streams from
//                    L2), element-local results to ybuf
//   asm_sum_kernel   per node: sum of the element-local results in ascending element order
//                    (fixed order: run-to-run reproducible, no atomics), optionally r.z and z.z
#include "asm_pc.cuh"
#include "asm_host.cuh"
#include "vec.cuh"

namespace pyn {

constexpr int kAsmKC = 16;         // k-chunk of the inverse staged in shared memory
constexpr int kAsmRG = 64;         // row groups (threads along the rows of a panel)
constexpr int kAsmRM = 3;          // rows per thread
constexpr int kAsmRP = kAsmRG * kAsmRM;   // rows per panel
constexpr int kAsmEM = 4;          // elements per thread (kAsmTile / 4 thread columns)
static_assert(kAsmTile == 4 * kAsmEM, "thread layout: 4 element groups x 64 row groups");

// MODE 0: hash the block of every element of the list into hash[e][2]
// MODE 1: write the dense block of list element li to dense[li][ld][ld]
// One warp per (list element, local row node).
template <int B, int MODE>
__global__ void __launch_bounds__(256)
asm_blocks_kernel(int64_t nlist, const int32_t* __restrict__ list, const int32_t* __restrict__ elem_nodes,
                  int nn, const int64_t* __restrict__ rowptr, const int32_t* __restrict__ colidx,
                  const double* __restrict__ val, unsigned long long* __restrict__ hash,
                  double* __restrict__ dense) {
  constexpr int BB = B * B;
  const int ld = nn * B;
  const int lane = threadIdx.x & 31;
  const int64_t warp = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
  for (int64_t w = warp; w < nlist * nn; w += nwarps) {
    const int64_t li = w / nn;
    const int a = (int)(w - li * nn);
    const int64_t e = list ? (int64_t)list[li] : li;
    const int32_t* en = elem_nodes + e * nn;
    const int64_t r = en[a];
    const int64_t p0 = rowptr[r];
    const int nb = (int)(rowptr[r + 1] - p0);
    unsigned long long h0 = 0ull, h1 = 0ull;
    for (int b = lane; b < nn; b += 32) {
      const int32_t c = en[b];
      int lo = 0, hi = nb;                       // first block of the row with column >= c
      while (lo < hi) {
        const int mid = (lo + hi) >> 1;
        if (colidx[p0 + mid] < c) lo = mid + 1;
        else hi = mid;
      }
      const bool found = lo < nb && colidx[p0 + lo] == c;
#pragma unroll
      for (int q = 0; q < BB; ++q) {
        const double v = found ? val[p0 * BB + (int64_t)q * nb + lo] : 0.0;
        const int i = q / B, j = q - i * B;
        const int64_t pos = (int64_t)(a * B + i) * ld + (b * B + j);
        if (MODE == 0) {
          if (v != 0.0) {
            const unsigned long long bits = (unsigned long long)__double_as_longlong(v);
            h0 += asm_hash_entry(bits, (uint64_t)pos, 0);
            h1 += asm_hash_entry(bits, (uint64_t)pos, 1);
          }
        } else {
          dense[li * (int64_t)ld * ld + pos] = v;
        }
      }
    }
    if (MODE == 0) {
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) {
        h0 += __shfl_xor_sync(0xffffffffu, h0, o);
        h1 += __shfl_xor_sync(0xffffffffu, h1, o);
      }
      if (lane == 0) {                           // wrapping integer sums: order-independent
        atomicAdd(hash + 2 * e, h0);
        atomicAdd(hash + 2 * e + 1, h1);
      }
    }
  }
}

__global__ void asm_weights_kernel(int64_t nnodes, const int64_t* __restrict__ n2e_ptr,
                                   double* __restrict__ wnode) {
  for (int64_t n = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; n < nnodes;
       n += (int64_t)gridDim.x * blockDim.x) {
    const double m = (double)(n2e_ptr[n + 1] - n2e_ptr[n]);
    wnode[n] = m > 0.0 ? 1.0 / sqrt(m) : 0.0;
  }
}

// Y[:, el] = inv(K_c) (W r)[dofs of el] for the 16 elements of a tile; 256 threads = 4 element
// groups (4 elements each) x 64 row groups (3 rows each, 64 apart) per panel of 192 rows.
// Shared memory: Xs[ld][16] gathered right-hand sides, As[16][192] one k-chunk of the inverse
// (symmetric: row k of the inverse is read for column k).
template <int B>
__global__ void __launch_bounds__(256)
asm_mult_kernel(int64_t ntiles, const int32_t* __restrict__ tile_elems, const int32_t* __restrict__ tile_class,
                const int32_t* __restrict__ elem_nodes, int nn, const double* __restrict__ inv,
                const double* __restrict__ wnode, const double* __restrict__ r, double* __restrict__ ybuf,
                const double* __restrict__ done) {
  extern __shared__ __align__(16) double asm_sm[];
  if (done && *done != 0.0) return;
  const int ld = nn * B;
  double* Xs = asm_sm;                              // [ld][kAsmTile]
  double* As = asm_sm + (size_t)ld * kAsmTile;      // [kAsmKC][kAsmRP]
  const int tid = threadIdx.x;
  const int tx = tid & 3, ty = tid >> 2;
  for (int64_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
    const int32_t* te = tile_elems + tile * kAsmTile;
    const double* A = inv + (int64_t)tile_class[tile] * ld * ld;
    __syncthreads();                                // the previous tile is done with Xs
    for (int idx = tid; idx < nn * kAsmTile; idx += 256) {
      const int el = idx % kAsmTile, a = idx / kAsmTile;
      const int e = te[el];
      double x[B];
#pragma unroll
      for (int c = 0; c < B; ++c) x[c] = 0.0;
      if (e >= 0) {
        const int64_t node = elem_nodes[(int64_t)e * nn + a];
        const double w = wnode[node];
#pragma unroll
        for (int c = 0; c < B; ++c) x[c] = w * r[node * B + c];
      }
#pragma unroll
      for (int c = 0; c < B; ++c) Xs[(a * B + c) * kAsmTile + el] = x[c];
    }
    for (int row0 = 0; row0 < ld; row0 += kAsmRP) {
      const int prows = ld - row0 < kAsmRP ? ld - row0 : kAsmRP;
      double acc[kAsmRM][kAsmEM];
#pragma unroll
      for (int m = 0; m < kAsmRM; ++m)
#pragma unroll
        for (int j = 0; j < kAsmEM; ++j) acc[m][j] = 0.0;
      for (int k0 = 0; k0 < ld; k0 += kAsmKC) {
        __syncthreads();                            // As free again; Xs complete on the first pass
        for (int idx = tid; idx < kAsmKC * kAsmRP; idx += 256) {
          const int kk = idx / kAsmRP, i = idx - kk * kAsmRP;
          As[idx] = (k0 + kk < ld && i < prows) ? A[(int64_t)(k0 + kk) * ld + row0 + i] : 0.0;
        }
        __syncthreads();
        const int kmax = ld - k0 < kAsmKC ? ld - k0 : kAsmKC;
        for (int kk = 0; kk < kmax; ++kk) {
          const double2 xa = *reinterpret_cast<const double2*>(Xs + (k0 + kk) * kAsmTile + tx * kAsmEM);
          const double2 xb = *reinterpret_cast<const double2*>(Xs + (k0 + kk) * kAsmTile + tx * kAsmEM + 2);
#pragma unroll
          for (int m = 0; m < kAsmRM; ++m) {
            const double av = As[kk * kAsmRP + ty + kAsmRG * m];
            acc[m][0] += av * xa.x;
            acc[m][1] += av * xa.y;
            acc[m][2] += av * xb.x;
            acc[m][3] += av * xb.y;
          }
        }
      }
#pragma unroll
      for (int m = 0; m < kAsmRM; ++m) {
        const int i = ty + kAsmRG * m;
        if (i < prows) {
#pragma unroll
          for (int j = 0; j < kAsmEM; ++j) {
            const int e = te[tx * kAsmEM + j];
            if (e >= 0) {
              const int64_t node = elem_nodes[(int64_t)e * nn + (row0 + i) / B];
              ybuf[(int64_t)e * ld + row0 + i] = acc[m][j] * wnode[node];
            }
          }
        }
      }
    }
  }
}

// z[node] = sum over the elements of the node, ascending, of their local results
template <int B>
__global__ void __launch_bounds__(256)
asm_sum_kernel(int64_t nnodes, const int64_t* __restrict__ n2e_ptr, const int2* __restrict__ n2e,
               const double* __restrict__ ybuf, int ld, const double* __restrict__ r, double* __restrict__ z,
               double* partials, unsigned int* ticket, double* dots, const double* __restrict__ done) {
  __shared__ double sm[64];
  if (done && *done != 0.0) return;
  double v[2] = {0.0, 0.0};
  for (int64_t n = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; n < nnodes;
       n += (int64_t)gridDim.x * blockDim.x) {
    double zz[B];
#pragma unroll
    for (int c = 0; c < B; ++c) zz[c] = 0.0;
    for (int64_t k = n2e_ptr[n]; k < n2e_ptr[n + 1]; ++k) {
      const int2 ea = n2e[k];
      const double* y = ybuf + (int64_t)ea.x * ld + ea.y * B;
#pragma unroll
      for (int c = 0; c < B; ++c) zz[c] += y[c];
    }
#pragma unroll
    for (int c = 0; c < B; ++c) {
      z[n * B + c] = zz[c];
      if (dots) {
        v[0] += r[n * B + c] * zz[c];
        v[1] += zz[c] * zz[c];
      }
    }
  }
  if (dots) {
    block_reduce<2>(v, sm);
    grid_reduce_finish<2>(v, partials, ticket, dots);
  }
}

template <int B>
static int asm_setup_impl(pyn_ctx* ctx, const pyn_mat* A, AsmPc* pc) {
  pyn_mesh* mesh = A->mesh;
  const int nn = mesh->nn, ld = nn * B;
  const int64_t nel = mesh->nel;
  pc->mesh = mesh;
  pc->B = B;
  pc->ld = ld;
  // 1. hash every element block
  DevBuf<unsigned long long> hash;
  PYN_CHECK(hash.alloc(ctx, 2 * nel));
  PYN_CHECK(hash.zero());
  const int grid = grid_for(ctx, nel * nn * 32, 256);
  asm_blocks_kernel<B, 0><<<grid, 256, 0, ctx->stream>>>(nel, nullptr, mesh->elem_nodes.p, nn, A->rowptr.p,
                                                         A->colidx.p, A->val.p, hash.p, nullptr);
  PYN_LAUNCHED(ctx);
  std::vector<unsigned long long> hh((size_t)(2 * nel));
  PYN_CHECK(hash.download(hh.data(), 2 * nel));
  // 2. classes of equal blocks, their representatives
  std::vector<int32_t> cls, rep;
  static_assert(sizeof(unsigned long long) == sizeof(uint64_t), "hash word");
  pc->nclass = asm_group_by_hash(nel, reinterpret_cast<const uint64_t*>(hh.data()), &cls, &rep);
  const int64_t inv_bytes = pc->nclass * (int64_t)ld * ld * (int64_t)sizeof(double);
  PYN_REQUIRE(inv_bytes <= kAsmMaxInverseBytes,
              "pc asm: %lld distinct element blocks of %d^2 (%lld MB of inverses): the elements of this mesh "
              "are not congruent enough for the element-block Schwarz preconditioner; use bjacobi",
              (long long)pc->nclass, ld, (long long)(inv_bytes >> 20));
  // 3. dense blocks of the representatives -> host -> inverse -> device
  DevBuf<int32_t> drep;
  PYN_CHECK(drep.alloc(ctx, pc->nclass));
  PYN_CHECK(drep.upload(rep.data(), pc->nclass));
  PYN_CHECK(pc->inv.alloc(ctx, pc->nclass * (int64_t)ld * ld));
  asm_blocks_kernel<B, 1><<<grid_for(ctx, pc->nclass * nn * 32, 256), 256, 0, ctx->stream>>>(
      pc->nclass, drep.p, mesh->elem_nodes.p, nn, A->rowptr.p, A->colidx.p, A->val.p, nullptr, pc->inv.p);
  PYN_LAUNCHED(ctx);
  std::vector<double> blocks((size_t)(pc->nclass * (int64_t)ld * ld));
  PYN_CHECK(pc->inv.download(blocks.data(), (int64_t)blocks.size()));
  for (int64_t c = 0; c < pc->nclass; ++c) {
    double* a = blocks.data() + c * (int64_t)ld * ld;
    PYN_REQUIRE(asm_dense_inverse(ld, a) == 0, "pc asm: the block of element %d is singular", (int)rep[(size_t)c]);
    asm_symmetrise(ld, a);
  }
  PYN_CHECK(pc->inv.upload(blocks.data(), (int64_t)blocks.size()));
  // 4. tiles of elements of one class
  std::vector<int32_t> tile_elems, tile_class;
  pc->ntiles = asm_make_tiles(nel, cls.data(), pc->nclass, kAsmTile, &tile_elems, &tile_class);
  PYN_CHECK(pc->tile_elems.alloc(ctx, (int64_t)tile_elems.size()));
  PYN_CHECK(pc->tile_elems.upload(tile_elems.data(), (int64_t)tile_elems.size()));
  PYN_CHECK(pc->tile_class.alloc(ctx, pc->ntiles));
  PYN_CHECK(pc->tile_class.upload(tile_class.data(), pc->ntiles));
  // 5. weights, element-local result buffer
  PYN_CHECK(pc->wnode.alloc(ctx, mesh->nnodes));
  asm_weights_kernel<<<grid_for(ctx, mesh->nnodes, 256), 256, 0, ctx->stream>>>(mesh->nnodes, mesh->n2e_ptr.p,
                                                                                pc->wnode.p);
  PYN_LAUNCHED(ctx);
  PYN_CHECK(pc->ybuf.alloc(ctx, nel * ld));
  pc->smem = sizeof(double) * ((size_t)ld * kAsmTile + (size_t)kAsmKC * kAsmRP);
  PYN_REQUIRE(pc->smem <= ctx->smem_optin, "pc asm: %zu B of shared memory per CTA needed, %zu available",
              pc->smem, ctx->smem_optin);
  PYN_CUDA(cudaFuncSetAttribute(asm_mult_kernel<B>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)pc->smem));
  PYN_CUDA(cudaStreamSynchronize(ctx->stream));     // the host vectors above leave scope
  return PYN_OK;
}

int asm_setup(pyn_ctx* ctx, const pyn_mat* A, AsmPc** out) {
  *out = nullptr;
  PYN_REQUIRE(A->mesh && !A->free_pattern, "pc asm needs an operator assembled from a mesh");
  PYN_REQUIRE(A->br == A->bc && A->nbrows == A->mesh->nnodes, "pc asm: operator is not square over the mesh nodes");
  if (A->halo && ctx->nranks > 1) {
    set_error("pc asm is not available on row-partitioned operators yet (single GPU only); use bjacobi");
    return PYN_E_UNSUPPORTED;
  }
  AsmPc* pc = new AsmPc();
  int rc;
  if (A->br == 3) rc = asm_setup_impl<3>(ctx, A, pc);
  else if (A->br == 2) rc = asm_setup_impl<2>(ctx, A, pc);
  else {
    set_error("pc asm: unsupported block size %d", A->br);
    rc = PYN_E_UNSUPPORTED;
  }
  if (rc != PYN_OK) {
    delete pc;
    return rc;
  }
  *out = pc;
  return PYN_OK;
}

template <int B>
static int asm_apply_impl(pyn_ctx* ctx, AsmPc* pc, const double* r, double* z, double* dots, const double* done) {
  pyn_mesh* mesh = pc->mesh;
  const int64_t cap = (int64_t)ctx->num_sms * 2;
  const int gridm = (int)(pc->ntiles < cap ? pc->ntiles : cap);
  asm_mult_kernel<B><<<gridm > 0 ? gridm : 1, 256, pc->smem, ctx->stream>>>(
      pc->ntiles, pc->tile_elems.p, pc->tile_class.p, mesh->elem_nodes.p, mesh->nn, pc->inv.p, pc->wnode.p, r,
      pc->ybuf.p, done);
  PYN_LAUNCHED(ctx);
  asm_sum_kernel<B><<<grid_for(ctx, mesh->nnodes, 256), 256, 0, ctx->stream>>>(
      mesh->nnodes, mesh->n2e_ptr.p, mesh->n2e.p, pc->ybuf.p, pc->ld, r, z, ctx->red_partials, ctx->red_ticket,
      dots, done);
  PYN_LAUNCHED(ctx);
  return PYN_OK;
}

int asm_apply(pyn_ctx* ctx, AsmPc* pc, const double* r, double* z, double* dots, const double* done) {
  if (pc->B == 3) return asm_apply_impl<3>(ctx, pc, r, z, dots, done);
  return asm_apply_impl<2>(ctx, pc, r, z, dots, done);
}

}  // namespace pyn
