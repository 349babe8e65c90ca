// asm_kernels.cuh -- the kernels of the element-block Schwarz preconditioner (asm_pc.cu).  In a
// header so that tests/hostcheck can run them on CPU threads through its CUDA emulation shim
// (tests/hostcheck/cuda_emu.h) in the build container, where there is no GPU.
#pragma once
#include "asm_host.cuh"
#include "vec.cuh"

#ifdef PYN_CUDA_EMU
#define PYN_DYN_SMEM(T, name) T* name = static_cast<T*>(cuda_emu::ts.dyn_smem)
#else
#define PYN_DYN_SMEM(T, name) extern __shared__ __align__(16) T name[]
#endif

namespace pyn {

constexpr int kAsmTile = 16;       // elements per CTA tile of the dense product
constexpr int kAsmKC = 16;         // k-chunk of the inverse staged in shared memory
constexpr int kAsmRG = 64;         // row groups (threads along the rows of a panel)
constexpr int kAsmRM = 3;          // rows per thread
constexpr int kAsmRP = kAsmRG * kAsmRM;   // rows per panel
constexpr int kAsmEM = 4;          // elements per thread (kAsmTile / 4 thread columns)
static_assert(kAsmTile == 4 * kAsmEM, "thread layout: 4 element groups x 64 row groups");

// MODE 0: hash the block of every element of the list into hash[e][2] (quantum > 0: of the values
//         rounded to multiples of it)
// MODE 1: write the dense block of list element li to dense[li][ld][ld]
// One warp per (list element, local row node).
template <int B, int MODE>
__global__ void __launch_bounds__(256)
asm_blocks_kernel(int64_t nlist, const int32_t* __restrict__ list, const int32_t* __restrict__ elem_nodes,
                  int nn, const int64_t* __restrict__ rowptr, const int32_t* __restrict__ colidx,
                  const double* __restrict__ val, unsigned long long* __restrict__ hash,
                  double* __restrict__ dense, double quantum) {
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
            // quantum > 0: values are compared as multiples of it (blocks equal up to rounding share a
            // class; what rounds to zero is a structural zero)
            unsigned long long bits = (unsigned long long)__double_as_longlong(v);
            if (quantum > 0.0) {
              bits = (unsigned long long)llrint(v / quantum);
              if (bits == 0ull) continue;
            }
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

// *out = bit pattern of max |val[i]| (non-negative doubles order like their bit patterns)
__global__ void asm_absmax_kernel(int64_t n, const double* __restrict__ val, unsigned long long* out) {
  double m = 0.0;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const double a = fabs(val[i]);
    if (a > m) m = a;                       // NaNs never win
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const double t = __shfl_xor_sync(0xffffffffu, m, o);
    if (t > m) m = t;
  }
  if ((threadIdx.x & 31) == 0) atomicMax(out, (unsigned long long)__double_as_longlong(m));
}

// valence[n] = elements of the local mesh holding node n (a row-partitioned run then refreshes
// the ghost entries from the owners, whose counts are complete)
__global__ void asm_valence_kernel(int64_t nnodes, const int64_t* __restrict__ n2e_ptr,
                                   double* __restrict__ valence) {
  for (int64_t n = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; n < nnodes;
       n += (int64_t)gridDim.x * blockDim.x)
    valence[n] = (double)(n2e_ptr[n + 1] - n2e_ptr[n]);
}

__global__ void asm_weights_kernel(int64_t nnodes, const double* __restrict__ valence,
                                   double* __restrict__ wnode) {
  for (int64_t n = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; n < nnodes;
       n += (int64_t)gridDim.x * blockDim.x) {
    const double m = valence[n];
    wnode[n] = m > 0.0 ? 1.0 / sqrt(m) : 0.0;
  }
}

// Signature of an element on a mesh whose cells are all congruent: its assembled block is then
// fixed by the dof classes and the global valences of its nodes (which neighbours exist, which
// rows and columns the boundary conditions replaced).  Row-partitioned runs use it to give an
// element whose block is incomplete here (ghost rows live on their owners) the inverse of a
// complete local element with the same signature.  One thread per element.
__global__ void asm_signature_kernel(int64_t nel, const int32_t* __restrict__ elem_nodes, int nn,
                                     const uint8_t* __restrict__ nodecls, const double* __restrict__ valence,
                                     unsigned long long* __restrict__ sig) {
  for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < nel; e += (int64_t)gridDim.x * blockDim.x) {
    unsigned long long h0 = 0ull, h1 = 0ull;
    for (int a = 0; a < nn; ++a) {
      const int64_t n = elem_nodes[e * nn + a];
      const unsigned long long key =
          (unsigned long long)(nodecls ? nodecls[n] : 0) | ((unsigned long long)(long long)valence[n] << 8);
      h0 += asm_hash_entry(key, (uint64_t)a, 0);
      h1 += asm_hash_entry(key, (uint64_t)a, 1);
    }
    sig[2 * e] = h0;
    sig[2 * e + 1] = h1;
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
  PYN_DYN_SMEM(double, asm_sm);
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

// second level: prolongation of the coarse correction xc (Q1 on the element corners), added by
// asm_sum_kernel.  xc == NULL: one-level preconditioner.
struct AsmProlong {
  const double* xc;             // [nv][B]
  const int32_t* elem_verts;    // [nel][ncorner] coarse vertex of corner q of element e
  const double* phi;            // [nn][ncorner] corner (multilinear) basis at the local nodes
  const double* keep;           // [nnodes*B] 1 free dof, 0 constrained (identity row of K)
  int ncorner;
};

// z[node] = sum over the elements of the node, ascending, of their local results (+ P xc)
template <int B>
__global__ void __launch_bounds__(256)
asm_sum_kernel(int64_t nnodes, const int64_t* __restrict__ n2e_ptr, const int2* __restrict__ n2e,
               const double* __restrict__ ybuf, int ld, const double* __restrict__ r, double* __restrict__ z,
               double* partials, unsigned int* ticket, double* dots, const double* __restrict__ done,
               AsmProlong cp) {
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
    if (cp.xc && n2e_ptr[n + 1] > n2e_ptr[n]) {    // + P xc through the node's first element
      const int2 ea = n2e[n2e_ptr[n]];
      const int32_t* ev = cp.elem_verts + (int64_t)ea.x * cp.ncorner;
      for (int q = 0; q < cp.ncorner; ++q) {
        const double w = cp.phi[ea.y * cp.ncorner + q];
        if (w != 0.0) {
          const double* x = cp.xc + (int64_t)ev[q] * B;
#pragma unroll
          for (int c = 0; c < B; ++c) zz[c] += cp.keep[n * B + c] * w * x[c];
        }
      }
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

// keep[n*B+i] = 1 when row i of node n couples to anything but its own diagonal entry, else 0
// (the identity rows K keeps for constrained dofs).  One warp per node.
template <int B>
__global__ void __launch_bounds__(256)
asm_free_dofs_kernel(int64_t nnodes, const int64_t* __restrict__ rowptr, const int32_t* __restrict__ colidx,
                     const double* __restrict__ val, double* __restrict__ keep) {
  constexpr int BB = B * B;
  const int lane = threadIdx.x & 31;
  const int64_t warp = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
  for (int64_t n = warp; n < nnodes; n += nwarps) {
    const int64_t p0 = rowptr[n];
    const int nb = (int)(rowptr[n + 1] - p0);
    int flag[B];
#pragma unroll
    for (int i = 0; i < B; ++i) flag[i] = 0;
    for (int u = lane; u < nb; u += 32) {
      const int64_t c = colidx[p0 + u];
#pragma unroll
      for (int q = 0; q < BB; ++q) {
        const int i = q / B, j = q - i * B;
        if (val[p0 * BB + (int64_t)q * nb + u] != 0.0 && !(c == n && i == j)) flag[i] = 1;
      }
    }
#pragma unroll
    for (int i = 0; i < B; ++i) {
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) flag[i] |= __shfl_xor_sync(0xffffffffu, flag[i], o);
      if (lane == 0) keep[n * B + i] = flag[i] ? 1.0 : 0.0;
    }
  }
}

// rc = P^T (keep r): one warp per coarse vertex over its (element, corner) list; a fine node
// contributes through its first element only (the one asm_sum_kernel prolongs through), lanes
// over the local nodes, fixed reduction tree
template <int B>
__global__ void __launch_bounds__(256)
asm_restrict_kernel(int64_t nv, const int64_t* __restrict__ v2e_ptr, const int2* __restrict__ v2e,
                    const int32_t* __restrict__ elem_nodes, int nn, int ncorner, const double* __restrict__ phi,
                    const int64_t* __restrict__ n2e_ptr, const int2* __restrict__ n2e,
                    const double* __restrict__ keep, const double* __restrict__ r, double* __restrict__ rc,
                    const double* __restrict__ done) {
  if (done && *done != 0.0) return;
  const int lane = threadIdx.x & 31;
  const int64_t warp = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
  for (int64_t v = warp; v < nv; v += nwarps) {
    double acc[B];
#pragma unroll
    for (int i = 0; i < B; ++i) acc[i] = 0.0;
    for (int64_t k = v2e_ptr[v]; k < v2e_ptr[v + 1]; ++k) {
      const int2 ec = v2e[k];
      for (int a = lane; a < nn; a += 32) {
        const double w = phi[a * ncorner + ec.y];
        if (w == 0.0) continue;
        const int64_t n = elem_nodes[(int64_t)ec.x * nn + a];
        const int2 first = n2e[n2e_ptr[n]];
        if (first.x != ec.x || first.y != a) continue;
#pragma unroll
        for (int i = 0; i < B; ++i) acc[i] += w * keep[n * B + i] * r[n * B + i];
      }
    }
#pragma unroll
    for (int i = 0; i < B; ++i) {
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) acc[i] += __shfl_xor_sync(0xffffffffu, acc[i], o);
      if (lane == 0) rc[v * B + i] = acc[i];
    }
  }
}

// Chebyshev iteration on the coarse problem Ac xc = rc with point Jacobi (Ac: block CSR over the
// vertex adjacency, B x B blocks row-major).  first: xc = 0, d_new = c2 D^-1 rc.  Otherwise one
// step: xc += d_old, rc -= Ac d_old, d_new = c1 d_old + c2 D^-1 rc.  One thread per vertex.
template <int B>
__global__ void __launch_bounds__(256)
asm_cheb_kernel(int64_t nv, const int64_t* __restrict__ ptr, const int32_t* __restrict__ col,
                const double* __restrict__ val, const double* __restrict__ dinv, int first, double c1, double c2,
                double* __restrict__ rc, double* __restrict__ xc, const double* __restrict__ d_old,
                double* __restrict__ d_new, const double* __restrict__ done) {
  if (done && *done != 0.0) return;
  for (int64_t v = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; v < nv; v += (int64_t)gridDim.x * blockDim.x) {
    if (first) {
#pragma unroll
      for (int i = 0; i < B; ++i) {
        xc[v * B + i] = 0.0;
        d_new[v * B + i] = c2 * dinv[v * B + i] * rc[v * B + i];
      }
      continue;
    }
    double ad[B];
#pragma unroll
    for (int i = 0; i < B; ++i) ad[i] = 0.0;
    for (int64_t k = ptr[v]; k < ptr[v + 1]; ++k) {
      const double* a = val + k * B * B;
      const double* d = d_old + (int64_t)col[k] * B;
#pragma unroll
      for (int i = 0; i < B; ++i)
#pragma unroll
        for (int j = 0; j < B; ++j) ad[i] += a[i * B + j] * d[j];
    }
#pragma unroll
    for (int i = 0; i < B; ++i) {
      const double dv = d_old[v * B + i];
      xc[v * B + i] += dv;
      const double rn = rc[v * B + i] - ad[i];
      rc[v * B + i] = rn;
      d_new[v * B + i] = c1 * dv + c2 * dinv[v * B + i] * rn;
    }
  }
}

// xf = keep (P xc): prolongation alone (set-up: probing the Galerkin coarse operator)
template <int B>
__global__ void __launch_bounds__(256)
asm_prolong_kernel(int64_t nnodes, const int64_t* __restrict__ n2e_ptr, const int2* __restrict__ n2e,
                   AsmProlong cp, double* __restrict__ xf) {
  for (int64_t n = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; n < nnodes;
       n += (int64_t)gridDim.x * blockDim.x) {
    double zz[B];
#pragma unroll
    for (int c = 0; c < B; ++c) zz[c] = 0.0;
    if (n2e_ptr[n + 1] > n2e_ptr[n]) {
      const int2 ea = n2e[n2e_ptr[n]];
      const int32_t* ev = cp.elem_verts + (int64_t)ea.x * cp.ncorner;
      for (int q = 0; q < cp.ncorner; ++q) {
        const double w = cp.phi[ea.y * cp.ncorner + q];
        if (w != 0.0) {
          const double* x = cp.xc + (int64_t)ev[q] * B;
#pragma unroll
          for (int c = 0; c < B; ++c) zz[c] += cp.keep[n * B + c] * w * x[c];
        }
      }
    }
#pragma unroll
    for (int c = 0; c < B; ++c) xf[n * B + c] = zz[c];
  }
}

}  // namespace pyn
