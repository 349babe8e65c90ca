// ksp.cu -- Krylov solve of the KLE system (k7): replaces the PETSc KSP/PC the
// reference configures in KspSolver.createSolver (kle_solver.py:49-64) and calls
// in KleSolver.solve (kle_solver.py:33-41).
//
// Preconditioned CG (K is SPD after Dirichlet elimination) with Jacobi or
// node-block Jacobi.  All Krylov scalars live on the device; an iteration is
//   (1) w = A p fused with p.w            (SpMV, one HBM pass over A)
//   (2) x += a p; r -= a w; z = M^-1 r; r.z and z.z reduced in the same pass
//   (3) p = z + b p
// and the convergence test (PETSc's default: preconditioned residual norm
// against max(rtol*||M^-1 b||, atol)) is evaluated on the device by the last
// CTA of (2).  The host only polls a pinned status word every few iterations;
// iterations launched after convergence return immediately.
#include "ksp.cuh"
#include <stdlib.h>
#include "vec.cuh"
#include "comm.cuh"
#include "asm_pc.cuh"
#include <memory>
#include <algorithm>

namespace pyn {
int g_ksp_check_every = 8;
bool g_force_idx64 = false;
bool g_asm_coarse_unfused = false;
bool g_halo_wait_at_start = true;    // r2f/r2g: the deferred per-warp wait measured 16 % slower at 2 GPUs
bool g_ksp_adaptive_chunks = true;
bool g_ksp_graphs = true;

// device scalar block
enum {
  S_RZ0 = 0,   // r.z, ping
  S_RZ1 = 1,   // r.z, pong
  S_PW = 2,    // p.Ap
  S_ZZ = 3,    // z.z (preconditioned residual norm^2)
  S_DONE = 4,  // != 0 once converged / broken down
  S_ITS = 5,   // iterations performed
  S_TTOL = 6,  // absolute threshold on sqrt(S_ZZ)
  S_REASON = 7,
  S_RNORM = 8,
  S_DTOL = 9,
  S_LRZ = 10,  // this rank's r.z / z.z partials (row-partitioned runs), summed in place by
  S_LZZ = 11,  // an allreduce before the bookkeeping kernel
  S_SR = 12,   // single-reduction CG: {w.u, r.u, u.u} of the current iterate (3 slots)
  S_COUNT = 16
};

template <int B>
__device__ __forceinline__ void apply_pc_node(int pc, const double* __restrict__ pcd, int64_t node,
                                              const double* r, double* z) {
  if (pc == PYN_PC_NONE) {
#pragma unroll
    for (int i = 0; i < B; ++i) z[i] = r[i];
  } else if (pc == PYN_PC_JACOBI) {
#pragma unroll
    for (int i = 0; i < B; ++i) z[i] = pcd[node * B + i] * r[i];
  } else {
#pragma unroll
    for (int i = 0; i < B; ++i) {
      double s = 0.0;
#pragma unroll
      for (int j = 0; j < B; ++j) s += pcd[(node * B + i) * B + j] * r[j];
      z[i] = s;
    }
  }
}

// z = M^-1 r ; out[0] = r.z ; out[1] = z.z          (set-up and residual restarts)
template <int B>
__global__ void __launch_bounds__(256)
pc_apply_dots_kernel(int pc, const double* __restrict__ pcd, const double* __restrict__ r,
                     double* __restrict__ z, int64_t nnodes, double* partials, unsigned int* ticket,
                     double* out) {
  __shared__ double sm[64];
  double v[2] = {0.0, 0.0};
  for (int64_t n = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; n < nnodes;
       n += (int64_t)gridDim.x * blockDim.x) {
    double rr[B], zz[B];
#pragma unroll
    for (int i = 0; i < B; ++i) rr[i] = r[n * B + i];
    apply_pc_node<B>(pc, pcd, n, rr, zz);
#pragma unroll
    for (int i = 0; i < B; ++i) {
      z[n * B + i] = zz[i];
      v[0] += rr[i] * zz[i];
      v[1] += zz[i] * zz[i];
    }
  }
  block_reduce<2>(v, sm);
  grid_reduce_finish<2>(v, partials, ticket, out);
}

// scalar bookkeeping of one iteration: new r.z, residual norm, convergence / breakdown test
__device__ __forceinline__ void cg_bookkeeping(double* S, double rz_new, double zz, int parity, int maxit) {
  const double pw = S[S_PW];
  S[parity ? S_RZ0 : S_RZ1] = rz_new;
  S[S_ZZ] = zz;
  const double its = S[S_ITS] + 1.0;
  S[S_ITS] = its;
  const double rn = sqrt(zz);
  S[S_RNORM] = rn;
  if (!(rn == rn) || !(pw == pw) || pw == 0.0) {       // NaN / breakdown
    S[S_REASON] = -9.0;
    S[S_DONE] = 1.0;
  } else if (pw < 0.0) {                                 // p.Ap < 0: KSP_DIVERGED_INDEFINITE_MAT
    S[S_REASON] = -10.0;
    S[S_DONE] = 1.0;
  } else if (rn <= S[S_TTOL]) {
    S[S_REASON] = S[S_REASON] == 0.0 ? 2.0 : S[S_REASON];
    S[S_DONE] = 1.0;
  } else if (rn >= S[S_DTOL]) {
    S[S_REASON] = -4.0;                                  // DIVERGED_DTOL
    S[S_DONE] = 1.0;
  } else if (its >= (double)maxit) {
    S[S_REASON] = -3.0;                                  // DIVERGED_ITS
    S[S_DONE] = 1.0;
  }
}

// row-partitioned runs: the partials S_LRZ/S_LZZ were summed over ranks by an allreduce
__global__ void cg_scalars_kernel(double* S, int parity, int maxit) {
  if (S[S_DONE] != 0.0) return;
  cg_bookkeeping(S, S[S_LRZ], S[S_LZZ], parity, maxit);
}

// the same with the allreduce of (r.z, z.z) over peer memory: this rank's partials are pushed into every
// rank's window and summed in rank order (one warp; used after the Schwarz preconditioner, whose apply
// sits between the vector update and this reduction)
__global__ void cg_scalars_p2p_kernel(double* S, int parity, int maxit, P2PRed red) {
  if (S[S_DONE] != 0.0) return;
  double t[2] = {S[S_LRZ], S[S_LZZ]}, g[2];
  p2p_push<2>(red, 1, t);
  const bool ok = p2p_wait_sum<2>(red, 1, g);
  if (threadIdx.x == 0) {
    if (ok) {
      cg_bookkeeping(S, g[0], g[1], parity, maxit);
    } else {
      S[S_REASON] = -100.0;              // a peer never answered
      S[S_DONE] = 1.0;
    }
  }
}

// step (2) of the iteration + convergence test in the last CTA
template <int B>
__global__ void __launch_bounds__(256)
cg_update_kernel(int pc, const double* __restrict__ pcd, const double* __restrict__ p,
                 const double* __restrict__ w, double* __restrict__ x, double* __restrict__ r,
                 double* __restrict__ z, int64_t nnodes, double* S, int parity, int maxit,
                 double* partials, unsigned int* ticket, int dist, P2PRed red) {
  __shared__ double sm[64];
  __shared__ bool is_last;
  __shared__ double s_pw;
  __shared__ int s_ok;
  if (S[S_DONE] != 0.0) return;
  const double rz = S[parity ? S_RZ1 : S_RZ0];
  double pw;
  if (red.nranks > 1) {                    // p.Ap partials of all ranks arrive in this rank's window
    if (threadIdx.x < 32) {
      double o[1];
      const bool ok = p2p_wait_sum<1>(red, 0, o);
      if (threadIdx.x == 0) {
        s_pw = o[0];
        s_ok = ok ? 1 : 0;
      }
    }
    __syncthreads();
    pw = s_pw;
    if (!s_ok) {
      if (threadIdx.x == 0 && blockIdx.x == 0) {
        S[S_REASON] = -100.0;              // a peer never answered
        S[S_DONE] = 1.0;
      }
      return;
    }
  } else {
    pw = S[S_PW];
  }
  const double alpha = rz / pw;
  double v[2] = {0.0, 0.0};
  for (int64_t n = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; n < nnodes;
       n += (int64_t)gridDim.x * blockDim.x) {
    double rr[B], zz[B];
#pragma unroll
    for (int i = 0; i < B; ++i) {
      x[n * B + i] += alpha * p[n * B + i];
      rr[i] = r[n * B + i] - alpha * w[n * B + i];
      r[n * B + i] = rr[i];
    }
    apply_pc_node<B>(pc, pcd, n, rr, zz);
#pragma unroll
    for (int i = 0; i < B; ++i) {
      z[n * B + i] = zz[i];
      v[0] += rr[i] * zz[i];
      v[1] += zz[i] * zz[i];
    }
  }
  block_reduce<2>(v, sm);
  // grid reduction, then the scalar bookkeeping by the last CTA
  if (threadIdx.x == 0) {
    partials[(size_t)blockIdx.x * 2 + 0] = v[0];
    partials[(size_t)blockIdx.x * 2 + 1] = v[1];
    __threadfence();
    is_last = (atomicAdd(ticket, 1u) == gridDim.x - 1);
  }
  __syncthreads();
  if (!is_last) return;
  __threadfence();
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  if (wid < 2) {
    double t = 0.0;
    for (unsigned int b = lane; b < gridDim.x; b += 32) t += partials[(size_t)b * 2 + wid];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) t += __shfl_xor_sync(0xffffffffu, t, o);
    if (lane == 0) sm[wid] = t;
  }
  __syncthreads();
  if (dist == 2) {                 // Schwarz preconditioner: z, r.z and the bookkeeping follow in other kernels
    if (threadIdx.x == 0) {
      *ticket = 0u;
      S[S_PW] = pw;
    }
    return;
  }
  if (red.nranks > 1) {            // fused allreduce over peer memory, then the bookkeeping
    if (threadIdx.x < 32) {
      double t[2] = {sm[0], sm[1]}, g[2];
      p2p_push<2>(red, 1, t);
      const bool ok = p2p_wait_sum<2>(red, 1, g);
      if (threadIdx.x == 0) {
        *ticket = 0u;
        S[S_PW] = pw;
        if (ok) {
          cg_bookkeeping(S, g[0], g[1], parity, maxit);
        } else {
          S[S_REASON] = -100.0;
          S[S_DONE] = 1.0;
        }
      }
    }
    return;
  }
  if (threadIdx.x == 0) {
    *ticket = 0u;
    if (dist) {                    // partial sums of this rank; allreduce + cg_scalars_kernel follow
      S[S_LRZ] = sm[0];
      S[S_LZZ] = sm[1];
    } else {
      cg_bookkeeping(S, sm[0], sm[1], parity, maxit);
    }
  }
}

// ---- single-synchronisation CG (PETSc -ksp_cg_single_reduction; Chronopoulos & Gear) ---------
// u = M^-1 r, w = A u.  delta = w.u is fused into the product kernel as before; gamma = r.u and
// u.u of the NEXT iterate are reduced by the vector kernel that produces r and u, so they travel
// while the next product runs and an iteration has ONE point where it waits for reductions:
//   beta = gamma/gamma_old, alpha = gamma/(delta - beta*gamma/alpha_old),
//   p = u + beta p, s = w + beta s, x += alpha p, r -= alpha s, u = M^-1 r.
// Two launches per iteration instead of three.  Scalars of iteration i live in slots [i & 1]
// (S_RZ0/1: gamma, S_LRZ/S_LZZ: alpha) so that the CTA doing the bookkeeping never overwrites
// what the others still read; S_SR+1/+2 are rewritten by the last CTA only after every CTA has
// read them (they deposit their partials after reading).
template <int B>
__global__ void __launch_bounds__(256)
cg_sr_update_kernel(int pc, const double* __restrict__ pcd, double* __restrict__ u,
                    const double* __restrict__ w, double* __restrict__ p, double* __restrict__ s,
                    double* __restrict__ x, double* __restrict__ r, int64_t nnodes, double* S, int it,
                    int maxit, double* partials, unsigned int* ticket, P2PRed red, HaloPush hp) {
  __shared__ double sm[64];
  __shared__ double s_sum[3];
  __shared__ int s_ok;
  __shared__ bool is_last;
  if (S[S_DONE] != 0.0) return;
  double delta, gamma, uu;
  if (red.nranks > 1) {                    // the ranks' partials arrive in this rank's window
    if (threadIdx.x < 32) {
      double o[1], g[2] = {0.0, 0.0};
      bool ok = p2p_wait_sum<1>(red, 0, o);
      if (it > 0) {                        // pushed by the previous iteration's vector kernel
        P2PRed prev = red;
        prev.seq = red.seq - 1;
        ok = p2p_wait_sum<2>(prev, 1, g) && ok;
      }
      if (threadIdx.x == 0) {
        s_sum[0] = o[0];
        s_sum[1] = g[0];
        s_sum[2] = g[1];
        s_ok = ok ? 1 : 0;
      }
    }
    __syncthreads();
    if (!s_ok) {
      if (threadIdx.x == 0 && blockIdx.x == 0) {
        S[S_REASON] = -100.0;              // a peer never answered
        S[S_DONE] = 1.0;
      }
      return;
    }
    delta = s_sum[0];
    gamma = it > 0 ? s_sum[1] : S[S_SR + 1];
    uu = it > 0 ? s_sum[2] : S[S_SR + 2];
  } else {
    delta = S[S_SR];
    gamma = S[S_SR + 1];
    uu = S[S_SR + 2];
  }
  const int cur = it & 1;
  const double gamma_old = S[cur ? S_RZ0 : S_RZ1];
  const double alpha_old = S[cur ? S_LRZ : S_LZZ];
  const double beta = it == 0 ? 0.0 : gamma / gamma_old;
  const double denom = it == 0 ? delta : delta - beta * gamma / alpha_old;
  const double alpha = gamma / denom;
  const double rn = sqrt(uu);
  // every CTA evaluates the same test on the same numbers; CTA 0 records the outcome
  int stop = 0;
  double reason = 0.0;
  if (!(rn == rn) || !(delta == delta)) {
    stop = 1; reason = -9.0;
  } else if (rn <= S[S_TTOL]) {
    stop = 1; reason = S[S_REASON] == 3.0 ? 3.0 : 2.0;
  } else if (rn >= S[S_DTOL]) {
    stop = 1; reason = -4.0;
  } else if (it >= maxit) {
    stop = 1; reason = -3.0;
  } else if (!(denom == denom) || denom == 0.0) {
    stop = 1; reason = -9.0;
  } else if (denom < 0.0) {                  // p.Kp < 0: KSP_DIVERGED_INDEFINITE_MAT
    stop = 1; reason = -10.0;
  }
  if (blockIdx.x == 0 && threadIdx.x == 0) {
    S[S_ZZ] = uu;
    S[S_RNORM] = rn;
    S[S_PW] = delta;
    if (stop) {
      S[S_REASON] = reason;
      S[S_DONE] = 1.0;
    } else {
      S[cur ? S_RZ1 : S_RZ0] = gamma;
      S[cur ? S_LZZ : S_LRZ] = alpha;
      S[S_ITS] = (double)(it + 1);
    }
  }
  if (stop) return;
  double v[2] = {0.0, 0.0};
  for (int64_t n = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; n < nnodes;
       n += (int64_t)gridDim.x * blockDim.x) {
    double rr[B], zz[B];
#pragma unroll
    for (int i = 0; i < B; ++i) {
      const int64_t j = n * B + i;
      const double pj = it == 0 ? u[j] : u[j] + beta * p[j];
      const double sj = it == 0 ? w[j] : w[j] + beta * s[j];
      p[j] = pj;
      s[j] = sj;
      x[j] += alpha * pj;
      rr[i] = r[j] - alpha * sj;
      r[j] = rr[i];
    }
    apply_pc_node<B>(pc, pcd, n, rr, zz);
#pragma unroll
    for (int i = 0; i < B; ++i) {
      u[n * B + i] = zz[i];
      v[0] += rr[i] * zz[i];
      v[1] += zz[i] * zz[i];
    }
    // fused halo push: u is the input of the next product; neighbours read it from their windows
    if (hp.nneigh > 0) halo_push_node(hp, n, zz);
  }
  if (hp.nneigh > 0) __threadfence_system();       // this thread's remote stores, before the ticket
  // gamma and u.u of the next iterate: grid reduction, then to S (or to every rank's window)
  block_reduce<2>(v, sm);
  if (threadIdx.x == 0) {
    partials[(size_t)blockIdx.x * 2 + 0] = v[0];
    partials[(size_t)blockIdx.x * 2 + 1] = v[1];
    __threadfence();
    is_last = (atomicAdd(ticket, 1u) == gridDim.x - 1);
  }
  __syncthreads();
  if (!is_last) return;
  __threadfence();
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  if (wid < 2) {
    double t = 0.0;
    for (unsigned int b = lane; b < gridDim.x; b += 32) t += partials[(size_t)b * 2 + wid];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) t += __shfl_xor_sync(0xffffffffu, t, o);
    if (lane == 0) sm[wid] = t;
  }
  __syncthreads();
  if (hp.nneigh > 0) halo_release_flags(hp);       // every CTA's pushes are visible: raise the flags
  if (threadIdx.x < 32) {
    if (red.nranks > 1) {
      double t[2] = {sm[0], sm[1]};
      p2p_push<2>(red, 1, t);
    } else if (threadIdx.x == 0) {
      S[S_SR + 1] = sm[0];
      S[S_SR + 2] = sm[1];
    }
    if (threadIdx.x == 0) *ticket = 0u;
  }
}

// step (3): p = z + (rz_new/rz_old) p
__global__ void __launch_bounds__(256)
cg_direction_kernel(const double* __restrict__ z, double* __restrict__ p, int64_t n, const double* S,
                    int parity) {
  if (S[S_DONE] != 0.0) return;
  const double beta = S[parity ? S_RZ0 : S_RZ1] / S[parity ? S_RZ1 : S_RZ0];
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n;
       i += (int64_t)gridDim.x * blockDim.x)
    p[i] = z[i] + beta * p[i];
}

// start-up scalars: out2 = {r.z, z.z} of the initial residual, bnorm2 = {.., ||M^-1 b||^2}
__global__ void cg_init_scalars_kernel(double* S, const double* rz_zz, const double* b_zz, double rtol,
                                       double atol, double dtol, int zero_guess) {
  const double bn = sqrt(b_zz[1]);
  const double rn = sqrt(rz_zz[1]);
  double ttol = rtol * bn;
  double reason_if = 2.0;
  if (atol > ttol) {
    ttol = atol;
    reason_if = 3.0;
  }
  S[S_RZ0] = rz_zz[0];
  S[S_SR + 1] = rz_zz[0];      // single-reduction CG reads gamma_0 = r.u and u.u of the start here
  S[S_SR + 2] = rz_zz[1];
  S[S_RZ1] = 0.0;
  S[S_PW] = 0.0;
  S[S_ZZ] = rz_zz[1];
  S[S_ITS] = 0.0;
  S[S_TTOL] = ttol;
  S[S_DTOL] = dtol * (bn > 0.0 ? bn : 1.0);
  S[S_RNORM] = rn;
  S[S_REASON] = reason_if == 3.0 ? 3.0 : 0.0;
  S[S_DONE] = 0.0;
  if (rn <= ttol) {            // converged before the first iteration (also b = 0)
    S[S_REASON] = reason_if;
    S[S_DONE] = 1.0;
  }
  (void)zero_guess;
}

// Jacobi / node-block Jacobi set-up from the diagonal blocks
template <int B>
__global__ void pc_setup_kernel(int pc, const double* __restrict__ dblk, double* __restrict__ pcd,
                                int64_t nnodes) {
  for (int64_t n = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; n < nnodes;
       n += (int64_t)gridDim.x * blockDim.x) {
    const double* d = dblk + n * B * B;
    if (pc == PYN_PC_JACOBI) {
      for (int i = 0; i < B; ++i) {
        const double v = d[i * B + i];
        pcd[n * B + i] = v != 0.0 ? 1.0 / v : 1.0;   // PCJacobi replaces zero diagonals by 1
      }
    } else {
      double inv[B * B];
      bool ok = true;
      if (B == 1) {
        ok = d[0] != 0.0;
        inv[0] = ok ? 1.0 / d[0] : 1.0;
      } else if (B == 2) {
        const double det = d[0] * d[3] - d[1] * d[2];
        ok = det != 0.0;
        if (ok) {
          inv[0] = d[3] / det; inv[1] = -d[1] / det; inv[2] = -d[2] / det; inv[3] = d[0] / det;
        }
      } else {
        const double c0 = d[4] * d[8] - d[5] * d[7], c1 = d[5] * d[6] - d[3] * d[8],
                     c2 = d[3] * d[7] - d[4] * d[6];
        const double det = d[0] * c0 + d[1] * c1 + d[2] * c2;
        ok = det != 0.0;
        if (ok) {
          const double id = 1.0 / det;
          inv[0] = c0 * id;
          inv[1] = (d[2] * d[7] - d[1] * d[8]) * id;
          inv[2] = (d[1] * d[5] - d[2] * d[4]) * id;
          inv[3] = c1 * id;
          inv[4] = (d[0] * d[8] - d[2] * d[6]) * id;
          inv[5] = (d[2] * d[3] - d[0] * d[5]) * id;
          inv[6] = c2 * id;
          inv[7] = (d[1] * d[6] - d[0] * d[7]) * id;
          inv[8] = (d[0] * d[4] - d[1] * d[3]) * id;
        }
      }
      for (int q = 0; q < B * B; ++q)
        pcd[n * B * B + q] = ok ? inv[q] : ((q / B == q % B) ? 1.0 : 0.0);
    }
  }
}

// ---- Fischer initial guess (KSPSetUseFischerGuess model 1: A-conjugate projection) ---------
// PETSc option -ksp_fischer_guess 1,<size>.  The last <size> solution increments are kept
// A-orthonormal (x~_i, A x~_i); the guess for a new right-hand side b is sum_i (x~_i . b) x~_i,
// the A-norm-optimal element of their span.  Successive KLE systems differ only through
// time-dependent boundary values and the slowly changing vorticity, so most of the solution is
// recovered before the first Krylov iteration.
constexpr int kFischerMax = 8;        // basis vectors handled by one multi_dot / lincomb launch
constexpr int kFischerCap = 32;       // capacity of the Fischer basis (chunks of kFischerMax)
struct BasisPtrs {
  const double* v[kFischerMax];
  int k;
};

// out[i] = basis_i . y, i < k (one pass over y and the k basis vectors)
__global__ void __launch_bounds__(256)
multi_dot_kernel(BasisPtrs b, const double* __restrict__ y, int64_t n, double* partials,
                 unsigned int* ticket, double* out) {
  __shared__ double sm[kFischerMax * 32];
  double v[kFischerMax];
#pragma unroll
  for (int i = 0; i < kFischerMax; ++i) v[i] = 0.0;
  for (int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; j < n; j += (int64_t)gridDim.x * blockDim.x) {
    const double yj = y[j];
#pragma unroll
    for (int i = 0; i < kFischerMax; ++i)
      if (i < b.k) v[i] += b.v[i][j] * yj;
  }
  block_reduce<kFischerMax>(v, sm);
  grid_reduce_finish<kFischerMax>(v, partials, ticket, out);
}

// y = (keep ? y : 0) + sign * sum_i coef[i] * basis_i
__global__ void lincomb_kernel(double* __restrict__ y, BasisPtrs b, const double* __restrict__ coef, double sign,
                               int keep, int64_t n) {
  double c[kFischerMax];
#pragma unroll
  for (int i = 0; i < kFischerMax; ++i) c[i] = i < b.k ? sign * coef[i] : 0.0;
  for (int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; j < n; j += (int64_t)gridDim.x * blockDim.x) {
    double s = keep ? y[j] : 0.0;
#pragma unroll
    for (int i = 0; i < kFischerMax; ++i)
      if (i < b.k) s += c[i] * b.v[i][j];
    y[j] = s;
  }
}

// d = x - x0 ; ad = r0 - r      (A d = A x - A x0 = (b - r) - (b - r0))
__global__ void fischer_delta_kernel(const double* __restrict__ x, const double* __restrict__ x0,
                                     const double* __restrict__ r0, const double* __restrict__ r,
                                     double* __restrict__ d, double* __restrict__ ad, int64_t n) {
  for (int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; j < n; j += (int64_t)gridDim.x * blockDim.x) {
    d[j] = x[j] - x0[j];
    ad[j] = r0[j] - r[j];
  }
}

// nrm2[0] = d.Ad after the orthogonalisation, nrm2[1] = d.Ad before it, nrm2[2] = x.b ~ x.Ax.
// Keep the new direction only if its A-norm is well above the solver's noise floor (tol
// relative to the solution's A-norm): a direction made of round-off would be scaled up to unit
// norm and poison every later guess.  Scale to unit A-norm.
__global__ void fischer_append_kernel(const double* __restrict__ d, const double* __restrict__ ad,
                                      const double* __restrict__ nrm2, double tol2, double* __restrict__ bx,
                                      double* __restrict__ bax, int64_t n, double* accepted) {
  const double a = nrm2[0], a0 = nrm2[1];
  const bool ok = a > 0.0 && a0 > 0.0 && a > tol2 * fabs(nrm2[2]) && a == a;
  if (blockIdx.x == 0 && threadIdx.x == 0) *accepted = ok ? 1.0 : 0.0;
  if (!ok) return;
  const double s = 1.0 / sqrt(a);
  for (int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; j < n; j += (int64_t)gridDim.x * blockDim.x) {
    bx[j] = s * d[j];
    bax[j] = s * ad[j];
  }
}

}  // namespace pyn

using namespace pyn;

struct pyn_ksp {
  pyn_ctx* ctx = nullptr;
  pyn_mat* A = nullptr;
  int type = PYN_KSP_CG, pc = PYN_PC_BJACOBI;
  double rtol = 1e-5, atol = 1e-50, dtol = 1e5;
  int maxit = 10000, restart = 30;
  int check_every = pyn::g_ksp_check_every;
  int last_its = -1;              // iterations of the previous solve: sizes the first chunk of the next
  int hint_its = -1;              // pyn_ksp_hint_iterations: the caller's expectation for the NEXT solve
  // one GPU: the launches of a CG iteration differ from one iteration to the next only in the parity of
  // the iteration, so they are captured once per parity in a CUDA graph and replayed (one graph launch
  // instead of 3 - 7 kernel launches; the gaps between dependent launches were ~75 us of a 358 us
  // iteration under the two-level preconditioner, r2n trace).  Re-captured when what the kernels were
  // given changes (solution vector, stored / matrix-free product, iteration limit).
  cudaGraphExec_t gexec[2] = {nullptr, nullptr};
  int gnodes[2] = {0, 0};
  const void* g_x = nullptr;
  const void* g_mf = nullptr;
  int g_maxit = -1;
  bool g_failed = false;          // capture was refused once: plain launches from then on
  int B = 0;
  int64_t n = 0, nnodes = 0;
  DevBuf<double> pcd, r, z, p, w, S, tmp2;
  bool single_reduction = false;        // -ksp_cg_single_reduction
  DevBuf<double> sr_s;                  // s = A p of the single-reduction recurrence
  double* status_host = nullptr;  // pinned [2][S_COUNT]
  cudaEvent_t ev[2] = {nullptr, nullptr};
  bool pc_ready = false;
  // Fischer guess: basis of <= fischer_m A-orthonormal increments and their images
  int fischer_m = 0, fischer_n = 0;
  DevBuf<double> fx[kFischerCap], fax[kFischerCap];
  DevBuf<double> x0, r0, fd, fad, fcoef;     // fcoef: [kFischerCap + 4]
  int64_t fischer_used = 0, fischer_restarts = 0;
  // work vectors of BiCGStab / GMRES(m)
  DevBuf<double> kv[4];
  std::vector<std::unique_ptr<DevBuf<double>>> gv;
  DevBuf<double> gcoef;
  AsmPc* asmpc = nullptr;               // -pc_type asm (asm_pc.cuh)
  int asm_coarse = 0;                   // -pc_asm_coarse <k>: Chebyshev steps of the coarse level (0: off)
  void drop_graphs() {
    for (int i = 0; i < 2; ++i) {
      if (gexec[i]) cudaGraphExecDestroy(gexec[i]);
      gexec[i] = nullptr;
    }
  }
  ~pyn_ksp() {
    drop_graphs();
    delete asmpc;
  }
};

template <int B>
static int ksp_setup_pc(pyn_ksp* k) {
  pyn_ctx* ctx = k->ctx;
  k->drop_graphs();                 // captured iterations hold pointers into the old preconditioner
  if (k->pc == PYN_PC_NONE) {
    k->pc_ready = true;
    return PYN_OK;
  }
  if (k->pc == PYN_PC_ASM) {
    delete k->asmpc;
    k->asmpc = nullptr;
    PYN_CHECK(asm_setup(ctx, k->A, k->asm_coarse, &k->asmpc));
    k->pc_ready = true;
    return PYN_OK;
  }
  DevBuf<double> dblk;
  PYN_CHECK(dblk.alloc(ctx, k->nnodes * B * B));
  PYN_CHECK(diag_blocks(k->A, dblk.p));
  PYN_CHECK(k->pcd.alloc(ctx, k->pc == PYN_PC_JACOBI ? k->n : k->nnodes * B * B));
  pc_setup_kernel<B><<<grid_for(ctx, k->nnodes, 128), 128, 0, ctx->stream>>>(k->pc, dblk.p, k->pcd.p, k->nnodes);
  PYN_LAUNCHED(ctx);
  PYN_CUDA(cudaStreamSynchronize(ctx->stream));
  k->pc_ready = true;
  return PYN_OK;
}

// z = M^-1 r ; out[0] = r.z ; out[1] = z.z
template <int B>
static int ksp_pc_dots(pyn_ksp* k, const double* r, double* z, int64_t nn, int gridn, double* out) {
  pyn_ctx* ctx = k->ctx;
  // (asm refreshes r's ghost entries in place on a row-partitioned operator: copies of the owners' values)
  if (k->pc == PYN_PC_ASM) return asm_apply(ctx, k->asmpc, const_cast<double*>(r), z, out, nullptr);
  pc_apply_dots_kernel<B><<<gridn, 256, 0, ctx->stream>>>(k->pc, k->pcd.p, r, z, nn, ctx->red_partials,
                                                          ctx->red_ticket, out);
  PYN_LAUNCHED(ctx);
  return PYN_OK;
}

// PYN_KSP_TRACE=1: one line per solve on stderr with the host wall time of its phases (us): enqueue of
// the set-up (guess, residual, preconditioner), the iteration loop up to the moment the host sees the
// convergence flag, the epilogue (Fischer update + final read-back), and the gap since the previous solve
// returned (the caller's time: right-hand-side evaluation, Python)
static double trace_now_us() {
  timespec ts;
  clock_gettime(CLOCK_MONOTONIC, &ts);
  return ts.tv_sec * 1e6 + ts.tv_nsec * 1e-3;
}
static double g_trace_last_return = 0.0;

template <int B>
static int ksp_solve_cg(pyn_ksp* k, const double* b, double* x, int* its, int* reason, double* rnorm) {
  static const bool trace = getenv("PYN_KSP_TRACE") != nullptr;
  const double tr0 = trace ? trace_now_us() : 0.0;
  double tr1 = 0.0, tr2 = 0.0;
  int tr_chunks = 0;
  pyn_ctx* ctx = k->ctx;
  const int dist = k->A->halo != nullptr && ctx->nranks > 1;
  // fused peer-memory allreduce instead of NCCL.  Under the Schwarz preconditioner, whose apply sits
  // between the vector update and the reduction of r.z, p.Ap still rides in the product / update kernels
  // and (r.z, z.z) are reduced by cg_scalars_p2p_kernel after the apply
  const bool p2p = dist && p2p_enabled(ctx);
  P2PRed red_off;
  memset(&red_off, 0, sizeof(red_off));
  red_off.nranks = 1;
  // row-partitioned: Krylov vectors hold owned + ghost entries, the kernels touch the owned ones
  const int64_t nn = k->A->halo ? k->A->halo->n_own : k->nnodes;
  const int64_t n = nn * B;
  double* S = k->S.p;
  const int gridn = grid_for(ctx, nn > 0 ? nn : 1, 256);
  if (!k->pc_ready) PYN_CHECK(ksp_setup_pc<B>(k));
  // The Schwarz preconditioner is not node-local, so it cannot ride inside the vector kernel:
  // the update runs unpreconditioned and leaves the bookkeeping to cg_scalars_kernel, exactly as
  // on the row-partitioned path (there an allreduce, here the preconditioner sits in between).
  const bool asmpc = k->pc == PYN_PC_ASM;
  const bool single_reduction = k->single_reduction && !asmpc;
  if (single_reduction && k->sr_s.n < k->n) PYN_CHECK(k->sr_s.alloc(ctx, k->n));
  const bool fischer = k->fischer_m > 0;
  const int gridv = grid_for(ctx, n > 0 ? n : 1, 256);
  if (fischer && k->fischer_n > 0) {        // x <- sum_i (x~_i . b) x~_i
    for (int i0 = 0; i0 < k->fischer_n; i0 += kFischerMax) {
      BasisPtrs bp;
      bp.k = k->fischer_n - i0 < kFischerMax ? k->fischer_n - i0 : kFischerMax;
      for (int i = 0; i < bp.k; ++i) bp.v[i] = k->fx[i0 + i].p;
      multi_dot_kernel<<<gridv, 256, 0, ctx->stream>>>(bp, b, n, ctx->red_partials, ctx->red_ticket,
                                                       k->fcoef.p + i0);
      PYN_LAUNCHED(ctx);
      if (dist) PYN_CHECK(comm_allreduce_sum(ctx, k->fcoef.p + i0, kFischerMax));
      lincomb_kernel<<<gridv, 256, 0, ctx->stream>>>(x, bp, k->fcoef.p + i0, 1.0, i0 > 0 ? 1 : 0, n);
      PYN_LAUNCHED(ctx);
    }
    k->fischer_used++;
  } else if (fischer) {
    // empty basis: start from zero so that the increments span the solutions themselves
    PYN_CHECK(vec_set(ctx, x, n, 0.0));
  }
  // ||M^-1 b|| for the threshold (KSPConvergedDefault with the preconditioned norm)
  PYN_CHECK(ksp_pc_dots<B>(k, b, k->z.p, nn, gridn, k->tmp2.p + 2));
  // r = b - A x ; z = M^-1 r ; p = z
  PYN_CHECK(spmv(k->A, -1.0, x, 1.0, b, k->r.p));
  if (fischer) {                            // keep the guess and its residual for the basis update
    PYN_CUDA(cudaMemcpyAsync(k->x0.p, x, sizeof(double) * n, cudaMemcpyDeviceToDevice, ctx->stream));
    PYN_CUDA(cudaMemcpyAsync(k->r0.p, k->r.p, sizeof(double) * n, cudaMemcpyDeviceToDevice, ctx->stream));
  }
  PYN_CHECK(ksp_pc_dots<B>(k, k->r.p, k->z.p, nn, gridn, k->tmp2.p));
  if (dist) PYN_CHECK(comm_allreduce_sum(ctx, k->tmp2.p, 4));
  cg_init_scalars_kernel<<<1, 1, 0, ctx->stream>>>(S, k->tmp2.p, k->tmp2.p + 2, k->rtol, k->atol, k->dtol, 0);
  PYN_LAUNCHED(ctx);
  PYN_CUDA(cudaMemcpyAsync(k->p.p, k->z.p, sizeof(double) * n, cudaMemcpyDeviceToDevice, ctx->stream));

  // single-reduction CG over peer memory: halo exchange fused into the two kernels of an iteration
  const bool fuse_halo = single_reduction && p2p && k->A->halo->p2p && k->A->halo->nneigh > 0 &&
                         getenv("PYN_FUSED_HALO_OFF") == nullptr;
  bool halo_ready = false;
  HaloWait hw;
  memset(&hw, 0, sizeof(hw));
  // Chunks of `check_every` iterations; the status of chunk c-1 is read while
  // chunk c runs, so the GPU never idles on the host.
  // the launches of iteration `it`
  auto enqueue_iteration = [&](int it, int parity) -> int {
      P2PRed red = p2p ? p2p_next(ctx) : red_off;
      if (single_reduction) {
        // w = A u with w.u fused; then the whole recurrence (and r.u, u.u of the next iterate)
        // in one vector kernel.  Over peer memory the halo exchange of u rides in the same two
        // kernels: pushed by the vector kernel, awaited by the next product.
        PYN_CHECK(spmv_dot(k->A, k->z.p, k->w.p, S + S_SR, S + S_DONE, &red, halo_ready ? &hw : nullptr));
        if (dist && !p2p) PYN_CHECK(comm_allreduce_sum(ctx, S + S_SR, 1));
        HaloPush hp;
        memset(&hp, 0, sizeof(hp));
        if (fuse_halo) {
          halo_fused_next(k->A->halo, B, &hp, &hw);
          halo_ready = true;
        }
        prof_begin(ctx, PROF_CG_UPDATE);
        cg_sr_update_kernel<B><<<gridn, 256, 0, ctx->stream>>>(k->pc, k->pcd.p, k->z.p, k->w.p, k->p.p,
                                                               k->sr_s.p, x, k->r.p, nn, S, it, k->maxit,
                                                               ctx->red_partials, ctx->red_ticket, red, hp);
        prof_end(ctx);
        PYN_LAUNCHED(ctx);
        if (dist && !p2p) PYN_CHECK(comm_allreduce_sum(ctx, S + S_SR + 1, 2));
        return PYN_OK;
      }
      PYN_CHECK(spmv_dot(k->A, k->p.p, k->w.p, S + S_PW, S + S_DONE, &red));
      if (dist && !p2p) PYN_CHECK(comm_allreduce_sum(ctx, S + S_PW, 1));
      prof_begin(ctx, PROF_CG_UPDATE);
      cg_update_kernel<B><<<gridn, 256, 0, ctx->stream>>>(asmpc ? PYN_PC_NONE : k->pc, k->pcd.p, k->p.p,
                                                          k->w.p, x, k->r.p, k->z.p, nn, S, parity, k->maxit,
                                                          ctx->red_partials, ctx->red_ticket,
                                                          asmpc ? 2 : (dist ? 1 : 0), red);
      prof_end(ctx);
      PYN_LAUNCHED(ctx);
      // z = M^-1 r with r.z and z.z (this rank's part) into the partial slots
      if (asmpc) PYN_CHECK(asm_apply(ctx, k->asmpc, k->r.p, k->z.p, S + S_LRZ, S + S_DONE));
      if (dist && !p2p) PYN_CHECK(comm_allreduce_sum(ctx, S + S_LRZ, 2));
      if (asmpc && p2p) {
        cg_scalars_p2p_kernel<<<1, 32, 0, ctx->stream>>>(S, parity, k->maxit, red);
        PYN_LAUNCHED(ctx);
      } else if (asmpc || (dist && !p2p)) {
        cg_scalars_kernel<<<1, 1, 0, ctx->stream>>>(S, parity, k->maxit);
        PYN_LAUNCHED(ctx);
      }
      prof_begin(ctx, PROF_CG_DIR);
      cg_direction_kernel<<<grid_for(ctx, n > 0 ? n : 1, 256), 256, 0, ctx->stream>>>(k->z.p, k->p.p, n, S, parity);
      prof_end(ctx);
      PYN_LAUNCHED(ctx);
      return PYN_OK;
  };
  // one GPU, plain CG recurrence, no per-launch profiling events: replay captured graphs
  bool use_graph = g_ksp_graphs && !dist && !single_reduction && !ctx->prof_on && !k->g_failed;
  if (use_graph && (k->g_x != (const void*)x || k->g_mf != (const void*)k->A->mf || k->g_maxit != k->maxit)) {
    k->drop_graphs();
    k->g_x = x;
    k->g_mf = k->A->mf;
    k->g_maxit = k->maxit;
  }
  int it = 0;
  if (trace) tr1 = trace_now_us();
  for (int chunk = 0;; ++chunk) {
    tr_chunks = chunk + 1;
    // Successive KLE systems need about the same number of iterations, so the first chunk runs to
    // just short of the previous count and the following ones are short: iterations launched after
    // convergence return at once but still cost their launches (r2c: ~7 iterations per solve under
    // the two-level preconditioner, up to 9 dead iterations of 7 launches each in chunks of 8).
    // Every rank sees the same counts, so the launch sequences stay identical across ranks.
    const int tail = std::max(1, std::min(k->check_every, asmpc ? 2 : 4));
    // the caller may know better than "like the last solve" (the RK stages of a time step need
    // 9, 17, 0, 3, 5, 6, 9 iterations, the same pattern every step: r2l trace)
    const int expect = k->hint_its >= 0 ? k->hint_its : (k->last_its > 2 ? k->last_its - 1 : -1);
    const int first = expect >= 0 ? std::max(1, std::min(expect, 4 * k->check_every)) : k->check_every;
    const int want = !g_ksp_adaptive_chunks ? k->check_every : (chunk == 0 ? first : tail);
    const int todo = (k->maxit - it) < want ? (k->maxit - it) : want;
    for (int j = 0; j < todo; ++j, ++it) {
      const int parity = it & 1;
      if (use_graph) {
        if (!k->gexec[parity]) {
          const int64_t l0 = ctx->launches;
          cudaGraph_t g = nullptr;
          bool ok = cudaStreamBeginCapture(ctx->stream, cudaStreamCaptureModeThreadLocal) == cudaSuccess;
          if (ok) {
            const int rc = enqueue_iteration(it, parity);
            ok = cudaStreamEndCapture(ctx->stream, &g) == cudaSuccess && rc == PYN_OK && g != nullptr;
          }
          if (ok) ok = cudaGraphInstantiate(&k->gexec[parity], g, 0) == cudaSuccess;
          if (g) cudaGraphDestroy(g);
          k->gnodes[parity] = (int)(ctx->launches - l0);
          ctx->launches = l0;
          if (!ok) {                       // e.g. a driver that will not capture the cooperative launch
            cudaGetLastError();
            k->drop_graphs();
            k->g_failed = true;
            use_graph = false;
          }
        }
        if (use_graph) {
          PYN_CUDA(cudaGraphLaunch(k->gexec[parity], ctx->stream));
          ctx->launches += k->gnodes[parity];
          continue;
        }
      }
      PYN_CHECK(enqueue_iteration(it, parity));
    }
    const int slot = chunk & 1;
    PYN_CUDA(cudaMemcpyAsync(k->status_host + slot * S_COUNT, S, sizeof(double) * S_COUNT,
                             cudaMemcpyDeviceToHost, ctx->stream));
    PYN_CUDA(cudaEventRecord(k->ev[slot], ctx->stream));
    if (chunk >= 1) {
      PYN_CUDA(cudaEventSynchronize(k->ev[slot ^ 1]));
      if (k->status_host[(slot ^ 1) * S_COUNT + S_DONE] != 0.0) break;
    }
    if (it >= k->maxit) break;
  }
  if (trace) tr2 = trace_now_us();
  if (fischer) {
    // d = x - x0, A d = r0 - r; A-orthogonalise against the basis; append (restart when full)
    fischer_delta_kernel<<<gridv, 256, 0, ctx->stream>>>(x, k->x0.p, k->r0.p, k->r.p, k->fd.p, k->fad.p, n);
    PYN_LAUNCHED(ctx);
    if (k->fischer_n == k->fischer_m) {     // restart: keep only the current solution (A x = b - r)
      k->fischer_n = 0;
      k->fischer_restarts++;
      PYN_CUDA(cudaMemcpyAsync(k->fd.p, x, sizeof(double) * n, cudaMemcpyDeviceToDevice, ctx->stream));
      PYN_CHECK(vec_waxpy(ctx, k->fad.p, -1.0, k->r.p, b, n));
    }
    double* nrm2 = k->fcoef.p + kFischerCap;
    PYN_CHECK(vec_dot_dev(ctx, k->fd.p, k->fad.p, n, nrm2 + 1));
    // classical Gram-Schmidt in the A inner product: all coefficients from the same A d, so the
    // chunks are independent
    for (int i0 = 0; i0 < k->fischer_n; i0 += kFischerMax) {
      BasisPtrs bp;
      bp.k = k->fischer_n - i0 < kFischerMax ? k->fischer_n - i0 : kFischerMax;
      for (int i = 0; i < bp.k; ++i) bp.v[i] = k->fx[i0 + i].p;
      multi_dot_kernel<<<gridv, 256, 0, ctx->stream>>>(bp, k->fad.p, n, ctx->red_partials, ctx->red_ticket,
                                                       k->fcoef.p + i0);
      PYN_LAUNCHED(ctx);
      if (dist) PYN_CHECK(comm_allreduce_sum(ctx, k->fcoef.p + i0, kFischerMax));
    }
    for (int i0 = 0; i0 < k->fischer_n; i0 += kFischerMax) {
      BasisPtrs bp, bap;
      bp.k = bap.k = k->fischer_n - i0 < kFischerMax ? k->fischer_n - i0 : kFischerMax;
      for (int i = 0; i < bp.k; ++i) {
        bp.v[i] = k->fx[i0 + i].p;
        bap.v[i] = k->fax[i0 + i].p;
      }
      lincomb_kernel<<<gridv, 256, 0, ctx->stream>>>(k->fd.p, bp, k->fcoef.p + i0, -1.0, 1, n);
      PYN_LAUNCHED(ctx);
      lincomb_kernel<<<gridv, 256, 0, ctx->stream>>>(k->fad.p, bap, k->fcoef.p + i0, -1.0, 1, n);
      PYN_LAUNCHED(ctx);
    }
    PYN_CHECK(vec_dot_dev(ctx, k->fd.p, k->fad.p, n, nrm2));
    PYN_CHECK(vec_dot_dev(ctx, x, b, n, nrm2 + 2));
    if (dist) PYN_CHECK(comm_allreduce_sum(ctx, nrm2, 3));
    const double ftol = k->rtol * 100.0 > 1e-12 ? k->rtol * 100.0 : 1e-12;
    fischer_append_kernel<<<gridv, 256, 0, ctx->stream>>>(k->fd.p, k->fad.p, nrm2, ftol * ftol,
                                                          k->fx[k->fischer_n].p, k->fax[k->fischer_n].p, n,
                                                          nrm2 + 3);
    PYN_LAUNCHED(ctx);
  }
  PYN_CUDA(cudaStreamSynchronize(ctx->stream));
  PYN_CUDA(cudaMemcpyAsync(k->status_host, S, sizeof(double) * S_COUNT, cudaMemcpyDeviceToHost, ctx->stream));
  if (fischer)
    PYN_CUDA(cudaMemcpyAsync(k->status_host + S_COUNT, k->fcoef.p + kFischerCap + 3, sizeof(double),
                             cudaMemcpyDeviceToHost, ctx->stream));
  PYN_CUDA(cudaStreamSynchronize(ctx->stream));
  if (fischer && k->status_host[S_COUNT] != 0.0) k->fischer_n++;
  const double* st = k->status_host;
  k->last_its = (int)st[S_ITS];
  k->hint_its = -1;
  if (trace) {
    const double tr3 = trace_now_us();
    fprintf(stderr, "[ksp_trace] its %d launched %d chunks %d gap_before %.0f setup %.0f loop %.0f epilogue %.0f total %.0f\n",
            (int)st[S_ITS], it, tr_chunks, g_trace_last_return > 0 ? tr0 - g_trace_last_return : 0.0, tr1 - tr0,
            tr2 - tr1, tr3 - tr2, tr3 - tr0);
    g_trace_last_return = tr3;
  }
  if (its) *its = (int)st[S_ITS];
  if (reason) *reason = st[S_DONE] != 0.0 ? (int)st[S_REASON] : -3;
  if (rnorm) *rnorm = st[S_RNORM];
  return PYN_OK;
}


// ---- BiCGStab and GMRES(m) ------------------------------------------------------------------
// The KLE operator is SPD, so CG is the production solver; these two exist because the
// reference's KspSolver defaults to gmres (kle_solver.py:58) and honours -ksp_type from the
// options database.  Both are left-preconditioned like PETSc's defaults and test the
// preconditioned residual norm against max(rtol*||M^-1 b||, atol).  Vector work runs on the
// device (SpMV, fused multi-dot / linear-combination kernels); the handful of recurrence
// scalars is kept on the host (one small read-back per dot group).
template <int B>
__global__ void pc_apply_kernel(int pc, const double* __restrict__ pcd, const double* __restrict__ r,
                                double* __restrict__ z, int64_t nnodes) {
  for (int64_t n = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; n < nnodes;
       n += (int64_t)gridDim.x * blockDim.x) {
    double rr[B], zz[B];
#pragma unroll
    for (int i = 0; i < B; ++i) rr[i] = r[n * B + i];
    apply_pc_node<B>(pc, pcd, n, rr, zz);
#pragma unroll
    for (int i = 0; i < B; ++i) z[n * B + i] = zz[i];
  }
}

template <int B>
struct KrylovOps {
  pyn_ksp* k;
  pyn_ctx* ctx;
  int64_t nn, n;
  bool dist;
  int gridn;
  // z = M^-1 (A x)     (tmp = k->w)
  int apply(const double* x, double* z) {
    PYN_CHECK(spmv(k->A, 1.0, x, 0.0, nullptr, k->w.p));
    return pc(k->w.p, z);
  }
  int pc(const double* r, double* z) {
    if (k->pc == PYN_PC_ASM) return asm_apply(ctx, k->asmpc, const_cast<double*>(r), z, nullptr, nullptr);
    pc_apply_kernel<B><<<gridn, 256, 0, ctx->stream>>>(k->pc, k->pcd.p, r, z, nn);
    PYN_LAUNCHED(ctx);
    return PYN_OK;
  }
  int dot(const double* a, const double* b, double* out) { return vec_dot(ctx, a, b, n, out, dist); }
};

template <int B>
static int ksp_solve_bcgs(pyn_ksp* k, const double* b, double* x, int* its, int* reason, double* rnorm) {
  pyn_ctx* ctx = k->ctx;
  KrylovOps<B> op;
  op.k = k;
  op.ctx = ctx;
  op.dist = k->A->halo != nullptr && ctx->nranks > 1;
  op.nn = k->A->halo ? k->A->halo->n_own : k->nnodes;
  op.n = op.nn * B;
  op.gridn = grid_for(ctx, op.nn > 0 ? op.nn : 1, 256);
  const int64_t n = op.n;
  if (!k->pc_ready) PYN_CHECK(ksp_setup_pc<B>(k));
  for (DevBuf<double>* v : {&k->kv[0], &k->kv[1], &k->kv[2], &k->kv[3]})
    if (v->n != k->n) PYN_CHECK(v->alloc(ctx, k->n));
  double *r = k->r.p, *rh = k->kv[0].p, *p = k->p.p, *v = k->kv[1].p, *s = k->kv[2].p, *t = k->kv[3].p;
  double bn2 = 0, rn2 = 0;
  PYN_CHECK(op.pc(b, k->z.p));
  PYN_CHECK(op.dot(k->z.p, k->z.p, &bn2));
  PYN_CHECK(spmv(k->A, -1.0, x, 1.0, b, k->w.p));          // w = b - A x
  PYN_CHECK(op.pc(k->w.p, r));
  PYN_CHECK(op.dot(r, r, &rn2));
  double ttol = k->rtol * sqrt(bn2);
  int why_ok = 2;
  if (k->atol > ttol) {
    ttol = k->atol;
    why_ok = 3;
  }
  const double dtol = k->dtol * (bn2 > 0 ? sqrt(bn2) : 1.0);
  PYN_CUDA(cudaMemcpyAsync(rh, r, sizeof(double) * n, cudaMemcpyDeviceToDevice, ctx->stream));
  PYN_CHECK(vec_set(ctx, p, n, 0.0));
  PYN_CHECK(vec_set(ctx, v, n, 0.0));
  double rho = 1, alpha = 1, omega = 1, rn = sqrt(rn2);
  int it = 0, why = 0;
  if (rn <= ttol) why = why_ok;
  while (!why) {
    double rho1 = 0, rhv = 0, ts[2] = {0, 0};
    PYN_CHECK(op.dot(rh, r, &rho1));
    if (rho1 == 0.0 || omega == 0.0) {
      why = -9;
      break;
    }
    const double beta = (rho1 / rho) * (alpha / omega);
    // p = r + beta (p - omega v)
    PYN_CHECK(vec_axpby(ctx, p, -beta * omega, v, beta, n));
    PYN_CHECK(vec_axpby(ctx, p, 1.0, r, 1.0, n));
    PYN_CHECK(op.apply(p, v));
    PYN_CHECK(op.dot(rh, v, &rhv));
    if (rhv == 0.0) {
      why = -9;
      break;
    }
    alpha = rho1 / rhv;
    PYN_CHECK(vec_waxpy(ctx, s, -alpha, v, r, n));           // s = r - alpha v
    PYN_CHECK(op.apply(s, t));
    PYN_CHECK(op.dot(t, s, &ts[0]));
    PYN_CHECK(op.dot(t, t, &ts[1]));
    omega = ts[1] != 0.0 ? ts[0] / ts[1] : 0.0;
    PYN_CHECK(vec_axpby(ctx, x, alpha, p, 1.0, n));
    PYN_CHECK(vec_axpby(ctx, x, omega, s, 1.0, n));
    PYN_CHECK(vec_waxpy(ctx, r, -omega, t, s, n));           // r = s - omega t
    PYN_CHECK(op.dot(r, r, &rn2));
    rn = sqrt(rn2);
    rho = rho1;
    ++it;
    if (!(rn == rn)) why = -9;
    else if (rn <= ttol) why = why_ok;
    else if (rn >= dtol) why = -4;
    else if (it >= k->maxit) why = -3;
  }
  if (its) *its = it;
  if (reason) *reason = why;
  if (rnorm) *rnorm = rn;
  return PYN_OK;
}

template <int B>
static int ksp_solve_gmres(pyn_ksp* k, const double* b, double* x, int* its, int* reason, double* rnorm) {
  pyn_ctx* ctx = k->ctx;
  KrylovOps<B> op;
  op.k = k;
  op.ctx = ctx;
  op.dist = k->A->halo != nullptr && ctx->nranks > 1;
  op.nn = k->A->halo ? k->A->halo->n_own : k->nnodes;
  op.n = op.nn * B;
  op.gridn = grid_for(ctx, op.nn > 0 ? op.nn : 1, 256);
  const int64_t n = op.n;
  const int m = k->restart > 0 ? k->restart : 30;
  if (!k->pc_ready) PYN_CHECK(ksp_setup_pc<B>(k));
  if ((int)k->gv.size() != m + 1) {
    k->gv.clear();
    for (int i = 0; i <= m; ++i) {
      k->gv.emplace_back(new DevBuf<double>());
      PYN_CHECK(k->gv.back()->alloc(ctx, k->n));
    }
    PYN_CHECK(k->gcoef.alloc(ctx, kFischerMax));
  }
  const int gridv = grid_for(ctx, n > 0 ? n : 1, 256);
  double bn2 = 0;
  PYN_CHECK(op.pc(b, k->z.p));
  PYN_CHECK(op.dot(k->z.p, k->z.p, &bn2));
  double ttol = k->rtol * sqrt(bn2);
  int why_ok = 2;
  if (k->atol > ttol) {
    ttol = k->atol;
    why_ok = 3;
  }
  const double dtol = k->dtol * (bn2 > 0 ? sqrt(bn2) : 1.0);
  std::vector<double> H((size_t)(m + 1) * m, 0.0), cs(m, 0.0), sn(m, 0.0), g(m + 1, 0.0), y(m, 0.0), hcol(m + 2, 0.0);
  int it = 0, why = 0;
  double rn = 0;
  while (!why) {
    // r = M^-1 (b - A x); v0 = r / ||r||
    PYN_CHECK(spmv(k->A, -1.0, x, 1.0, b, k->w.p));
    PYN_CHECK(op.pc(k->w.p, k->gv[0]->p));
    double rn2 = 0;
    PYN_CHECK(op.dot(k->gv[0]->p, k->gv[0]->p, &rn2));
    rn = sqrt(rn2);
    if (!(rn == rn)) { why = -9; break; }
    if (rn <= ttol) { why = why_ok; break; }
    if (rn >= dtol) { why = -4; break; }
    if (it >= k->maxit) { why = -3; break; }
    PYN_CHECK(vec_scale(ctx, k->gv[0]->p, n, 1.0 / rn));
    std::fill(g.begin(), g.end(), 0.0);
    g[0] = rn;
    int j = 0;
    for (; j < m && !why; ++j) {
      double* w = k->gv[j + 1]->p;
      PYN_CHECK(op.apply(k->gv[j]->p, w));
      // classical Gram-Schmidt, two passes (h = V^T w; w -= V h), coefficients summed
      for (int i = 0; i <= j; ++i) hcol[i] = 0.0;
      for (int pass = 0; pass < 2; ++pass) {
        for (int i0 = 0; i0 <= j; i0 += kFischerMax) {
          BasisPtrs bp;
          bp.k = (j + 1 - i0) < kFischerMax ? (j + 1 - i0) : kFischerMax;
          for (int i = 0; i < bp.k; ++i) bp.v[i] = k->gv[i0 + i]->p;
          multi_dot_kernel<<<gridv, 256, 0, ctx->stream>>>(bp, w, n, ctx->red_partials, ctx->red_ticket, k->gcoef.p);
          PYN_LAUNCHED(ctx);
          if (op.dist) PYN_CHECK(comm_allreduce_sum(ctx, k->gcoef.p, kFischerMax));
          lincomb_kernel<<<gridv, 256, 0, ctx->stream>>>(w, bp, k->gcoef.p, -1.0, 1, n);
          PYN_LAUNCHED(ctx);
          double hc[kFischerMax];
          PYN_CHECK(k->gcoef.download(hc, kFischerMax));
          for (int i = 0; i < bp.k; ++i) hcol[i0 + i] += hc[i];
        }
      }
      double wn2 = 0;
      PYN_CHECK(op.dot(w, w, &wn2));
      const double hn = sqrt(wn2);
      hcol[j + 1] = hn;
      if (hn > 0) PYN_CHECK(vec_scale(ctx, w, n, 1.0 / hn));
      // Givens rotations on the new column
      for (int i = 0; i < j; ++i) {
        const double t0 = cs[i] * hcol[i] + sn[i] * hcol[i + 1];
        hcol[i + 1] = -sn[i] * hcol[i] + cs[i] * hcol[i + 1];
        hcol[i] = t0;
      }
      const double den = sqrt(hcol[j] * hcol[j] + hcol[j + 1] * hcol[j + 1]);
      cs[j] = den > 0 ? hcol[j] / den : 1.0;
      sn[j] = den > 0 ? hcol[j + 1] / den : 0.0;
      hcol[j] = den;
      g[j + 1] = -sn[j] * g[j];
      g[j] = cs[j] * g[j];
      for (int i = 0; i <= j; ++i) H[(size_t)i * m + j] = hcol[i];
      rn = fabs(g[j + 1]);
      ++it;
      if (!(rn == rn)) why = -9;
      else if (rn <= ttol) why = why_ok;
      else if (rn >= dtol) why = -4;
      else if (it >= k->maxit) why = -3;
      else if (hn == 0.0) { ++j; break; }                      // happy breakdown
    }
    // x += V y, H y = g (upper triangular, j columns)
    for (int i = j - 1; i >= 0; --i) {
      double sum = g[i];
      for (int l = i + 1; l < j; ++l) sum -= H[(size_t)i * m + l] * y[l];
      y[i] = H[(size_t)i * m + i] != 0.0 ? sum / H[(size_t)i * m + i] : 0.0;
    }
    for (int i0 = 0; i0 < j; i0 += kFischerMax) {
      BasisPtrs bp;
      bp.k = (j - i0) < kFischerMax ? (j - i0) : kFischerMax;
      double yc[kFischerMax] = {0, 0, 0, 0, 0, 0, 0, 0};
      for (int i = 0; i < bp.k; ++i) {
        bp.v[i] = k->gv[i0 + i]->p;
        yc[i] = y[i0 + i];
      }
      PYN_CHECK(k->gcoef.upload(yc, kFischerMax));
      lincomb_kernel<<<gridv, 256, 0, ctx->stream>>>(x, bp, k->gcoef.p, 1.0, 1, n);
      PYN_LAUNCHED(ctx);
    }
    PYN_CUDA(cudaStreamSynchronize(ctx->stream));
  }
  if (its) *its = it;
  if (reason) *reason = why;
  if (rnorm) *rnorm = rn;
  return PYN_OK;
}

extern "C" {

int pyn_ksp_create(pyn_ctx* ctx, pyn_mat* a, int type, int pc, pyn_ksp** out) {
  PYN_REQUIRE(ctx && a && out, "pyn_ksp_create: null argument");
  PYN_REQUIRE(a->br == a->bc && a->nbrows == a->nbcols, "KSP operator must be square");
  PYN_REQUIRE(pc >= PYN_PC_NONE && pc <= PYN_PC_ASM, "unknown pc type %d", pc);
  if (pc == PYN_PC_ASM) {
    PYN_REQUIRE(a->mesh && !a->free_pattern && (a->br == 2 || a->br == 3),
                "pc asm needs a 2-D or 3-D operator assembled from a mesh");
  }
  if (type != PYN_KSP_CG && type != PYN_KSP_GMRES && type != PYN_KSP_BCGS) {
    set_error("KSP type %d is not available (cg, gmres, bcgs)", type);
    return PYN_E_UNSUPPORTED;
  }
  pyn_ksp* k = new pyn_ksp();
  k->ctx = ctx;
  k->A = a;
  k->type = type;
  k->pc = pc;
  k->B = a->br;
  k->nnodes = a->nbrows;
  k->n = a->nbrows * a->br;
  int rc = PYN_OK;
  if (rc == PYN_OK) rc = k->r.alloc(ctx, k->n);
  if (rc == PYN_OK) rc = k->z.alloc(ctx, k->n);
  if (rc == PYN_OK) rc = k->p.alloc(ctx, k->n);
  if (rc == PYN_OK) rc = k->w.alloc(ctx, k->n);
  if (rc == PYN_OK) rc = k->S.alloc(ctx, S_COUNT);
  if (rc == PYN_OK) rc = k->tmp2.alloc(ctx, 4);
  if (rc == PYN_OK && cudaMallocHost((void**)&k->status_host, sizeof(double) * 2 * S_COUNT) != cudaSuccess) {
    set_error("cudaMallocHost failed");
    rc = PYN_E_NOMEM;
  }
  if (rc == PYN_OK) {
    cudaEventCreateWithFlags(&k->ev[0], cudaEventDisableTiming);
    cudaEventCreateWithFlags(&k->ev[1], cudaEventDisableTiming);
  }
  if (rc != PYN_OK) {
    pyn_ksp_destroy(k);
    return rc;
  }
  *out = k;
  return PYN_OK;
}

int pyn_ksp_set_tolerances(pyn_ksp* k, double rtol, double atol, int maxit, int restart) {
  PYN_REQUIRE(k, "null ksp");
  if (rtol >= 0) k->rtol = rtol;
  if (atol >= 0) k->atol = atol;
  if (maxit > 0) k->maxit = maxit;
  if (restart > 0) k->restart = restart;
  return PYN_OK;
}

int pyn_ksp_set_asm_coarse(pyn_ksp* k, int steps) {
  PYN_REQUIRE(k, "null ksp");
  PYN_REQUIRE(steps >= 0 && steps <= 64, "pc asm: coarse Chebyshev steps %d out of range [0,64]", steps);
  if (steps != k->asm_coarse) k->pc_ready = false;
  k->asm_coarse = steps;
  return PYN_OK;
}

int pyn_ksp_hint_iterations(pyn_ksp* k, int n) {
  PYN_REQUIRE(k, "null ksp");
  k->hint_its = n;
  return PYN_OK;
}
int pyn_ksp_set_cg_single_reduction(pyn_ksp* k, int on) {
  PYN_REQUIRE(k, "null ksp");
  k->single_reduction = on != 0;
  return PYN_OK;
}

int pyn_ksp_set_fischer_guess(pyn_ksp* k, int model, int size) {
  PYN_REQUIRE(k, "null ksp");
  PYN_REQUIRE(size == 0 || model == 1, "KSP Fischer guess: only model 1 (A-conjugate projection) is implemented");
  PYN_REQUIRE(size >= 0 && size <= kFischerCap, "KSP Fischer guess: size %d (max %d)", size, kFischerCap);
  pyn_ctx* ctx = k->ctx;
  k->fischer_m = size;
  k->fischer_n = 0;
  if (size == 0) return PYN_OK;
  for (int i = 0; i < size; ++i) {
    PYN_CHECK(k->fx[i].alloc(ctx, k->n));
    PYN_CHECK(k->fax[i].alloc(ctx, k->n));
  }
  PYN_CHECK(k->x0.alloc(ctx, k->n));
  PYN_CHECK(k->r0.alloc(ctx, k->n));
  PYN_CHECK(k->fd.alloc(ctx, k->n));
  PYN_CHECK(k->fad.alloc(ctx, k->n));
  PYN_CHECK(k->fcoef.alloc(ctx, kFischerCap + 4 + kFischerMax));
  PYN_CHECK(k->fcoef.zero());
  return PYN_OK;
}

int pyn_ksp_solve(pyn_ksp* k, pyn_vec* b, pyn_vec* x, int* its, int* reason, double* rnorm) {
  PYN_REQUIRE(k && b && x, "pyn_ksp_solve: null argument");
  PYN_REQUIRE(b->n == k->n && x->n == k->n, "pyn_ksp_solve: vector size %lld/%lld, operator %lld",
              (long long)b->n, (long long)x->n, (long long)k->n);
  PYN_REQUIRE(b != x, "pyn_ksp_solve: b and x must differ");
#define SOLVE(FN)                                                         \
  switch (k->B) {                                                         \
    case 1: return FN<1>(k, b->d.p, x->d.p, its, reason, rnorm);          \
    case 2: return FN<2>(k, b->d.p, x->d.p, its, reason, rnorm);          \
    case 3: return FN<3>(k, b->d.p, x->d.p, its, reason, rnorm);          \
  }
  if (k->type == PYN_KSP_GMRES) {
    SOLVE(ksp_solve_gmres)
  } else if (k->type == PYN_KSP_BCGS) {
    SOLVE(ksp_solve_bcgs)
  } else {
    SOLVE(ksp_solve_cg)
  }
#undef SOLVE
  set_error("unsupported block size %d", k->B);
  return PYN_E_UNSUPPORTED;
}

int pyn_ksp_destroy(pyn_ksp* k) {
  if (!k) return PYN_OK;
  if (k->status_host) cudaFreeHost(k->status_host);
  if (k->ev[0]) cudaEventDestroy(k->ev[0]);
  if (k->ev[1]) cudaEventDestroy(k->ev[1]);
  delete k;
  return PYN_OK;
}

}  // extern "C"
