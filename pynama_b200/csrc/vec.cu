// vec.cu -- device Vec operations (PETSc.Vec surface of the hot path)
// and the pointwise pieces of the vorticity-transport RHS (k8, k9, k10).
#include "vec.cuh"
#include "comm.cuh"

namespace pyn {

#define GRID_STRIDE(i, n)                                                              \
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < (n);           \
       i += (int64_t)gridDim.x * blockDim.x)

__global__ void __launch_bounds__(256)
dot_kernel(const double* __restrict__ x, const double* __restrict__ y, int64_t n,
           double* partials, unsigned int* ticket, double* out) {
  __shared__ double sm[32];
  double v[1] = {0.0};
  GRID_STRIDE(i, n) v[0] += x[i] * y[i];
  block_reduce<1>(v, sm);
  grid_reduce_finish<1>(v, partials, ticket, out);
}

__global__ void __launch_bounds__(256)
sum_kernel(const double* __restrict__ x, int64_t n, double* partials, unsigned int* ticket, double* out) {
  __shared__ double sm[32];
  double v[1] = {0.0};
  GRID_STRIDE(i, n) v[0] += x[i];
  block_reduce<1>(v, sm);
  grid_reduce_finish<1>(v, partials, ticket, out);
}

int vec_dot_dev(pyn_ctx* ctx, const double* x, const double* y, int64_t n, double* out_dev) {
  dot_kernel<<<grid_for(ctx, n, 256 * 4), 256, 0, ctx->stream>>>(x, y, n, ctx->red_partials,
                                                                 ctx->red_ticket, out_dev);
  PYN_LAUNCHED(ctx);
  return PYN_OK;
}

static int fetch_scalar(pyn_ctx* ctx, double* out_host) {
  PYN_CUDA(cudaMemcpyAsync(ctx->red_result_host, ctx->red_result_dev, sizeof(double),
                           cudaMemcpyDeviceToHost, ctx->stream));
  PYN_CUDA(cudaStreamSynchronize(ctx->stream));
  *out_host = ctx->red_result_host[0];
  return PYN_OK;
}

int vec_dot(pyn_ctx* ctx, const double* x, const double* y, int64_t n, double* out_host, bool dist) {
  PYN_CHECK(vec_dot_dev(ctx, x, y, n, ctx->red_result_dev));
  if (dist) PYN_CHECK(comm_allreduce_sum(ctx, ctx->red_result_dev, 1));
  return fetch_scalar(ctx, out_host);
}

int vec_sum(pyn_ctx* ctx, const double* x, int64_t n, double* out_host, bool dist) {
  sum_kernel<<<grid_for(ctx, n, 256 * 4), 256, 0, ctx->stream>>>(x, n, ctx->red_partials,
                                                                 ctx->red_ticket, ctx->red_result_dev);
  PYN_LAUNCHED(ctx);
  if (dist) PYN_CHECK(comm_allreduce_sum(ctx, ctx->red_result_dev, 1));
  return fetch_scalar(ctx, out_host);
}

__global__ void set_kernel(double* x, int64_t n, double a) { GRID_STRIDE(i, n) x[i] = a; }
__global__ void scale_kernel(double* x, int64_t n, double a) { GRID_STRIDE(i, n) x[i] *= a; }
__global__ void axpby_kernel(double* y, double a, const double* x, double b, int64_t n) {
  if (b == 1.0) {
    GRID_STRIDE(i, n) y[i] += a * x[i];          // VecAXPY
  } else if (a == 1.0) {
    GRID_STRIDE(i, n) y[i] = x[i] + b * y[i];    // VecAYPX
  } else {
    GRID_STRIDE(i, n) y[i] = a * x[i] + b * y[i];
  }
}
__global__ void waxpy_kernel(double* w, double a, const double* x, const double* y, int64_t n) {
  GRID_STRIDE(i, n) w[i] = a * x[i] + y[i];
}
__global__ void pmult_kernel(double* w, const double* x, const double* y, int64_t n) {
  GRID_STRIDE(i, n) w[i] = x[i] * y[i];
}
__global__ void recip_kernel(double* x, int64_t n) {
  GRID_STRIDE(i, n) {
    const double v = x[i];
    if (v != 0.0) x[i] = 1.0 / v;     // VecReciprocal leaves zeros alone
  }
}

#define EW_LAUNCH(kernel, n, ...)                                                     \
  do {                                                                                \
    if ((n) > 0) {                                                                    \
      kernel<<<grid_for(ctx, (n), 256), 256, 0, ctx->stream>>>(__VA_ARGS__);          \
      PYN_LAUNCHED(ctx);                                                              \
    }                                                                                 \
  } while (0)

int vec_set(pyn_ctx* ctx, double* x, int64_t n, double a) {
  EW_LAUNCH(set_kernel, n, x, n, a);
  return PYN_OK;
}
int vec_scale(pyn_ctx* ctx, double* x, int64_t n, double a) {
  EW_LAUNCH(scale_kernel, n, x, n, a);
  return PYN_OK;
}
int vec_axpby(pyn_ctx* ctx, double* y, double a, const double* x, double b, int64_t n) {
  EW_LAUNCH(axpby_kernel, n, y, a, x, b, n);
  return PYN_OK;
}
int vec_waxpy(pyn_ctx* ctx, double* w, double a, const double* x, const double* y, int64_t n) {
  EW_LAUNCH(waxpy_kernel, n, w, a, x, y, n);
  return PYN_OK;
}
int vec_pointwise_mult(pyn_ctx* ctx, double* w, const double* x, const double* y, int64_t n) {
  EW_LAUNCH(pmult_kernel, n, w, x, y, n);
  return PYN_OK;
}
int vec_reciprocal(pyn_ctx* ctx, double* x, int64_t n) {
  EW_LAUNCH(recip_kernel, n, x, n);
  return PYN_OK;
}

// y += sum_k a[k] x_k in one pass over y (RK stage combination, k11)
constexpr int kMaxMaxpy = 8;
struct MaxpyArgs {
  const double* x[kMaxMaxpy];
  double a[kMaxMaxpy];
  int k;
};
__global__ void maxpy_kernel(double* y, MaxpyArgs args, int64_t n) {
  GRID_STRIDE(i, n) {
    double s = y[i];
    for (int k = 0; k < args.k; ++k) s += args.a[k] * args.x[k][i];
    y[i] = s;
  }
}
int vec_maxpy(pyn_ctx* ctx, double* y, int k, const double* a, const double* const* x, int64_t n) {
  for (int k0 = 0; k0 < k; k0 += kMaxMaxpy) {
    MaxpyArgs args;
    args.k = (k - k0) < kMaxMaxpy ? (k - k0) : kMaxMaxpy;
    for (int j = 0; j < args.k; ++j) {
      args.x[j] = x[k0 + j];
      args.a[j] = a[k0 + j];
    }
    EW_LAUNCH(maxpy_kernel, n, y, args, n);
  }
  return PYN_OK;
}

__global__ void scatter_kernel(double* v, int64_t k, const int64_t* idx, const double* vals, int add) {
  GRID_STRIDE(i, k) {
    if (add)
      atomicAdd(&v[idx[i]], vals[i]);
    else
      v[idx[i]] = vals[i];
  }
}
__global__ void gather_kernel(const double* v, int64_t k, const int64_t* idx, double* vals) {
  GRID_STRIDE(i, k) vals[i] = v[idx[i]];
}
int vec_scatter(pyn_ctx* ctx, double* v, int64_t k, const int64_t* idx, const double* vals, int add) {
  EW_LAUNCH(scatter_kernel, k, v, k, idx, vals, add);
  return PYN_OK;
}
int vec_gather(pyn_ctx* ctx, const double* v, int64_t k, const int64_t* idx, double* vals) {
  EW_LAUNCH(gather_kernel, k, v, k, idx, vals);
  return PYN_OK;
}

// ---- RHS pointwise pieces ---------------------------------------------------------
// computeVtensV (base_problem.py:139-154): [xx, xy, yy] / [xx, xy, yy, yz, zz, zx]
template <int DIM>
__device__ __forceinline__ void vtensv_node(const double* v, double* t) {
  if (DIM == 2) {
    t[0] = v[0] * v[0];
    t[1] = v[0] * v[1];
    t[2] = v[1] * v[1];
  } else {
    t[0] = v[0] * v[0];
    t[1] = v[0] * v[1];
    t[2] = v[1] * v[1];
    t[3] = v[1] * v[2];
    t[4] = v[2] * v[2];
    t[5] = v[2] * v[0];
  }
}
template <int DIM>
__global__ void vtensv_kernel(const double* __restrict__ vel, double* __restrict__ out, int64_t nnodes) {
  constexpr int DS = DIM == 2 ? 3 : 6;
  GRID_STRIDE(i, nnodes) {
    double v[DIM], t[DS];
    for (int d = 0; d < DIM; ++d) v[d] = vel[i * DIM + d];
    vtensv_node<DIM>(v, t);
    for (int s = 0; s < DS; ++s) out[i * DS + s] = t[s];
  }
}
// aux = (2 mu) * aux + (-rho) * vtensv(vel)    (base_problem.py:127-130)
template <int DIM>
__global__ void stress_kernel(double* __restrict__ aux, const double* __restrict__ vel, double two_mu,
                              double mrho, int64_t nnodes) {
  constexpr int DS = DIM == 2 ? 3 : 6;
  GRID_STRIDE(i, nnodes) {
    double v[DIM], t[DS];
    for (int d = 0; d < DIM; ++d) v[d] = vel[i * DIM + d];
    vtensv_node<DIM>(v, t);
    for (int s = 0; s < DS; ++s) {
      double a = aux[i * DS + s] * two_mu;       // _Aux1 *= 2 mu
      aux[i * DS + s] = a + mrho * t[s];         // _Aux1.axpy(-rho, VtensV)
    }
  }
}

int vtensv(pyn_ctx* ctx, const double* vel, double* out, int64_t nnodes, int dim) {
  if (dim == 2)
    EW_LAUNCH(vtensv_kernel<2>, nnodes, vel, out, nnodes);
  else
    EW_LAUNCH(vtensv_kernel<3>, nnodes, vel, out, nnodes);
  return PYN_OK;
}
int stress_combine(pyn_ctx* ctx, double* aux, const double* vel, double mu, double rho,
                   int64_t nnodes, int dim) {
  if (dim == 2)
    EW_LAUNCH(stress_kernel<2>, nnodes, aux, vel, 2.0 * mu, -1.0 * rho, nnodes);
  else
    EW_LAUNCH(stress_kernel<3>, nnodes, aux, vel, 2.0 * mu, -1.0 * rho, nnodes);
  return PYN_OK;
}

}  // namespace pyn
