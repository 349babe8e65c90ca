// matfree.cu -- matrix-free product with the KLE operator K (see matfree.cuh).
//
// The bilinear form is the one the element kernel integrates (spectral.py:116-159):
//   full rule     sum_g w|J| grad(v):grad(u)
//   reduced rule  sum_g w|J| (alpha_d div(v) div(u) + alpha_w curl(v).curl(u))
// with Dirichlet / no-slip dofs eliminated the way MatFS.buildFS / MatNS.buildNS do it
// (mat_fs.py:179-189, mat_ns.py:116-130): constrained columns dropped, constrained rows = identity.
#include "matfree.cuh"
#include "elem_math.cuh"
#include "vec.cuh"

pyn_mesh::~pyn_mesh() { delete mf; }

namespace pyn {

// 1-D tables [point][node] of the two rules in constant memory: with the contraction loops
// fully unrolled every table entry is a constant-bank operand of its DFMA (no load at all).
struct MfTables {
  double Hf[kMaxPts1d][kMaxNgl], Df[kMaxPts1d][kMaxNgl];
  double Hr[kMaxPts1d][kMaxNgl], Dr[kMaxPts1d][kMaxNgl];
};
__constant__ MfTables c_mf;
static int g_mf_key = -1;     // ngl whose tables c_mf holds

static int mf_upload_tables(const pyn_mesh* mesh) {
  if (g_mf_key == mesh->ngl) return PYN_OK;
  MfTables t;
  memset(&t, 0, sizeof(t));
  const Rule1D& f = mesh->htab.rule[RULE_FULL];
  const Rule1D& r = mesh->htab.rule[RULE_RED];
  memcpy(t.Hf, f.h, sizeof(t.Hf));
  memcpy(t.Df, f.dh, sizeof(t.Df));
  memcpy(t.Hr, r.h, sizeof(t.Hr));
  memcpy(t.Dr, r.dh, sizeof(t.Dr));
  // stream-ordered: products already queued keep the tables they were launched with
  PYN_CUDA(cudaMemcpyToSymbolAsync(c_mf, &t, sizeof(t), 0, cudaMemcpyHostToDevice, mesh->ctx->stream));
  PYN_CUDA(cudaStreamSynchronize(mesh->ctx->stream));      // t is a stack object
  g_mf_key = mesh->ngl;
  return PYN_OK;
}

// J^-1 and w|J| at every quadrature point of every element, points in tensor order
// (p0 fastest), full rule first.  Same arithmetic as the element kernel (point_geometry).
__global__ void mf_geometry_kernel(const ElemTables* __restrict__ tab, const double* __restrict__ corners,
                                   int64_t nel, int npf, int npr, double* __restrict__ geo) {
  const int64_t npts = (int64_t)npf * npf * npf + (int64_t)npr * npr * npr;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < nel * npts;
       i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t e = i / npts;
    int pt = (int)(i - e * npts);
    const bool full = pt < npf * npf * npf;
    const Rule1D& rule = tab->rule[full ? RULE_FULL : RULE_RED];
    const int np = full ? npf : npr;
    int q = full ? pt : pt - npf * npf * npf;
    uint8_t gt[3];
    gt[0] = (uint8_t)(q % np);
    gt[1] = (uint8_t)((q / np) % np);
    gt[2] = (uint8_t)(q / (np * np));
    double X[24], jinv[9], det;
    for (int k = 0; k < 24; ++k) X[k] = corners[e * 24 + k];
    point_geometry_t<3>(rule, tab->corn_t, gt, X, jinv, &det);
    const double wd = ((rule.w[gt[0]] * rule.w[gt[1]]) * rule.w[gt[2]]) * det;
    double* g = geo + e * 10 * npts + pt;
#pragma unroll
    for (int k = 0; k < 9; ++k) g[k * npts] = jinv[k];
    g[9 * npts] = wd;
  }
}

int matfree_create(pyn_mesh* mesh, pyn_matfree** out) {
  pyn_ctx* ctx = mesh->ctx;
  PYN_REQUIRE(mesh->dim == 3, "matrix-free K: 3-D meshes only");
  PYN_REQUIRE(mesh->ngl >= 2 && mesh->ngl <= 6, "matrix-free K: ngl 2..6");
  PYN_REQUIRE(mesh->has_cls, "matrix-free K: set the dof classes first");
  pyn_matfree* mf = new pyn_matfree();
  mf->mesh = mesh;
  mf->npf = mesh->htab.rule[RULE_FULL].npts1d;
  mf->npr = mesh->htab.rule[RULE_RED].npts1d;
  mf->npts = (int64_t)mf->npf * mf->npf * mf->npf + (int64_t)mf->npr * mf->npr * mf->npr;
  const int nn = mesh->nn, n = mesh->ngl;
  int rc = PYN_OK;
  // tensor position of every reference-local node, and the element node table in tensor order
  std::vector<int32_t> loc(nn);
  for (int a = 0; a < nn; ++a) {
    const uint8_t* t = mesh->htab.node_t[a];
    loc[t[0] + n * (t[1] + n * t[2])] = a;
  }
  std::vector<int32_t> en((size_t)mesh->nel * nn), ent((size_t)mesh->nel * nn);
  if (rc == PYN_OK) rc = mesh->elem_nodes.download(en.data(), mesh->nel * nn);
  for (int64_t e = 0; e < mesh->nel && rc == PYN_OK; ++e)
    for (int t = 0; t < nn; ++t) ent[e * nn + t] = en[e * nn + loc[t]];
  if (rc == PYN_OK) rc = mf->loc_of_t.alloc(ctx, nn);
  if (rc == PYN_OK) rc = mf->loc_of_t.upload(loc.data(), nn);
  if (rc == PYN_OK) rc = mf->nodes_t.alloc(ctx, mesh->nel * nn);
  if (rc == PYN_OK) rc = mf->nodes_t.upload(ent.data(), mesh->nel * nn);
  if (rc == PYN_OK) rc = mf->geo.alloc(ctx, mesh->nel * 10 * mf->npts);
  if (rc == PYN_OK) rc = mf->ye.alloc(ctx, mesh->nel * nn * 3);
  if (rc == PYN_OK) {
    mf_geometry_kernel<<<grid_for(ctx, mesh->nel * mf->npts, 128), 128, 0, ctx->stream>>>(
        mesh->dtab.p, mesh->corners.p, mesh->nel, mf->npf, mf->npr, mf->geo.p);
    ctx->launches++;
    if (cudaGetLastError() != cudaSuccess || cudaStreamSynchronize(ctx->stream) != cudaSuccess) {
      set_error("matrix-free K: geometry kernel failed");
      rc = PYN_E_CUDA;
    }
  }
  if (rc != PYN_OK) {
    delete mf;
    return rc;
  }
  *out = mf;
  return PYN_OK;
}

// ---- element kernel ---------------------------------------------------------------------------
// One warp per element, everything of the element in a private shared-memory slice.  Scalar
// arrays are N x N x N with strides (1, S1, S2), S1 odd so that lanes walking the second index
// hit different banks.
template <int N>
struct MfShape {
  static constexpr int S1 = N | 1;
  static constexpr int S2 = S1 * N;
  static constexpr int SC = S2 * N;             // one padded scalar array
  static constexpr int NN = N * N * N;
  static constexpr int WARP_DOUBLES = 21 * SC;  // u (3) + buffer A (9) + buffer B (9)
  static constexpr int SLOTS = (3 * N * N + 31) / 32;
};

template <bool FULL>
__device__ __forceinline__ double tH(int p, int i) { return FULL ? c_mf.Hf[p][i] : c_mf.Hr[p][i]; }
template <bool FULL>
__device__ __forceinline__ double tD(int p, int i) { return FULL ? c_mf.Df[p][i] : c_mf.Dr[p][i]; }

// nodes -> points along x: UH = H u, UD = D u.  in: u[c] (3 arrays); out: UH[c], UD[c] (6 arrays)
template <int N, int NP, bool FULL>
__device__ __forceinline__ void mf_fwd_x(const double* __restrict__ su, double* __restrict__ s1, int lane) {
  using S = MfShape<N>;
#pragma unroll
  for (int s = 0; s < S::SLOTS; ++s) {
    const int l = lane + 32 * s;
    if (l < 3 * N * N) {
      const int j = l % N, k = (l / N) % N, c = l / (N * N);
      const int off = c * S::SC + S::S1 * j + S::S2 * k;
      double u[N];
#pragma unroll
      for (int i = 0; i < N; ++i) u[i] = su[off + i];
#pragma unroll
      for (int p = 0; p < NP; ++p) {
        double a = 0.0, b = 0.0;
#pragma unroll
        for (int i = 0; i < N; ++i) {
          a = fma(tH<FULL>(p, i), u[i], a);
          b = fma(tD<FULL>(p, i), u[i], b);
        }
        s1[off + p] = a;
        s1[off + 3 * S::SC + p] = b;
      }
    }
  }
}

// along y: HH = H UH, HD = D UH, DH = H UD.  in: s1 (6 arrays); out: s2 = HH[c], HD[c], DH[c] (9 arrays)
template <int N, int NP, bool FULL>
__device__ __forceinline__ void mf_fwd_y(const double* __restrict__ s1, double* __restrict__ s2, int lane) {
  using S = MfShape<N>;
  constexpr int L = 3 * NP * N;
  for (int l = lane; l < L; l += 32) {
    const int p0 = l % NP, k = (l / NP) % N, c = l / (NP * N);
    const int off = c * S::SC + p0 + S::S2 * k;
    double uh[N], ud[N];
#pragma unroll
    for (int j = 0; j < N; ++j) {
      uh[j] = s1[off + S::S1 * j];
      ud[j] = s1[off + 3 * S::SC + S::S1 * j];
    }
#pragma unroll
    for (int p = 0; p < NP; ++p) {
      double hh = 0.0, hd = 0.0, dh = 0.0;
#pragma unroll
      for (int j = 0; j < N; ++j) {
        hh = fma(tH<FULL>(p, j), uh[j], hh);
        hd = fma(tD<FULL>(p, j), uh[j], hd);
        dh = fma(tH<FULL>(p, j), ud[j], dh);
      }
      s2[off + S::S1 * p] = hh;
      s2[off + 3 * S::SC + S::S1 * p] = hd;
      s2[off + 6 * S::SC + S::S1 * p] = dh;
    }
  }
}

// along z: R[c][0] = H DH, R[c][1] = H HD, R[c][2] = D HH.  out: sR[(c*3+d)] (9 arrays)
template <int N, int NP, bool FULL>
__device__ __forceinline__ void mf_fwd_z(const double* __restrict__ s2, double* __restrict__ sR, int lane) {
  using S = MfShape<N>;
  constexpr int L = 3 * NP * NP;
  for (int l = lane; l < L; l += 32) {
    const int p0 = l % NP, p1 = (l / NP) % NP, c = l / (NP * NP);
    const int off = p0 + S::S1 * p1;
    double hh[N], hd[N], dh[N];
#pragma unroll
    for (int k = 0; k < N; ++k) {
      hh[k] = s2[c * S::SC + off + S::S2 * k];
      hd[k] = s2[(3 + c) * S::SC + off + S::S2 * k];
      dh[k] = s2[(6 + c) * S::SC + off + S::S2 * k];
    }
#pragma unroll
    for (int p = 0; p < NP; ++p) {
      double r0 = 0.0, r1 = 0.0, r2 = 0.0;
#pragma unroll
      for (int k = 0; k < N; ++k) {
        r0 = fma(tH<FULL>(p, k), dh[k], r0);
        r1 = fma(tH<FULL>(p, k), hd[k], r1);
        r2 = fma(tD<FULL>(p, k), hh[k], r2);
      }
      sR[(c * 3 + 0) * S::SC + off + S::S2 * p] = r0;
      sR[(c * 3 + 1) * S::SC + off + S::S2 * p] = r1;
      sR[(c * 3 + 2) * S::SC + off + S::S2 * p] = r2;
    }
  }
}

// point-wise: reference gradients R[c][d] -> reference fluxes T[c][d] (in place)
template <int N, int NP, bool FULL>
__device__ __forceinline__ void mf_pointwise(double* __restrict__ sR, const double* __restrict__ geo,
                                             int64_t npts, int lane) {
  using S = MfShape<N>;
  constexpr int NPTS = NP * NP * NP;
  for (int g = lane; g < NPTS; g += 32) {
    const int p0 = g % NP, p1 = (g / NP) % NP, p2 = g / (NP * NP);
    const int off = p0 + S::S1 * p1 + S::S2 * p2;
    double ji[9], R[3][3], H[3][3], F[3][3];
#pragma unroll
    for (int q = 0; q < 9; ++q) ji[q] = __ldg(geo + (int64_t)q * npts + g);
    const double wd = __ldg(geo + 9 * npts + g);
#pragma unroll
    for (int c = 0; c < 3; ++c)
#pragma unroll
      for (int d = 0; d < 3; ++d) R[c][d] = sR[(c * 3 + d) * S::SC + off];
    // H[c][k] = d u_c / d x_k = sum_d J^-1[k][d] R[c][d]
#pragma unroll
    for (int c = 0; c < 3; ++c)
#pragma unroll
      for (int k = 0; k < 3; ++k)
        H[c][k] = fma(ji[k * 3 + 2], R[c][2], fma(ji[k * 3 + 1], R[c][1], ji[k * 3] * R[c][0]));
    if (FULL) {
#pragma unroll
      for (int c = 0; c < 3; ++c)
#pragma unroll
        for (int k = 0; k < 3; ++k) F[c][k] = wd * H[c][k];
    } else {
      const double dv = kAlphaD * (H[0][0] + H[1][1] + H[2][2]);
      const double w0 = kAlphaW * (H[2][1] - H[1][2]);
      const double w1 = kAlphaW * (H[0][2] - H[2][0]);
      const double w2 = kAlphaW * (H[1][0] - H[0][1]);
      // F[c][k] = w|J| (alpha_d delta_ck div u + alpha_w sum_r C_rc(e_k) curl_r u)
      F[0][0] = wd * dv;  F[0][1] = wd * -w2; F[0][2] = wd * w1;
      F[1][0] = wd * w2;  F[1][1] = wd * dv;  F[1][2] = wd * -w0;
      F[2][0] = wd * -w1; F[2][1] = wd * w0;  F[2][2] = wd * dv;
    }
    // T[c][d] = sum_k J^-1[k][d] F[c][k]
#pragma unroll
    for (int c = 0; c < 3; ++c)
#pragma unroll
      for (int d = 0; d < 3; ++d)
        sR[(c * 3 + d) * S::SC + off] = fma(ji[6 + d], F[c][2], fma(ji[3 + d], F[c][1], ji[d] * F[c][0]));
  }
}

// transposed passes: points -> nodes
template <int N, int NP, bool FULL>
__device__ __forceinline__ void mf_bwd_z(const double* __restrict__ sT, double* __restrict__ sA, int lane) {
  using S = MfShape<N>;
  constexpr int L = 3 * NP * NP;
  for (int l = lane; l < L; l += 32) {
    const int p0 = l % NP, p1 = (l / NP) % NP, c = l / (NP * NP);
    const int off = p0 + S::S1 * p1;
    double t0[NP], t1[NP], t2[NP];
#pragma unroll
    for (int p = 0; p < NP; ++p) {
      t0[p] = sT[(c * 3 + 0) * S::SC + off + S::S2 * p];
      t1[p] = sT[(c * 3 + 1) * S::SC + off + S::S2 * p];
      t2[p] = sT[(c * 3 + 2) * S::SC + off + S::S2 * p];
    }
#pragma unroll
    for (int k = 0; k < N; ++k) {
      double a0 = 0.0, a1 = 0.0, a2 = 0.0;
#pragma unroll
      for (int p = 0; p < NP; ++p) {
        a0 = fma(tH<FULL>(p, k), t0[p], a0);
        a1 = fma(tH<FULL>(p, k), t1[p], a1);
        a2 = fma(tD<FULL>(p, k), t2[p], a2);
      }
      sA[c * S::SC + off + S::S2 * k] = a0;
      sA[(3 + c) * S::SC + off + S::S2 * k] = a1;
      sA[(6 + c) * S::SC + off + S::S2 * k] = a2;
    }
  }
}

template <int N, int NP, bool FULL>
__device__ __forceinline__ void mf_bwd_y(const double* __restrict__ sA, double* __restrict__ sC, int lane) {
  using S = MfShape<N>;
  constexpr int L = 3 * NP * N;
  for (int l = lane; l < L; l += 32) {
    const int p0 = l % NP, k = (l / NP) % N, c = l / (NP * N);
    const int off = p0 + S::S2 * k;
    double a0[NP], a1[NP], a2[NP];
#pragma unroll
    for (int p = 0; p < NP; ++p) {
      a0[p] = sA[c * S::SC + off + S::S1 * p];
      a1[p] = sA[(3 + c) * S::SC + off + S::S1 * p];
      a2[p] = sA[(6 + c) * S::SC + off + S::S1 * p];
    }
#pragma unroll
    for (int j = 0; j < N; ++j) {
      double c0 = 0.0, c1 = 0.0;
#pragma unroll
      for (int p = 0; p < NP; ++p) {
        c0 = fma(tH<FULL>(p, j), a0[p], c0);
        c1 = fma(tD<FULL>(p, j), a1[p], c1);
        c1 = fma(tH<FULL>(p, j), a2[p], c1);
      }
      sC[c * S::SC + off + S::S1 * j] = c0;
      sC[(3 + c) * S::SC + off + S::S1 * j] = c1;
    }
  }
}

template <int N, int NP, bool FULL>
__device__ __forceinline__ void mf_bwd_x(const double* __restrict__ sC, double (&yacc)[MfShape<N>::SLOTS][N],
                                         int lane) {
  using S = MfShape<N>;
#pragma unroll
  for (int s = 0; s < S::SLOTS; ++s) {
    const int l = lane + 32 * s;
    if (l < 3 * N * N) {
      const int j = l % N, k = (l / N) % N, c = l / (N * N);
      const int off = c * S::SC + S::S1 * j + S::S2 * k;
      double c0[NP], c1[NP];
#pragma unroll
      for (int p = 0; p < NP; ++p) {
        c0[p] = sC[off + p];
        c1[p] = sC[off + 3 * S::SC + p];
      }
#pragma unroll
      for (int i = 0; i < N; ++i) {
        double y = yacc[s][i];
#pragma unroll
        for (int p = 0; p < NP; ++p) {
          y = fma(tD<FULL>(p, i), c0[p], y);
          y = fma(tH<FULL>(p, i), c1[p], y);
        }
        yacc[s][i] = y;
      }
    }
  }
}

template <int N, int NP, bool FULL>
__device__ __forceinline__ void mf_rule(const double* su, double* bufA, double* bufB,
                                        const double* geo, int64_t npts,
                                        double (&yacc)[MfShape<N>::SLOTS][N], int lane) {
  mf_fwd_x<N, NP, FULL>(su, bufB, lane);
  __syncwarp();
  mf_fwd_y<N, NP, FULL>(bufB, bufA, lane);
  __syncwarp();
  mf_fwd_z<N, NP, FULL>(bufA, bufB, lane);
  __syncwarp();
  mf_pointwise<N, NP, FULL>(bufB, geo, npts, lane);
  __syncwarp();
  mf_bwd_z<N, NP, FULL>(bufB, bufA, lane);
  __syncwarp();
  mf_bwd_y<N, NP, FULL>(bufA, bufB, lane);
  __syncwarp();
  mf_bwd_x<N, NP, FULL>(bufB, yacc, lane);
  __syncwarp();
}

template <int N, int WPC>
__global__ void __launch_bounds__(WPC * 32)
mf_apply_kernel(int64_t nel, const int32_t* __restrict__ nodes_t, const int32_t* __restrict__ loc_of_t,
                const uint8_t* __restrict__ cls, const double* __restrict__ geo, int64_t npts,
                const double* __restrict__ x, double* __restrict__ ye, const double* __restrict__ skip,
                HaloWait hw) {
  using S = MfShape<N>;
  extern __shared__ double mf_smem[];
  if (skip && *skip != 0.0) return;
  const double* xg = nullptr;          // ghost section of x in the halo window (fused exchange)
  if (hw.win) {
    if (!halo_wait_cta(hw)) return;
    xg = halo_buf(hw.win, (int)(hw.seq & 1ull), hw.nghost, hw.max_bs);
  }
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  const int64_t e = (int64_t)blockIdx.x * WPC + wid;
  if (e >= nel) return;
  double* su = mf_smem + (size_t)wid * S::WARP_DOUBLES;
  double* bufA = su + 3 * S::SC;
  double* bufB = bufA + 9 * S::SC;
  // gather, constrained dofs masked out (columns of K exist for free dofs only)
  for (int t = lane; t < S::NN; t += 32) {
    const int64_t node = nodes_t[e * S::NN + t];
    const int i = t % N, j = (t / N) % N, k = t / (N * N);
    const int off = i + S::S1 * j + S::S2 * k;
#pragma unroll
    for (int c = 0; c < 3; ++c) {
      const double xv = (xg && node >= hw.n_own) ? __ldcg(xg + (node - hw.n_own) * 3 + c) : x[node * 3 + c];
      su[c * S::SC + off] = cls[node * 3 + c] == 0 ? xv : 0.0;
    }
  }
  __syncwarp();
  double yacc[S::SLOTS][N];
#pragma unroll
  for (int s = 0; s < S::SLOTS; ++s)
#pragma unroll
    for (int i = 0; i < N; ++i) yacc[s][i] = 0.0;
  const double* ge = geo + e * 10 * npts;
  mf_rule<N, N, true>(su, bufA, bufB, ge, npts, yacc, lane);
  mf_rule<N, N - 1, false>(su, bufA, bufB, ge + S::NN, npts, yacc, lane);
#pragma unroll
  for (int s = 0; s < S::SLOTS; ++s) {
    const int l = lane + 32 * s;
    if (l < 3 * N * N) {
      const int j = l % N, k = (l / N) % N, c = l / (N * N);
#pragma unroll
      for (int i = 0; i < N; ++i) {
        const int a = loc_of_t[i + N * (j + N * k)];
        ye[(e * S::NN + a) * 3 + c] = yacc[s][i];
      }
    }
  }
}

// ---- node kernel: y[r] = alpha * (free ? sum of the element results : x[r]) + beta z[r] --------
template <int NDOT>
__global__ void __launch_bounds__(256)
mf_gather_kernel(int64_t nrows, const int64_t* __restrict__ n2e_ptr, const int2* __restrict__ n2e,
                 const double* __restrict__ ye, int nn, const uint8_t* __restrict__ cls,
                 const double* __restrict__ x, double alpha, double beta, const double* __restrict__ z,
                 double* __restrict__ y, SpmvFuse fuse) {
  constexpr int NV = NDOT > 0 ? NDOT : 1;
  __shared__ double red_sm[NV * 32];
  if (fuse.skip && *fuse.skip != 0.0) return;
  double dotacc[NV];
#pragma unroll
  for (int k = 0; k < NV; ++k) dotacc[k] = 0.0;
  for (int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; r < nrows;
       r += (int64_t)gridDim.x * blockDim.x) {
    double acc[3] = {0.0, 0.0, 0.0};
    const int64_t k1 = n2e_ptr[r + 1];
    for (int64_t k = n2e_ptr[r]; k < k1; ++k) {
      const int2 ea = n2e[k];
      const double* src = ye + ((int64_t)ea.x * nn + ea.y) * 3;
      acc[0] += src[0];
      acc[1] += src[1];
      acc[2] += src[2];
    }
#pragma unroll
    for (int c = 0; c < 3; ++c) {
      const int64_t dof = r * 3 + c;
      const double xv = x[dof];
      double out = alpha * (cls[dof] == 0 ? acc[c] : xv);
      if (beta != 0.0) out += beta * z[dof];
      y[dof] = out;
      if (NDOT > 0) {
        dotacc[0] += out * xv;
        if (NDOT == 3) {
          dotacc[NV > 1 ? 1 : 0] += fuse.dot_r[dof] * xv;
          dotacc[NV > 2 ? 2 : 0] += xv * xv;
        }
      }
    }
  }
  if (NDOT > 0) {
    block_reduce<NV>(dotacc, red_sm);
    if (fuse.red.nranks > 1)
      grid_reduce_push<NV>(dotacc, fuse.partials, fuse.ticket, fuse.red, 0);
    else
      grid_reduce_finish<NV>(dotacc, fuse.partials, fuse.ticket, fuse.dot_out);
  }
}

template <int N, int WPC>
static int mf_launch_apply(pyn_matfree* mf, const double* x, const double* skip, const HaloWait& hw) {
  pyn_mesh* mesh = mf->mesh;
  pyn_ctx* ctx = mesh->ctx;
  const size_t smem = (size_t)WPC * MfShape<N>::WARP_DOUBLES * sizeof(double);
  PYN_CUDA(cudaFuncSetAttribute(mf_apply_kernel<N, WPC>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  const unsigned grid = (unsigned)((mesh->nel + WPC - 1) / WPC);
  mf_apply_kernel<N, WPC><<<grid, WPC * 32, smem, ctx->stream>>>(
      mesh->nel, mf->nodes_t.p, mf->loc_of_t.p, mesh->dofcls.p, mf->geo.p, mf->npts, x, mf->ye.p, skip, hw);
  PYN_LAUNCHED(ctx);
  return PYN_OK;
}

int matfree_apply(pyn_mat* m, int64_t nrows, double alpha, const double* x, double beta,
                  const double* z, double* y, bool dot, const SpmvFuse& fuse) {
  pyn_matfree* mf = m->mf;
  pyn_mesh* mesh = mf->mesh;
  pyn_ctx* ctx = mesh->ctx;
  PYN_CHECK(mf_upload_tables(mesh));
  prof_begin(ctx, dot ? PROF_SPMV_KSP : PROF_SPMV);
  switch (mesh->ngl) {
    case 2: PYN_CHECK((mf_launch_apply<2, 4>(mf, x, fuse.skip, fuse.hw))); break;
    case 3: PYN_CHECK((mf_launch_apply<3, 4>(mf, x, fuse.skip, fuse.hw))); break;
    case 4: PYN_CHECK((mf_launch_apply<4, 4>(mf, x, fuse.skip, fuse.hw))); break;
    case 5: PYN_CHECK((mf_launch_apply<5, 4>(mf, x, fuse.skip, fuse.hw))); break;
    case 6: PYN_CHECK((mf_launch_apply<6, 2>(mf, x, fuse.skip, fuse.hw))); break;
    default:
      set_error("matrix-free K: ngl %d not supported", mesh->ngl);
      return PYN_E_UNSUPPORTED;
  }
  int grid = grid_for(ctx, nrows, 256, 8);
  if (grid > kRedMaxBlocks) grid = kRedMaxBlocks;
#define MF_GATHER(NDOT)                                                                                   \
  mf_gather_kernel<NDOT><<<grid, 256, 0, ctx->stream>>>(nrows, mesh->n2e_ptr.p, mesh->n2e.p, mf->ye.p,   \
                                                        mesh->nn, mesh->dofcls.p, x, alpha, beta, z, y, fuse)
  if (dot && fuse.dot_r) MF_GATHER(3);
  else if (dot) MF_GATHER(1);
  else MF_GATHER(0);
#undef MF_GATHER
  prof_end(ctx);
  PYN_LAUNCHED(ctx);
  return PYN_OK;
}

}  // namespace pyn
