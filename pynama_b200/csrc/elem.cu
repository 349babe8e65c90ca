// elem.cu -- batched per-element quadrature kernels (k1, k2, k3).
//
// k1+k2: K_e, Rw_e, Rd_e of Spectral.getElemKLEMatrices (spectral.py:92-160).
//   One CTA per (element, tile of row nodes).  Per chunk of quadrature points
//   the CTA stages J^-1, w|J| and the physical gradients G[g][k][b] of all
//   element nodes in shared memory (rebuilt from the 1-D tables), then every
//   thread accumulates its node pairs (a,b) in registers: 21 fp64
//   accumulators per pair in 3-D (K 3x3, Rw 3x3, Rd 3).
// k3: SrT_e, DivSrT_e, Curl_e, W_e of Spectral.getElemKLEOperators
//   (spectral.py:162-228).  The nodal rule is collocated (H[g][a] = delta_ga,
//   exact in the reference's tables), so row node a only sees point g = a and
//   only the nodes b on a's grid lines give non-zero blocks: the kernel
//   evaluates just those ("stencil" pairs); every other entry of the
//   reference's dense blocks is an exact 0.0.
//
// Element block store (EBS) layout: block-major, [element][pair][i][j], so the
// assembly gather reads whole node blocks contiguously.
#include "elem.cuh"
#include "elem_math.cuh"

namespace pyn {

template <int DIM, int P>
__global__ void __launch_bounds__(256)
elem_kle_kernel(const ElemTables* __restrict__ tab, const double* __restrict__ corners,
                int64_t e0, int natiles, int TA, int GC, double* __restrict__ ebsK,
                double* __restrict__ ebsRw, double* __restrict__ ebsRd) {
  constexpr int NC = 1 << DIM;
  constexpr int DW = DIM == 2 ? 1 : 3;
  extern __shared__ double smem[];
  const int nn = tab->nn;
  const int64_t el = blockIdx.x / natiles;          // local element in this batch
  const int atile = blockIdx.x % natiles;
  const int a0 = atile * TA;
  const int ta = min(TA, nn - a0);
  const int tid = threadIdx.x, nth = blockDim.x;

  double* sX = smem;                                 // [NC*DIM]
  double* sJinv = sX + NC * DIM;                     // [GC*DIM*DIM]
  double* sWd = sJinv + GC * DIM * DIM;              // [GC]
  double* sG = sWd + GC;                             // [GC*DIM*nn]
  double* sH = sG + (size_t)GC * DIM * nn;           // [GC*nn]

  if (tid < NC * DIM) sX[tid] = corners[(e0 + el) * NC * DIM + tid];

  PairAcc<DIM> acc[P];
#pragma unroll
  for (int m = 0; m < P; ++m) acc[m].clear();

  const int npairs = ta * nn;
  for (int q = 0; q < 2; ++q) {                      // 0: full rule, 1: reduced rule
    const Rule1D& rule = tab->rule[q == 0 ? RULE_FULL : RULE_RED];
    const int npts = rule.npts;
    for (int g0 = 0; g0 < npts; g0 += GC) {
      const int gc_n = min(GC, npts - g0);
      __syncthreads();                               // previous chunk fully consumed (and sX visible)
      for (int gc = tid; gc < gc_n; gc += nth) {
        double det;
        point_geometry<DIM>(rule, tab->corn_t, g0 + gc, sX, sJinv + gc * DIM * DIM, &det);
        sWd[gc] = point_weight<DIM>(rule, g0 + gc) * det;
      }
      __syncthreads();
      for (int i = tid; i < gc_n * nn; i += nth) {
        const int gc = i / nn, b = i - gc * nn;
        double G[DIM], H;
        node_gradient<DIM>(rule, tab->node_t[b], g0 + gc, sJinv + gc * DIM * DIM, G, &H);
#pragma unroll
        for (int k = 0; k < DIM; ++k) sG[((size_t)gc * DIM + k) * nn + b] = G[k];
        sH[gc * nn + b] = H;
      }
      __syncthreads();
#pragma unroll
      for (int m = 0; m < P; ++m) {
        const int p = tid + m * nth;
        if (p < npairs) {
          const int a = a0 + p / nn, b = p % nn;
          for (int gc = 0; gc < gc_n; ++gc) {
            double ga[DIM], gb[DIM];
#pragma unroll
            for (int k = 0; k < DIM; ++k) {
              ga[k] = sG[((size_t)gc * DIM + k) * nn + a];
              gb[k] = sG[((size_t)gc * DIM + k) * nn + b];
            }
            const double wd = sWd[gc];
            if (q == 0)
              acc[m].add_full(wd, ga, sH[gc * nn + a], gb);
            else
              acc[m].add_red(wd, ga, gb, sH[gc * nn + b]);
          }
        }
      }
    }
  }
#pragma unroll
  for (int m = 0; m < P; ++m) {
    const int p = tid + m * nth;
    if (p < npairs) {
      const int a = a0 + p / nn, b = p % nn;
      const size_t blk = ((size_t)el * nn + a) * nn + b;
#pragma unroll
      for (int c = 0; c < DIM; ++c) {
#pragma unroll
        for (int e = 0; e < DIM; ++e) ebsK[blk * DIM * DIM + c * DIM + e] = acc[m].K[c][e];
#pragma unroll
        for (int e = 0; e < DW; ++e) ebsRw[blk * DIM * DW + c * DW + e] = acc[m].Rw[c][e];
        ebsRd[blk * DIM + c] = acc[m].Rd[c];
      }
    }
  }
}

template <int DIM>
__global__ void __launch_bounds__(128)
elem_ops_kernel(const ElemTables* __restrict__ tab, const double* __restrict__ corners,
                const int32_t* __restrict__ pair_b, int ppa, int64_t e0,
                double* __restrict__ ebsSrT, double* __restrict__ ebsDiv,
                double* __restrict__ ebsCurl, double* __restrict__ ebsW) {
  constexpr int NC = 1 << DIM;
  constexpr int DW = DIM == 2 ? 1 : 3;
  constexpr int DS = DIM == 2 ? 3 : 6;
  extern __shared__ double smem[];
  const int nn = tab->nn;
  const int64_t el = blockIdx.x;
  const int tid = threadIdx.x, nth = blockDim.x;
  double* sX = smem;                       // [NC*DIM]
  double* sJinv = sX + NC * DIM;           // [nn*DIM*DIM]
  double* sWd = sJinv + nn * DIM * DIM;    // [nn]
  const Rule1D& rule = tab->rule[RULE_NODAL];

  if (tid < NC * DIM) sX[tid] = corners[(e0 + el) * NC * DIM + tid];
  __syncthreads();
  for (int g = tid; g < nn; g += nth) {
    double det;
    point_geometry<DIM>(rule, tab->corn_t, g, sX, sJinv + g * DIM * DIM, &det);
    const double wd = point_weight<DIM>(rule, g) * det;
    sWd[g] = wd;
    // W_e[a] = sum_b sum_g wd H[g][a] H[g][b] = wd[a]  (collocation; spectral.py:223-225)
    ebsW[el * nn + g] = wd;
  }
  __syncthreads();
  const int npairs = nn * ppa;
  for (int q = tid; q < npairs; q += nth) {
    const int a = q / ppa;
    const int b = pair_b[q];
    double G[DIM], Hb;
    node_gradient<DIM>(rule, tab->node_t[b], a, sJinv + a * DIM * DIM, G, &Hb);
    // H[g=a][a] from the tables (== 1.0 exactly)
    const double ha = tensor_val<DIM, kMaxNgl>(rule.h, rule.dh, rule.gp_t[a], tab->node_t[a], -1);
    OpsAcc<DIM> acc;
    acc.clear();
    acc.add(sWd[a], ha, G);
    const size_t blk = (size_t)el * npairs + q;
#pragma unroll
    for (int s = 0; s < DS; ++s)
#pragma unroll
      for (int c = 0; c < DIM; ++c) {
        ebsSrT[blk * DS * DIM + s * DIM + c] = acc.SrT[s][c];
        ebsDiv[blk * DS * DIM + c * DS + s] = acc.Div[c][s];
      }
#pragma unroll
    for (int r = 0; r < DW; ++r)
#pragma unroll
      for (int c = 0; c < DIM; ++c) ebsCurl[blk * DW * DIM + r * DIM + c] = acc.Curl[r][c];
  }
}

// ---- launchers ----------------------------------------------------------------
static void kle_launch_shape(const pyn_mesh* mesh, int* TA, int* GC, int* natiles, size_t* smem) {
  const int nn = mesh->nn, dim = mesh->dim;
  const int P = 2, nth = 256;
  int ta = (P * nth) / nn;
  if (ta < 1) ta = 1;
  if (ta > nn) ta = nn;
  const int npts_max = mesh->htab.rule[RULE_FULL].npts;
  const size_t per_pt = (size_t)(dim + 1) * nn * 8 + (dim * dim + 1) * 8;
  int gc = (int)((size_t)(64 * 1024) / per_pt);
  if (gc < 1) gc = 1;
  if (gc > npts_max) gc = npts_max;
  *TA = ta;
  *GC = gc;
  *natiles = (nn + ta - 1) / ta;
  *smem = (size_t)(1 << dim) * dim * 8 + gc * per_pt;
}

int launch_elem_kle(pyn_mesh* mesh, int64_t e0, int64_t ne, double* ebsK, double* ebsRw,
                    double* ebsRd) {
  pyn_ctx* ctx = mesh->ctx;
  if (ne <= 0) return PYN_OK;
  int TA, GC, natiles;
  size_t smem;
  kle_launch_shape(mesh, &TA, &GC, &natiles, &smem);
  PYN_REQUIRE(ne * natiles < (int64_t)2147483647, "element batch too large for one launch");
  dim3 grid((unsigned)(ne * natiles)), block(256);
  prof_begin(ctx, PROF_ELEM_KLE);
  if (mesh->dim == 2) {
    PYN_CUDA(cudaFuncSetAttribute(elem_kle_kernel<2, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    elem_kle_kernel<2, 2><<<grid, block, smem, ctx->stream>>>(mesh->dtab.p, mesh->corners.p, e0, natiles,
                                                              TA, GC, ebsK, ebsRw, ebsRd);
  } else {
    PYN_CUDA(cudaFuncSetAttribute(elem_kle_kernel<3, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    elem_kle_kernel<3, 2><<<grid, block, smem, ctx->stream>>>(mesh->dtab.p, mesh->corners.p, e0, natiles,
                                                              TA, GC, ebsK, ebsRw, ebsRd);
  }
  prof_end(ctx);
  PYN_LAUNCHED(ctx);
  return PYN_OK;
}

int launch_elem_ops(pyn_mesh* mesh, int64_t e0, int64_t ne, double* ebsSrT, double* ebsDiv,
                    double* ebsCurl, double* ebsW) {
  pyn_ctx* ctx = mesh->ctx;
  if (ne <= 0) return PYN_OK;
  PYN_REQUIRE(mesh->has_stencil, "stencil graph not built");
  const int dim = mesh->dim, nn = mesh->nn;
  size_t smem = ((size_t)(1 << dim) * dim + (size_t)nn * (dim * dim + 1)) * 8;
  dim3 grid((unsigned)ne), block(128);
  if (dim == 2) {
    PYN_CUDA(cudaFuncSetAttribute(elem_ops_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    elem_ops_kernel<2><<<grid, block, smem, ctx->stream>>>(mesh->dtab.p, mesh->corners.p,
                                                           mesh->stencil.pair_b.p, mesh->stencil.ppa, e0,
                                                           ebsSrT, ebsDiv, ebsCurl, ebsW);
  } else {
    PYN_CUDA(cudaFuncSetAttribute(elem_ops_kernel<3>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    elem_ops_kernel<3><<<grid, block, smem, ctx->stream>>>(mesh->dtab.p, mesh->corners.p,
                                                           mesh->stencil.pair_b.p, mesh->stencil.ppa, e0,
                                                           ebsSrT, ebsDiv, ebsCurl, ebsW);
  }
  PYN_LAUNCHED(ctx);
  return PYN_OK;
}

}  // namespace pyn
