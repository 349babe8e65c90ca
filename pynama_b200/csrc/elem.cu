// elem.cu -- batched per-element quadrature kernels (k1, k2, k3).
//
// k1+k2: K_e, Rw_e, Rd_e of Spectral.getElemKLEMatrices (spectral.py:92-160).
//   One CTA per (element, tile of row nodes).  Per chunk of quadrature points
//   the CTA stages J^-1, w|J| and the physical gradients of all element nodes
//   in shared memory (rebuilt from the 1-D tables), then every thread
//   accumulates the 16 moments (3-D) of its AT x 1 node pairs in registers
//   and expands them into K (3x3), Rw (3x3), Rd (3) once, at the end.
// k3: SrT_e, DivSrT_e, Curl_e, W_e of Spectral.getElemKLEOperators
//   (spectral.py:162-228).  The nodal rule is collocated (H[g][a] = delta_ga,
//   exact in the reference's tables), so row node a only sees point g = a and
//   only the nodes b on a's grid lines give non-zero blocks: the kernel
//   evaluates just those ("stencil" pairs); every other entry of the
//   reference's dense blocks is an exact 0.0.
//
// Element block store (EBS) layout of k1+k2: one record per (element, node pair),
// [K (dim x dim) | pad] [Rw (dim x dim_w) | Rd (dim) | pad], 12 + 12 doubles in 3-D, 4 + 4 in 2-D
// (KleRecord, elem.cuh): whole 32-byte sectors, so the assembly gather reads a pair's
// three blocks with 256-bit loads.  k3 stores block-major, [element][pair][i][j].
#include "elem.cuh"
#include "elem_math.cuh"

namespace pyn {

// 1-D tables of the full and the reduced rule, copied to shared memory once per CTA; tensor
// indices of nodes / points / corners packed into one word (t0 | t1<<8 | t2<<16).
struct KleSmemTab {
  double h[2][kMaxPts1d][kMaxNgl];
  double dh[2][kMaxPts1d][kMaxNgl];
  double hc[2][kMaxPts1d][2];
  double dhc[2][kMaxPts1d][2];
  double w[2][kMaxPts1d];
};
__device__ __forceinline__ uint32_t pack_t(const uint8_t* t) {
  return (uint32_t)t[0] | ((uint32_t)t[1] << 8) | ((uint32_t)t[2] << 16);
}
// value (which < 0) or derivative along `which` of a tensor-product basis function; factors
// multiplied in direction order as tensor_val does
template <int DIM>
__device__ __forceinline__ double tensor3(const double* v, const double* dv, int which) {
  double r = which == 0 ? dv[0] : v[0];
#pragma unroll
  for (int d = 1; d < DIM; ++d) r = r * (which == d ? dv[d] : v[d]);
  return r;
}
// J^-1 and det J at a point with packed tensor index gt (same arithmetic as point_geometry)
template <int DIM>
__device__ __forceinline__ void point_geometry_s(const KleSmemTab& T, int q, uint32_t gt,
                                                 const uint32_t* cornT, const double* X,
                                                 double* jinv, double* det) {
  constexpr int NC = 1 << DIM;
  double J[DIM][DIM];
#pragma unroll
  for (int d = 0; d < DIM; ++d)
#pragma unroll
    for (int e = 0; e < DIM; ++e) J[d][e] = 0.0;
#pragma unroll
  for (int c = 0; c < NC; ++c) {
    const uint32_t ct = cornT[c];
    double v[DIM], dv[DIM];
#pragma unroll
    for (int d = 0; d < DIM; ++d) {
      const int gi = (gt >> (8 * d)) & 255, ci = (ct >> (8 * d)) & 255;
      v[d] = T.hc[q][gi][ci];
      dv[d] = T.dhc[q][gi][ci];
    }
#pragma unroll
    for (int d = 0; d < DIM; ++d) {
      const double hv = tensor3<DIM>(v, dv, d);
#pragma unroll
      for (int e = 0; e < DIM; ++e) J[d][e] += hv * X[c * DIM + e];
    }
  }
  if (DIM == 2) {
    double dt = J[0][0] * J[1][1] - J[0][1] * J[1][0];
    double inv = 1.0 / dt;
    jinv[0] = J[1][1] * inv;
    jinv[1] = -J[0][1] * inv;
    jinv[2] = -J[1][0] * inv;
    jinv[3] = J[0][0] * inv;
    *det = dt;
  } else {
    double c00 = J[1][1] * J[2][2] - J[1][2] * J[2][1];
    double c01 = J[1][2] * J[2][0] - J[1][0] * J[2][2];
    double c02 = J[1][0] * J[2][1] - J[1][1] * J[2][0];
    double dt = J[0][0] * c00 + J[0][1] * c01 + J[0][2] * c02;
    double inv = 1.0 / dt;
    jinv[0] = c00 * inv;
    jinv[1] = (J[0][2] * J[2][1] - J[0][1] * J[2][2]) * inv;
    jinv[2] = (J[0][1] * J[1][2] - J[0][2] * J[1][1]) * inv;
    jinv[3] = c01 * inv;
    jinv[4] = (J[0][0] * J[2][2] - J[0][2] * J[2][0]) * inv;
    jinv[5] = (J[0][2] * J[1][0] - J[0][0] * J[1][2]) * inv;
    jinv[6] = c02 * inv;
    jinv[7] = (J[0][1] * J[2][0] - J[0][0] * J[2][1]) * inv;
    jinv[8] = (J[0][0] * J[1][1] - J[0][1] * J[1][0]) * inv;
    *det = dt;
  }
}

// Moments of one node pair (a,b) over the quadrature points; every entry of K_e, Rw_e and
// Rd_e of the pair is a fixed linear combination of them (C(g) = the curl matrix of g):
//   L   = sum_full wd (ga . gb)            M[k][l] = sum_red wd ga[k] gb[l]
//   N[k] = sum_full wd ha gb[k]            Q[k]    = sum_red wd hb ga[k]
//   K[c][e]  = delta_ce (L + aw tr M) + ad M[c][e] - aw M[e][c]     (C(ga)^T C(gb) = (ga.gb) I - gb ga^T)
//   Rw[r][c] = curl_w(r,c,N) + aw curl_v(c,r,Q)        Rd[c] = -N[c] + ad Q[c]
// 16 accumulators and 7 / 15 FMA-pipe instructions per full / reduced point in 3-D, against 21
// accumulators and ~25 / ~50 for the entry-by-entry form (PairAcc, kept for the host check).
template <int DIM>
struct PairMoments {
  double L, M[DIM][DIM], N[DIM], Q[DIM];
  __device__ __forceinline__ void clear() {
    L = 0.0;
#pragma unroll
    for (int k = 0; k < DIM; ++k) {
      N[k] = 0.0;
      Q[k] = 0.0;
#pragma unroll
      for (int l = 0; l < DIM; ++l) M[k][l] = 0.0;
    }
  }
};

// scr holds one BB-double block per lane (lane-major); write them to dst[blk*BB ..) so that
// consecutive lanes of a store instruction touch consecutive doubles.  All 32 lanes call.
// (lane l's block starts at scr + l*LD; LD = BB+1 keeps the lanes' writes off a common bank)
template <int BB, int LD>
__device__ __forceinline__ void warp_store_blocks(const double* scr, double* __restrict__ dst,
                                                  long long blk, bool valid, int lane) {
  __syncwarp();
#pragma unroll
  for (int j = 0; j < BB; ++j) {
    const int idx = j * 32 + lane;
    const int src = idx / BB, q = idx - src * BB;
    const long long bl = __shfl_sync(0xffffffffu, blk, src);
    const int ok = __shfl_sync(0xffffffffu, (int)valid, src);
    if (ok) __stcs(dst + bl * BB + q, scr[src * LD + q]);
  }
  __syncwarp();
}

// One CTA per (element, tile of row nodes).  Thread = AT row nodes x 1 column node b, so the
// column-side gradient is loaded once per point and reused AT times from registers; the
// row-side data is stored interleaved ([node][g0,g1,g2,h]) and read with two broadcast
// 16-byte loads, the column side planar ([k][node]) and read conflict-free.
template <int DIM, int AT, int NTH>
__global__ void __launch_bounds__(NTH, 1)
elem_kle_kernel(const ElemTables* __restrict__ tab, const double* __restrict__ corners,
                int64_t e0, const int32_t* __restrict__ rep, int natiles, int ntile, int GC,
                double* __restrict__ ebs) {
  constexpr int NC = 1 << DIM;
  constexpr int DW = DIM == 2 ? 1 : 3;
  extern __shared__ __align__(16) double smem[];
  const int nn = tab->nn;
  const int64_t el = blockIdx.x / natiles;          // position in this batch (= index in the store)
  const int64_t esrc = rep ? (int64_t)rep[el] : e0 + el;   // element whose corners are used
  const int atile = blockIdx.x % natiles;
  const int tid = threadIdx.x;

  double* sA = smem;                                 // [GC][nn][4]  interleaved g0,g1,g2,h
  double* sB = sA + (size_t)GC * nn * 4;             // [GC][4][nn]  planar
  double* sJinv = sB + (size_t)GC * nn * 4;          // [GC*DIM*DIM]
  double* sWd = sJinv + GC * DIM * DIM;              // [GC]
  double* sX = sWd + GC;                             // [NC*DIM]
  KleSmemTab& T = *reinterpret_cast<KleSmemTab*>(sX + NC * DIM);
  uint32_t* sNodeT = reinterpret_cast<uint32_t*>(&T + 1);   // [nn]
  uint32_t* sCornT = sNodeT + nn;                            // [8]
  uint32_t* sGpT = sCornT + 8;                               // [npts(full)] then [npts(red)]
  const int npts_full = tab->rule[RULE_FULL].npts;

  if (tid < NC * DIM) sX[tid] = corners[esrc * NC * DIM + tid];
  for (int i = tid; i < 2 * kMaxPts1d * kMaxNgl; i += NTH) {
    const int q = i / (kMaxPts1d * kMaxNgl), r = i % (kMaxPts1d * kMaxNgl);
    const Rule1D& rl = tab->rule[q == 0 ? RULE_FULL : RULE_RED];
    (&T.h[q][0][0])[r] = (&rl.h[0][0])[r];
    (&T.dh[q][0][0])[r] = (&rl.dh[0][0])[r];
    if (r < kMaxPts1d * 2) {
      (&T.hc[q][0][0])[r] = (&rl.hc[0][0])[r];
      (&T.dhc[q][0][0])[r] = (&rl.dhc[0][0])[r];
    }
    if (r < kMaxPts1d) T.w[q][r] = rl.w[r];
  }
  for (int i = tid; i < nn; i += NTH) sNodeT[i] = pack_t(tab->node_t[i]);
  if (tid < NC) sCornT[tid] = pack_t(tab->corn_t[tid]);
  for (int i = tid; i < npts_full; i += NTH) sGpT[i] = pack_t(tab->rule[RULE_FULL].gp_t[i]);
  for (int i = tid; i < tab->rule[RULE_RED].npts; i += NTH)
    sGpT[npts_full + i] = pack_t(tab->rule[RULE_RED].gp_t[i]);

  // this thread's pairs: rows a_first .. a_first+AT-1 (clipped), column b
  const int it = tid / nn;
  const int b = tid - it * nn;
  const int a_first = (atile * ntile + it) * AT;
  const bool active = it < ntile && a_first < nn;
  int arow[AT];
#pragma unroll
  for (int t = 0; t < AT; ++t) arow[t] = min(a_first + t, nn - 1);

  const unsigned nn_rcp = (unsigned)((0x100000000ull + (unsigned)nn - 1) / (unsigned)nn);  // i / nn = umulhi(i, nn_rcp), i < 2^32/nn

  PairMoments<DIM> acc[AT];
#pragma unroll
  for (int t = 0; t < AT; ++t) acc[t].clear();

  for (int q = 0; q < 2; ++q) {                      // 0: full rule, 1: reduced rule
    const int npts = q == 0 ? npts_full : tab->rule[RULE_RED].npts;
    const uint32_t* gpT = sGpT + (q == 0 ? 0 : npts_full);
    for (int g0 = 0; g0 < npts; g0 += GC) {
      const int gc_n = min(GC, npts - g0);
      __syncthreads();                               // previous chunk fully consumed (and tables visible)
      for (int gc = tid; gc < gc_n; gc += NTH) {
        const uint32_t gt = gpT[g0 + gc];
        double det;
        point_geometry_s<DIM>(T, q, gt, sCornT, sX, sJinv + gc * DIM * DIM, &det);
        double w = T.w[q][gt & 255];
#pragma unroll
        for (int d = 1; d < DIM; ++d) w = w * T.w[q][(gt >> (8 * d)) & 255];
        sWd[gc] = w * det;
      }
      __syncthreads();
      {
        // unit i -> (point gc, node nb) = (i / nn, i % nn); SU independent units per trip so the
        // table / J^-1 loads of one overlap the arithmetic of another (2 warps per scheduler only)
        constexpr int SU = 4;
        const int units = gc_n * nn;
        for (int i0 = tid; i0 < units; i0 += SU * NTH) {
          double G[SU][4];
          int gcs[SU], nbs[SU];
#pragma unroll
          for (int u = 0; u < SU; ++u) {
            const int i = min(i0 + u * NTH, units - 1);
            const int gc = (int)__umulhi((unsigned)i, nn_rcp), nb = i - gc * nn;
            gcs[u] = gc;
            nbs[u] = nb;
            const uint32_t gt = gpT[g0 + gc], nt = sNodeT[nb];
            double v[DIM], dv[DIM];
#pragma unroll
            for (int d = 0; d < DIM; ++d) {
              const int gi = (gt >> (8 * d)) & 255, ni = (nt >> (8 * d)) & 255;
              v[d] = T.h[q][gi][ni];
              dv[d] = T.dh[q][gi][ni];
            }
            double hr[DIM];
#pragma unroll
            for (int d = 0; d < DIM; ++d) hr[d] = tensor3<DIM>(v, dv, d);
            const double* ji = sJinv + gc * DIM * DIM;
#pragma unroll
            for (int k = 0; k < DIM; ++k) {
              double sum = ji[k * DIM] * hr[0];
#pragma unroll
              for (int d = 1; d < DIM; ++d) sum += ji[k * DIM + d] * hr[d];
              G[u][k] = sum;
            }
            if (DIM == 2) G[u][2] = 0.0;
            G[u][3] = tensor3<DIM>(v, dv, -1);
          }
#pragma unroll
          for (int u = 0; u < SU; ++u) {
            if (i0 + u * NTH < units) {
              double2* dst = reinterpret_cast<double2*>(sA + ((size_t)gcs[u] * nn + nbs[u]) * 4);
              dst[0] = make_double2(G[u][0], G[u][1]);
              dst[1] = make_double2(G[u][2], G[u][3]);
#pragma unroll
              for (int k = 0; k < 4; ++k)
                if (k < DIM || k == 3) sB[((size_t)gcs[u] * 4 + k) * nn + nbs[u]] = G[u][k];
            }
          }
        }
      }
      __syncthreads();
      if (active) {
        if (q == 0) {
#pragma unroll 2
          for (int gc = 0; gc < gc_n; ++gc) {
            const double wd = sWd[gc];
            double gb[DIM];
#pragma unroll
            for (int k = 0; k < DIM; ++k) gb[k] = sB[((size_t)gc * 4 + k) * nn + b];
#pragma unroll
            for (int t = 0; t < AT; ++t) {
              const double2* pa = reinterpret_cast<const double2*>(sA + ((size_t)gc * nn + arow[t]) * 4);
              const double2 g01 = pa[0], g2h = pa[1];
              double lap = g01.x * gb[0];
              lap = fma(g01.y, gb[1], lap);
              if (DIM == 3) lap = fma(g2h.x, gb[2], lap);
              acc[t].L = fma(wd, lap, acc[t].L);
              const double wha = wd * g2h.y;
#pragma unroll
              for (int k = 0; k < DIM; ++k) acc[t].N[k] = fma(wha, gb[k], acc[t].N[k]);
            }
          }
        } else {
#pragma unroll 2
          for (int gc = 0; gc < gc_n; ++gc) {
            const double wd = sWd[gc];
            double gb[DIM];
#pragma unroll
            for (int k = 0; k < DIM; ++k) gb[k] = sB[((size_t)gc * 4 + k) * nn + b];
            const double whb = wd * sB[((size_t)gc * 4 + 3) * nn + b];
#pragma unroll
            for (int t = 0; t < AT; ++t) {
              const double2* pa = reinterpret_cast<const double2*>(sA + ((size_t)gc * nn + arow[t]) * 4);
              const double2 g01 = pa[0];
              double ga[3];
              ga[0] = g01.x;
              ga[1] = g01.y;
              if (DIM == 3) ga[2] = pa[1].x;
#pragma unroll
              for (int k = 0; k < DIM; ++k) {
                const double wga = wd * ga[k];
#pragma unroll
                for (int l = 0; l < DIM; ++l) acc[t].M[k][l] = fma(wga, gb[l], acc[t].M[k][l]);
                acc[t].Q[k] = fma(whb, ga[k], acc[t].Q[k]);
              }
            }
          }
        }
      }
    }
  }
  // Epilogue: expand the moments into the pair's record [K | pad | Rw | Rd | pad] (KleRecord) and
  // store record-major.  A lane's record is RS consecutive doubles, so the warp transposes
  // through a private shared-memory tile and writes runs of 32 consecutive doubles (the lanes'
  // records are adjacent when their b are).
  __syncthreads();                                   // staging buffers are free now
  constexpr int RS = KleRecord<DIM>::RS, OW = KleRecord<DIM>::OFF_RW, OD = KleRecord<DIM>::OFF_RD;
  const int lane = tid & 31;
  double* scr = smem + (size_t)(tid >> 5) * 32 * (RS + 1);
#pragma unroll
  for (int t = 0; t < AT; ++t) {
    const int a = a_first + t;
    const bool valid = active && a < nn;
    const long long blk = ((long long)el * nn + a) * nn + b;
    const PairMoments<DIM>& m = acc[t];
    double tr = m.M[0][0];
#pragma unroll
    for (int k = 1; k < DIM; ++k) tr += m.M[k][k];
    const double dg = m.L + kAlphaW * tr;
    double* rec = scr + lane * (RS + 1);
#pragma unroll
    for (int c = 0; c < DIM; ++c)
#pragma unroll
      for (int e = 0; e < DIM; ++e) {
        double v = kAlphaD * m.M[c][e] - kAlphaW * m.M[e][c];
        if (c == e) v += dg;
        rec[c * DIM + e] = v;
      }
#pragma unroll
    for (int c = 0; c < DIM; ++c)
#pragma unroll
      for (int e = 0; e < DW; ++e) {
        double v = 0.0;
        if (DIM == 2 || c != e) v = curl_w<DIM>(c, e, m.N) + kAlphaW * curl_v<DIM>(e, c, m.Q);
        rec[OW + c * DW + e] = v;
      }
#pragma unroll
    for (int c = 0; c < DIM; ++c) rec[OD + c] = kAlphaD * m.Q[c] - m.N[c];
#pragma unroll
    for (int c = DIM * DIM; c < OW; ++c) rec[c] = 0.0;
#pragma unroll
    for (int c = OD + DIM; c < RS; ++c) rec[c] = 0.0;
    warp_store_blocks<RS, RS + 1>(scr, ebs, blk, valid, lane);
  }
}

template <int DIM>
__global__ void __launch_bounds__(128)
elem_ops_kernel(const ElemTables* __restrict__ tab, const double* __restrict__ corners,
                const int32_t* __restrict__ pair_b, int ppa, int64_t e0,
                const int32_t* __restrict__ rep, double* __restrict__ ebsSrT, double* __restrict__ ebsDiv,
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

  const int64_t esrc = rep ? (int64_t)rep[el] : e0 + el;
  if (tid < NC * DIM) sX[tid] = corners[esrc * NC * DIM + tid];
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
struct KleShape {
  int nth, at, ntile, natiles, gc;
  size_t smem;
};
static KleShape kle_launch_shape(const pyn_mesh* mesh) {
  const int nn = mesh->nn, dim = mesh->dim;
  KleShape s;
  s.nth = nn <= 256 ? 256 : 512;           // nn <= 512 (kMaxNN)
  s.at = nn <= 256 ? 4 : 2;
  s.ntile = s.nth / nn;
  s.natiles = (nn + s.ntile * s.at - 1) / (s.ntile * s.at);
  const int npts_max = mesh->htab.rule[RULE_FULL].npts;
  const size_t per_pt = (size_t)nn * 64 + (dim * dim + 1) * 8;
  int gc = (int)((size_t)(144 * 1024) / per_pt);
  if (gc < 1) gc = 1;
  if (gc > npts_max) gc = npts_max;
  const int nchunks = (npts_max + gc - 1) / gc;    // equal chunks: no short last chunk
  gc = (npts_max + nchunks - 1) / nchunks;
  s.gc = gc;
  s.smem = (size_t)(1 << dim) * dim * 8 + gc * per_pt + sizeof(KleSmemTab) +
           4 * ((size_t)nn + 8 + npts_max + mesh->htab.rule[RULE_RED].npts);
  const size_t scratch = (size_t)s.nth * (kle_record_stride(dim) + 1) * 8;      // epilogue transpose tiles
  if (s.smem < scratch) s.smem = scratch;
  return s;
}

int launch_elem_kle(pyn_mesh* mesh, int64_t e0, int64_t ne, const int32_t* rep, double* ebs) {
  pyn_ctx* ctx = mesh->ctx;
  if (ne <= 0) return PYN_OK;
  const KleShape s = kle_launch_shape(mesh);
  PYN_REQUIRE(ne * s.natiles < (int64_t)2147483647, "element batch too large for one launch");
  dim3 grid((unsigned)(ne * s.natiles));
  prof_begin(ctx, PROF_ELEM_KLE);
#define KLE_CALL(DIM, AT, NTH)                                                                       \
  do {                                                                                               \
    PYN_CUDA(cudaFuncSetAttribute(elem_kle_kernel<DIM, AT, NTH>,                                     \
                                  cudaFuncAttributeMaxDynamicSharedMemorySize, (int)s.smem));        \
    elem_kle_kernel<DIM, AT, NTH><<<grid, NTH, s.smem, ctx->stream>>>(                               \
        mesh->dtab.p, mesh->corners.p, e0, rep, s.natiles, s.ntile, s.gc, ebs);            \
  } while (0)
  if (mesh->dim == 2) {
    if (s.nth == 256) KLE_CALL(2, 4, 256); else KLE_CALL(2, 2, 512);
  } else {
    if (s.nth == 256) KLE_CALL(3, 4, 256); else KLE_CALL(3, 2, 512);
  }
#undef KLE_CALL
  prof_end(ctx);
  PYN_LAUNCHED(ctx);
  return PYN_OK;
}

int launch_elem_ops(pyn_mesh* mesh, int64_t e0, int64_t ne, const int32_t* rep, double* ebsSrT,
                    double* ebsDiv, double* ebsCurl, double* ebsW) {
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
                                                           rep, ebsSrT, ebsDiv, ebsCurl, ebsW);
  } else {
    PYN_CUDA(cudaFuncSetAttribute(elem_ops_kernel<3>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    elem_ops_kernel<3><<<grid, block, smem, ctx->stream>>>(mesh->dtab.p, mesh->corners.p,
                                                           mesh->stencil.pair_b.p, mesh->stencil.ppa, e0,
                                                           rep, ebsSrT, ebsDiv, ebsCurl, ebsW);
  }
  PYN_LAUNCHED(ctx);
  return PYN_OK;
}

}  // namespace pyn
