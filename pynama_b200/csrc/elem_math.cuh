// elem_math.cuh -- per-point / per-node-pair arithmetic of the element kernels.
//
// Host+device inline functions so the same code can be exercised on the CPU by
// the (test-only) host harness in tests/hostcheck.  Follows
// Spectral.getElemKLEMatrices (spectral.py:92-160) and
// Spectral.getElemKLEOperators (spectral.py:162-228) term by term; products are
// grouped like the reference's expressions ((w*det)*alpha)*(B^T B entry) so
// results agree to a few ulp.
#pragma once
#include "tables.cuh"

namespace pyn {

constexpr double kAlphaW = 1e2;  // spectral.py:96
constexpr double kAlphaD = 1e3;  // spectral.py:97

// Tensor-product value of a 1-D table set at local point g for local
// node/corner a; factors multiplied in direction order (spectral.py:283-345).
// which < 0: all basis values; which = d: derivative along d.
template <int DIM, int LD>
PYN_HD double tensor_val(const double (*h)[LD], const double (*dh)[LD], const uint8_t* gt,
                         const uint8_t* at, int which) {
  double v = (which == 0) ? dh[gt[0]][at[0]] : h[gt[0]][at[0]];
#pragma unroll
  for (int d = 1; d < DIM; ++d) v = v * ((which == d) ? dh[gt[d]][at[d]] : h[gt[d]][at[d]]);
  return v;
}

// J = Hrsc . X at the point with tensor index gt, its inverse (row-major jinv[k*DIM+d]) and det.
template <int DIM>
PYN_HD void point_geometry_t(const Rule1D& r, const uint8_t (*corn_t)[3], const uint8_t* gt,
                             const double* X, double* jinv, double* det) {
  constexpr int NC = 1 << DIM;
  double J[DIM][DIM];
#pragma unroll
  for (int d = 0; d < DIM; ++d)
#pragma unroll
    for (int e = 0; e < DIM; ++e) J[d][e] = 0.0;
  for (int c = 0; c < NC; ++c) {
#pragma unroll
    for (int d = 0; d < DIM; ++d) {
      double hv = tensor_val<DIM, 2>(r.hc, r.dhc, gt, corn_t[c], d);
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
// ... at local point g of the rule
template <int DIM>
PYN_HD void point_geometry(const Rule1D& r, const uint8_t (*corn_t)[3], int g, const double* X,
                           double* jinv, double* det) {
  point_geometry_t<DIM>(r, corn_t, r.gp_t[g], X, jinv, det);
}

// quadrature weight of local point g: w[c1]*w[c2](*w[c3]) left to right (utilities.py:15-41)
template <int DIM>
PYN_HD double point_weight(const Rule1D& r, int g) {
  double w = r.w[r.gp_t[g][0]];
#pragma unroll
  for (int d = 1; d < DIM; ++d) w = w * r.w[r.gp_t[g][d]];
  return w;
}

// Physical gradient of the basis of node a at point g: G[k] = sum_d jinv[k][d] dN_a/dr_d,
// and the basis value H.
template <int DIM>
PYN_HD void node_gradient(const Rule1D& r, const uint8_t* node_t, int g, const double* jinv,
                          double* G, double* H) {
  double hr[DIM];
#pragma unroll
  for (int d = 0; d < DIM; ++d) hr[d] = tensor_val<DIM, kMaxNgl>(r.h, r.dh, r.gp_t[g], node_t, d);
  *H = tensor_val<DIM, kMaxNgl>(r.h, r.dh, r.gp_t[g], node_t, -1);
#pragma unroll
  for (int k = 0; k < DIM; ++k) {
    double s = jinv[k * DIM] * hr[0];
#pragma unroll
    for (int d = 1; d < DIM; ++d) s += jinv[k * DIM + d] * hr[d];
    G[k] = s;
  }
}

// Entry [r][c] of the discrete curl built from a node gradient g[]:
//   3-D (indCurl / indWCurl, spectral.py:30-31): [[0,-g2,g1],[g2,0,-g0],[-g1,g0,0]]
//   2-D velocity curl (indCurl, :27): row 0 = [-g1, +g0]
//   2-D vorticity curl (indWCurl, :26): column 0 = [+g1, -g0]^T
template <int DIM>
PYN_HD double curl_v(int r, int c, const double* g) {  // rows: dim_w, cols: dim
  if (DIM == 2) return c == 0 ? -g[1] : g[0];
  if (r == c) return 0.0;
  const int k = 3 - r - c;
  const bool pos = ((r + 1) % 3) != c;  // (r, c=r+2 mod 3) is +, (r, c=r+1 mod 3) is -
  return pos ? g[k] : -g[k];
}
template <int DIM>
PYN_HD double curl_w(int r, int c, const double* g) {  // rows: dim, cols: dim_w
  if (DIM == 2) return r == 0 ? g[1] : -g[0];
  return curl_v<3>(r, c, g);
}

template <int DIM>
struct PairAcc {
  static constexpr int DW = DIM == 2 ? 1 : 3;
  double K[DIM][DIM];
  double Rw[DIM][DW];
  double Rd[DIM];
  PYN_HD void clear() {
#pragma unroll
    for (int i = 0; i < DIM; ++i) {
#pragma unroll
      for (int j = 0; j < DIM; ++j) K[i][j] = 0.0;
#pragma unroll
      for (int j = 0; j < DW; ++j) Rw[i][j] = 0.0;
      Rd[i] = 0.0;
    }
  }
  // full-rule point (spectral.py:116-136)
  PYN_HD void add_full(double wd, const double* ga, double ha, const double* gb) {
    double lap = ga[0] * gb[0];
#pragma unroll
    for (int k = 1; k < DIM; ++k) lap += ga[k] * gb[k];
#pragma unroll
    for (int c = 0; c < DIM; ++c) K[c][c] += wd * lap;
#pragma unroll
    for (int r = 0; r < DIM; ++r) {
#pragma unroll
      for (int c = 0; c < DW; ++c) {
        if (DIM == 2 || r != c) Rw[r][c] += wd * (ha * curl_w<DIM>(r, c, gb));
      }
      Rd[r] -= wd * (ha * gb[r]);
    }
  }
  // reduced-rule point (spectral.py:138-159)
  PYN_HD void add_red(double wd, const double* ga, const double* gb, double hb) {
    const double wdw = wd * kAlphaW, wdd = wd * kAlphaD;
#pragma unroll
    for (int c = 0; c < DIM; ++c) {
#pragma unroll
      for (int e = 0; e < DIM; ++e) {
        double cc = 0.0;
#pragma unroll
        for (int r = 0; r < DW; ++r) {
          if (DIM == 2 || (r != c && r != e)) cc += curl_v<DIM>(r, c, ga) * curl_v<DIM>(r, e, gb);
        }
        K[c][e] += wd * (kAlphaD * (ga[c] * gb[e]) + kAlphaW * cc);
      }
#pragma unroll
      for (int r = 0; r < DW; ++r) {
        if (DIM == 2 || r != c) Rw[c][r] += wdw * (curl_v<DIM>(r, c, ga) * hb);
      }
      Rd[c] += wdd * (ga[c] * hb);
    }
  }
};

// ---- nodal-rule operators (spectral.py:162-228) -----------------------------
// Packed symmetric index of T_{x i} (indBdiv, spectral.py:29,33).
template <int DIM>
PYN_HD int sym_index(int x, int i) {
  if (DIM == 2) return x + i;                      // [[0,1],[1,2]]
  if (x == i) return 2 * x;                        // xx,yy,zz -> 0,2,4
  return (x + i == 2) ? 5 : x + i;                 // xy->1, yz->3, zx->5
}

template <int DIM>
struct OpsAcc {
  static constexpr int DW = DIM == 2 ? 1 : 3;
  static constexpr int DS = DIM == 2 ? 3 : 6;
  double SrT[DS][DIM];
  double Div[DIM][DS];
  double Curl[DW][DIM];
  PYN_HD void clear() {
#pragma unroll
    for (int s = 0; s < DS; ++s)
#pragma unroll
      for (int c = 0; c < DIM; ++c) {
        SrT[s][c] = 0.0;
        Div[c][s] = 0.0;
      }
#pragma unroll
    for (int r = 0; r < DW; ++r)
#pragma unroll
      for (int c = 0; c < DIM; ++c) Curl[r][c] = 0.0;
  }
  // one nodal point: wd = w*det, ha = H[g][a], gb = grad of node b at g
  PYN_HD void add(double wd, double ha, const double* gb) {
#pragma unroll
    for (int x = 0; x < DIM; ++x) {
#pragma unroll
      for (int i = 0; i < DIM; ++i) {
        // B_srt (after *0.5): row sym(x,i), column comp i gets d/dx_x; diagonal 2*0.5
        const double bs = (x == i) ? 0.5 * (2.0 * gb[x]) : 0.5 * gb[x];
        SrT[sym_index<DIM>(x, i)][i] += wd * (ha * bs);
        // B_div: row i, packed column sym(x,i) gets d/dx_x
        Div[i][sym_index<DIM>(x, i)] += wd * (ha * gb[x]);
      }
    }
#pragma unroll
    for (int r = 0; r < DW; ++r)
#pragma unroll
      for (int c = 0; c < DIM; ++c)
        if (DIM == 2 || r != c) Curl[r][c] += wd * (ha * curl_v<DIM>(r, c, gb));
  }
};

}  // namespace pyn
