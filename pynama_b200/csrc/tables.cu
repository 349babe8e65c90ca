// tables.cu -- host-side reference-element tables (see tables.cuh).
#include <math.h>
#include <stdarg.h>

#include "tables.cuh"

namespace pyn {

static thread_local char g_err[1024] = "";

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}
const char* get_error() { return g_err; }

// Legendre P_n(x) and P_{n-1}(x) by the three-term recurrence.
static void legendre(int n, double x, double* pn, double* pnm1) {
  double p0 = 1.0, p1 = x;
  if (n == 0) {
    *pn = 1.0;
    *pnm1 = 0.0;
    return;
  }
  for (int k = 2; k <= n; ++k) {
    double p2 = ((2 * k - 1) * x * p1 - (k - 1) * p0) / k;
    p0 = p1;
    p1 = p2;
  }
  *pn = p1;
  *pnm1 = p0;
}

// kind 0: Gauss-Legendre (utilities.py:43-61 computes the same rule from the
// Jacobi matrix; here Newton on P_n).  kind 1: Gauss-Lobatto-Legendre
// (utilities.py:63-92).  Both are symmetrised like the reference.
int rule_1d(int kind, int n, double* x, double* w) {
  PYN_REQUIRE(n >= 1 && n <= 64, "rule_1d: n=%d out of range", n);
  std::vector<double> xs(n), ws(n);
  if (kind == 0) {
    for (int i = 0; i < n; ++i) {
      // ascending roots
      double z = -cos(M_PI * (i + 0.75) / (n + 0.5));
      double pn = 0, pm = 0, dp = 1;
      for (int it = 0; it < 100; ++it) {
        legendre(n, z, &pn, &pm);
        dp = n * (z * pn - pm) / (z * z - 1.0);
        double dz = pn / dp;
        z -= dz;
        if (fabs(dz) < 1e-16) break;
      }
      legendre(n, z, &pn, &pm);
      dp = n * (z * pn - pm) / (z * z - 1.0);
      xs[i] = z;
      ws[i] = 2.0 / ((1.0 - z * z) * dp * dp);
    }
  } else if (kind == 1) {
    PYN_REQUIRE(n >= 2, "Gauss-Lobatto needs n >= 2");
    // Newton from the Chebyshev-Gauss-Lobatto guess on q(x) = x P_{n-1} - P_{n-2}
    for (int i = 0; i < n; ++i) {
      double z = -cos(M_PI * i / (n - 1));
      double pn = 0, pm = 0;
      for (int it = 0; it < 200; ++it) {
        legendre(n - 1, z, &pn, &pm);
        double dz = (z * pn - pm) / (n * pn);
        z -= dz;
        if (fabs(dz) < 1e-16) break;
      }
      legendre(n - 1, z, &pn, &pm);
      xs[i] = z;
      ws[i] = 2.0 / ((double)(n - 1) * n * pn * pn);
    }
    xs[0] = -1.0;
    xs[n - 1] = 1.0;
  } else {
    set_error("rule_1d: unknown kind %d", kind);
    return PYN_E_INVALID;
  }
  for (int i = 0; i < n; ++i) {
    x[i] = (xs[i] - xs[n - 1 - i]) / 2.0;
    w[i] = (ws[i] + ws[n - 1 - i]) / 2.0;
  }
  return PYN_OK;
}

// Product form of the Lagrange basis (element.py:13-44): a point that sits on
// a node gives exact 0 / 1 because numerator and denominator factors coincide.
int lagrange_1d(int nn, const double* nodes, int npts, const double* pts, double* h, double* dh) {
  PYN_REQUIRE(nn >= 1 && npts >= 1, "lagrange_1d: empty input");
  for (int p = 0; p < npts; ++p) {
    for (int j = 0; j < nn; ++j) {
      double den = 1.0, num = 1.0;
      for (int k = 0; k < nn; ++k) {
        if (k == j) continue;
        den *= (-nodes[k] + nodes[j]);
        num *= (pts[p] - nodes[k]);
      }
      double dsum = 0.0;
      for (int m = 0; m < nn; ++m) {
        if (m == j) continue;
        double pr = 1.0;
        for (int k = 0; k < nn; ++k) {
          if (k == j || k == m) continue;
          pr *= (pts[p] - nodes[k]);
        }
        dsum += pr;
      }
      h[p * nn + j] = num / den;
      dh[p * nn + j] = dsum / den;
    }
  }
  return PYN_OK;
}

// Local order of an n^dim tensor grid: corners, edge interiors, face interiors,
// body (spectral.py:230-252 in 2-D, spectral.py:356-441 in 3-D).  out[a*dim+d]
// is the tensor index of local position a along reference direction d.
int local_order(int dim, int n, int32_t* out) {
  PYN_REQUIRE((dim == 2 || dim == 3) && n >= 1 && n <= 64, "local_order: bad dim/n");
  int cnt = 0;
  auto put2 = [&](int a, int b) {
    out[cnt * 2 + 0] = a;
    out[cnt * 2 + 1] = b;
    ++cnt;
  };
  auto put3 = [&](int a, int b, int c) {
    out[cnt * 3 + 0] = a;
    out[cnt * 3 + 1] = b;
    out[cnt * 3 + 2] = c;
    ++cnt;
  };
  if (n == 1) {
    for (int d = 0; d < dim; ++d) out[d] = 0;
    return PYN_OK;
  }
  const int hi = n - 1;
  if (dim == 2) {
    put2(hi, hi); put2(0, hi); put2(0, 0); put2(hi, 0);
    for (int t = n - 2; t >= 1; --t) put2(t, hi);
    for (int t = n - 2; t >= 1; --t) put2(0, t);
    for (int t = 1; t <= n - 2; ++t) put2(t, 0);
    for (int t = 1; t <= n - 2; ++t) put2(hi, t);
    for (int a = 1; a <= n - 2; ++a)
      for (int b = n - 2; b >= 1; --b) put2(a, b);
    return PYN_OK;
  }
  put3(0, 0, 0); put3(0, hi, 0); put3(hi, hi, 0); put3(hi, 0, 0);
  put3(0, 0, hi); put3(hi, 0, hi); put3(hi, hi, hi); put3(0, hi, hi);
#define UP(t) for (int t = 1; t <= n - 2; ++t)
#define DN(t) for (int t = n - 2; t >= 1; --t)
  UP(t) put3(0, t, 0);
  UP(t) put3(t, hi, 0);
  DN(t) put3(hi, t, 0);
  DN(t) put3(t, 0, 0);
  UP(t) put3(t, 0, hi);
  UP(t) put3(hi, t, hi);
  DN(t) put3(t, hi, hi);
  DN(t) put3(0, t, hi);
  UP(t) put3(hi, 0, t);
  DN(t) put3(0, 0, t);
  UP(t) put3(0, hi, t);
  DN(t) put3(hi, hi, t);
  DN(y) UP(x) put3(x, y, 0);
  UP(y) DN(x) put3(x, y, hi);
  DN(x) UP(z) put3(x, 0, z);
  UP(x) DN(z) put3(x, hi, z);
  DN(z) DN(y) put3(hi, y, z);
  UP(z) UP(y) put3(0, y, z);
  DN(z) DN(y) UP(x) put3(x, y, z);
#undef UP
#undef DN
  return PYN_OK;
}

static int fill_rule(Rule1D* r, int dim, int ngl, const double* nodes, int npts1d,
                     const double* pts, const double* w) {
  PYN_REQUIRE(npts1d <= kMaxPts1d, "too many quadrature points");
  memset(r, 0, sizeof(*r));
  r->npts1d = npts1d;
  r->npts = 1;
  for (int d = 0; d < dim; ++d) r->npts *= npts1d;
  std::vector<double> h(npts1d * ngl), dh(npts1d * ngl), hc(npts1d * 2), dhc(npts1d * 2);
  const double cn[2] = {-1.0, 1.0};
  PYN_CHECK(lagrange_1d(ngl, nodes, npts1d, pts, h.data(), dh.data()));
  PYN_CHECK(lagrange_1d(2, cn, npts1d, pts, hc.data(), dhc.data()));
  for (int p = 0; p < npts1d; ++p) {
    r->w[p] = w[p];
    for (int j = 0; j < ngl; ++j) {
      r->h[p][j] = h[p * ngl + j];
      r->dh[p][j] = dh[p * ngl + j];
    }
    for (int j = 0; j < 2; ++j) {
      r->hc[p][j] = hc[p * 2 + j];
      r->dhc[p][j] = dhc[p * 2 + j];
    }
  }
  std::vector<int32_t> ord((size_t)r->npts * dim);
  PYN_CHECK(local_order(dim, npts1d, ord.data()));
  for (int g = 0; g < r->npts; ++g)
    for (int d = 0; d < dim; ++d) r->gp_t[g][d] = (uint8_t)ord[g * dim + d];
  return PYN_OK;
}

int build_elem_tables(int dim, int ngl, ElemTables* t) {
  PYN_REQUIRE(dim == 2 || dim == 3, "dim must be 2 or 3");
  PYN_REQUIRE(ngl >= 2 && ngl <= kMaxNgl, "ngl=%d out of range [2,%d]", ngl, kMaxNgl);
  memset(t, 0, sizeof(*t));
  t->dim = dim;
  t->ngl = ngl;
  t->nn = 1;
  for (int d = 0; d < dim; ++d) t->nn *= ngl;
  PYN_REQUIRE(t->nn <= kMaxNN, "element too large");
  t->ncorner = 1 << dim;
  t->dim_w = dim == 2 ? 1 : 3;
  t->dim_s = dim == 2 ? 3 : 6;
  double nodes[kMaxNgl], wn[kMaxNgl], xf[kMaxPts1d], wf[kMaxPts1d], xr[kMaxPts1d], wr[kMaxPts1d];
  PYN_CHECK(rule_1d(1, ngl, nodes, wn));
  // full rule: Gauss for ngl <= 3, else Lobatto (spectral.py:44-45, 69-70)
  PYN_CHECK(rule_1d(ngl <= 3 ? 0 : 1, ngl, xf, wf));
  PYN_CHECK(rule_1d(0, ngl - 1, xr, wr));  // reduced rule (spectral.py:46, 71)
  PYN_CHECK(fill_rule(&t->rule[RULE_FULL], dim, ngl, nodes, ngl, xf, wf));
  PYN_CHECK(fill_rule(&t->rule[RULE_RED], dim, ngl, nodes, ngl - 1, xr, wr));
  PYN_CHECK(fill_rule(&t->rule[RULE_NODAL], dim, ngl, nodes, ngl, nodes, wn));
  std::vector<int32_t> ord((size_t)t->nn * dim);
  PYN_CHECK(local_order(dim, ngl, ord.data()));
  for (int a = 0; a < t->nn; ++a)
    for (int d = 0; d < dim; ++d) t->node_t[a][d] = (uint8_t)ord[a * dim + d];
  int32_t cord[8 * 3];
  PYN_CHECK(local_order(dim, 2, cord));
  for (int c = 0; c < t->ncorner; ++c)
    for (int d = 0; d < dim; ++d) t->corn_t[c][d] = (uint8_t)cord[c * dim + d];
  return PYN_OK;
}

}  // namespace pyn
