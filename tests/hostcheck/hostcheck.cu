// hostcheck.cu -- TEST-ONLY host harness.  Runs the host+device inline element
// arithmetic of pynama_b200/csrc/elem_math.cuh in plain CPU loops so that the
// formulas and the C++ reference-element tables can be checked against the
// golden vectors in the build container (no GPU there).  It is not part of
// libpynama_b200.so and nothing in the product loads it.
#include "../../pynama_b200/csrc/elem_math.cuh"
#include "../../pynama_b200/csrc/classes.cuh"
#include "../../pynama_b200/csrc/asm_host.cuh"

#include <string.h>

using namespace pyn;

template <int DIM>
static void kle_host(const ElemTables& t, const double* X, double* K, double* Rw, double* Rd) {
  constexpr int DW = DIM == 2 ? 1 : 3;
  const int nn = t.nn;
  std::vector<PairAcc<DIM>> acc((size_t)nn * nn);
  for (auto& a : acc) a.clear();
  for (int q = 0; q < 2; ++q) {
    const Rule1D& r = t.rule[q == 0 ? RULE_FULL : RULE_RED];
    std::vector<double> G((size_t)nn * DIM), H(nn);
    for (int g = 0; g < r.npts; ++g) {
      double jinv[9], det;
      point_geometry<DIM>(r, t.corn_t, g, X, jinv, &det);
      const double wd = point_weight<DIM>(r, g) * det;
      for (int b = 0; b < nn; ++b) node_gradient<DIM>(r, t.node_t[b], g, jinv, &G[(size_t)b * DIM], &H[b]);
      for (int a = 0; a < nn; ++a)
        for (int b = 0; b < nn; ++b) {
          if (q == 0) acc[(size_t)a * nn + b].add_full(wd, &G[(size_t)a * DIM], H[a], &G[(size_t)b * DIM]);
          else acc[(size_t)a * nn + b].add_red(wd, &G[(size_t)a * DIM], &G[(size_t)b * DIM], H[b]);
        }
    }
  }
  for (int a = 0; a < nn; ++a)
    for (int b = 0; b < nn; ++b) {
      const PairAcc<DIM>& p = acc[(size_t)a * nn + b];
      for (int c = 0; c < DIM; ++c) {
        for (int e = 0; e < DIM; ++e) K[((size_t)a * DIM + c) * nn * DIM + b * DIM + e] = p.K[c][e];
        for (int e = 0; e < DW; ++e) Rw[((size_t)a * DIM + c) * nn * DW + b * DW + e] = p.Rw[c][e];
        Rd[((size_t)a * DIM + c) * nn + b] = p.Rd[c];
      }
    }
}

template <int DIM>
static void ops_host(const ElemTables& t, const double* X, double* S, double* D, double* C, double* W) {
  constexpr int DW = DIM == 2 ? 1 : 3;
  constexpr int DS = DIM == 2 ? 3 : 6;
  const int nn = t.nn;
  const Rule1D& r = t.rule[RULE_NODAL];
  for (int a = 0; a < nn; ++a) {
    double jinv[9], det;
    point_geometry<DIM>(r, t.corn_t, a, X, jinv, &det);
    const double wd = point_weight<DIM>(r, a) * det;
    W[a] = wd;
    const double ha = tensor_val<DIM, kMaxNgl>(r.h, r.dh, r.gp_t[a], t.node_t[a], -1);
    for (int b = 0; b < nn; ++b) {
      double G[DIM], Hb;
      node_gradient<DIM>(r, t.node_t[b], a, jinv, G, &Hb);
      OpsAcc<DIM> acc;
      acc.clear();
      acc.add(wd, ha, G);
      for (int s = 0; s < DS; ++s)
        for (int c = 0; c < DIM; ++c) {
          S[((size_t)a * DS + s) * nn * DIM + b * DIM + c] = acc.SrT[s][c];
          D[((size_t)a * DIM + c) * nn * DS + b * DS + s] = acc.Div[c][s];
        }
      for (int rr = 0; rr < DW; ++rr)
        for (int c = 0; c < DIM; ++c) C[((size_t)a * DW + rr) * nn * DIM + b * DIM + c] = acc.Curl[rr][c];
    }
  }
}

extern "C" {
// congruence classes of element shapes (classes.cuh): cls[nel], rep[>= nclass]; returns nclass
long long hc_congruence_classes(int dim, long long nel, const double* corners, double tol, int* cls, int* rep) {
  std::vector<int32_t> c, r;
  const int64_t n = congruence_classes(dim, nel, corners, tol, &c, &r);
  for (long long e = 0; e < nel; ++e) cls[e] = c[(size_t)e];
  for (int64_t k = 0; k < n; ++k) rep[k] = r[(size_t)k];
  return n;
}
// host arithmetic of the element-block Schwarz preconditioner (asm_host.cuh)
int hc_asm_dense_inverse(int n, double* a, int symmetrise) {
  const int rc = asm_dense_inverse(n, a);
  if (rc == 0 && symmetrise) asm_symmetrise(n, a);
  return rc;
}
// the two hashes of a dense ld x ld block, entries visited in the order perm[] (NULL: natural)
void hc_asm_block_hash(int ld, const double* a, const long long* perm, unsigned long long* out) {
  uint64_t h0 = 0, h1 = 0;
  for (long long t = 0; t < (long long)ld * ld; ++t) {
    const long long pos = perm ? perm[t] : t;
    const double v = a[pos];
    if (v == 0.0) continue;
    uint64_t bits;
    memcpy(&bits, &v, 8);
    h0 += asm_hash_entry(bits, (uint64_t)pos, 0);
    h1 += asm_hash_entry(bits, (uint64_t)pos, 1);
  }
  out[0] = h0;
  out[1] = h1;
}
long long hc_asm_group_by_hash(long long nel, const unsigned long long* h, int* cls, int* rep) {
  std::vector<int32_t> c, r;
  const int64_t n = asm_group_by_hash(nel, reinterpret_cast<const uint64_t*>(h), &c, &r);
  for (long long e = 0; e < nel; ++e) cls[e] = c[(size_t)e];
  for (int64_t k = 0; k < n; ++k) rep[k] = r[(size_t)k];
  return n;
}
// tile_elems[>= (nel + nclass*te)], tile_class[>= nel + nclass]; returns the number of tiles
long long hc_asm_make_tiles(long long nel, const int* cls, long long nclass, int te, int* tile_elems,
                            int* tile_class) {
  std::vector<int32_t> tev, tcv;
  const int64_t nt = asm_make_tiles(nel, cls, nclass, te, &tev, &tcv);
  for (size_t i = 0; i < tev.size(); ++i) tile_elems[i] = tev[i];
  for (size_t i = 0; i < tcv.size(); ++i) tile_class[i] = tcv[i];
  return nt;
}
// coarse tables of the two-level preconditioner: returns nv; out arrays sized by the caller
// (vert_node[nnodes], elem_verts[nel*ncorner], color[nnodes], adj_ptr[nnodes+1], adj[nnodes*125]); info = {ncolors}
long long hc_asm_coarse_tables(int dim, int ngl, long long nel, long long nnodes, const int* elem_nodes, int* vert_node,
                               int* elem_verts, int* color, long long* adj_ptr, int* adj, int* info, double* phi_out) {
  ElemTables t;
  if (build_elem_tables(dim, ngl, &t) != 0) return -1;
  int corner_local[8];
  std::vector<double> phi;
  if (asm_coarse_phi(t, corner_local, &phi) != 0) return -2;
  for (size_t i = 0; i < phi.size(); ++i) phi_out[i] = phi[i];
  AsmCoarseTables T;
  asm_coarse_tables(nel, nnodes, t.nn, t.ncorner, elem_nodes, corner_local, &T);
  for (int64_t v = 0; v < T.nv; ++v) {
    vert_node[v] = T.vert_node[(size_t)v];
    color[v] = T.color[(size_t)v];
  }
  for (size_t i = 0; i < T.elem_verts.size(); ++i) elem_verts[i] = T.elem_verts[i];
  for (int64_t v = 0; v <= T.nv; ++v) adj_ptr[v] = T.adj_ptr[(size_t)v];
  for (size_t i = 0; i < T.adj.size(); ++i) adj[i] = T.adj[i];
  info[0] = T.ncolors;
  return T.nv;
}
void hc_asm_cheb_coefficients(int k, double lmin, double lmax, double* c1, double* c2) {
  std::vector<double> a, b;
  asm_cheb_coefficients(k, lmin, lmax, &a, &b);
  for (int s = 0; s <= k; ++s) {
    c1[s] = a[(size_t)s];
    c2[s] = b[(size_t)s];
  }
}
int hc_kle(int dim, int ngl, const double* X, double* K, double* Rw, double* Rd) {
  ElemTables t;
  if (build_elem_tables(dim, ngl, &t) != 0) return -1;
  if (dim == 2) kle_host<2>(t, X, K, Rw, Rd); else kle_host<3>(t, X, K, Rw, Rd);
  return 0;
}
int hc_ops(int dim, int ngl, const double* X, double* S, double* D, double* C, double* W) {
  ElemTables t;
  if (build_elem_tables(dim, ngl, &t) != 0) return -1;
  if (dim == 2) ops_host<2>(t, X, S, D, C, W); else ops_host<3>(t, X, S, D, C, W);
  return 0;
}
}
