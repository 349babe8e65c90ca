// asm_host.cuh -- host arithmetic of the element-block additive Schwarz preconditioner
// (asm_pc.cu): hashing of element blocks, grouping of elements with equal blocks, dense
// inversion, tiling of the elements by class.  Plain C++ so that tests/hostcheck runs it on CPU.
#pragma once
#include <stdint.h>
#include <math.h>
#include <algorithm>
#include <unordered_map>
#include <vector>
#include "tables.cuh"

namespace pyn {

// 64-bit finaliser (splitmix64)
PYN_HD uint64_t asm_mix(uint64_t x) {
  x += 0x9E3779B97F4A7C15ull;
  x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ull;
  x = (x ^ (x >> 27)) * 0x94D049BB133111EBull;
  return x ^ (x >> 31);
}

// Contribution of entry `pos` (row*ld + col of the dense element block) with value bits `bits`
// to hash `which` (0, 1) of the block.  The block hash is the wrapping SUM of these, so the
// order in which the entries are visited does not matter; zero entries are skipped by the caller.
PYN_HD uint64_t asm_hash_entry(uint64_t bits, uint64_t pos, int which) {
  const uint64_t seed = which ? 0xD6E8FEB86659FD93ull : 0xA0761D6478BD642Full;
  return asm_mix(bits ^ asm_mix(pos * 2 + 1 + seed));
}

// Elements whose two block hashes agree form one class.  cls[e] = class of element e, numbered
// by lowest member; rep[k] = lowest element of class k.  Returns the number of classes.
inline int64_t asm_group_by_hash(int64_t nel, const uint64_t* h /* [nel][2] */, std::vector<int32_t>* cls,
                                 std::vector<int32_t>* rep) {
  struct Key {
    uint64_t a, b;
    bool operator==(const Key& o) const { return a == o.a && b == o.b; }
  };
  struct KeyHash {
    size_t operator()(const Key& k) const { return (size_t)(k.a ^ (k.b * 0x9E3779B97F4A7C15ull)); }
  };
  std::unordered_map<Key, int32_t, KeyHash> seen;
  cls->assign((size_t)nel, 0);
  rep->clear();
  for (int64_t e = 0; e < nel; ++e) {
    const Key k{h[2 * e], h[2 * e + 1]};
    auto it = seen.find(k);
    if (it == seen.end()) {
      it = seen.emplace(k, (int32_t)rep->size()).first;
      rep->push_back((int32_t)e);
    }
    (*cls)[(size_t)e] = it->second;
  }
  return (int64_t)rep->size();
}

// Row-partitioned runs: classes from the block hashes of the COMPLETE elements (all rows local),
// the others by their signature (asm_signature_kernel).  hash/sig: [nel][2].  Returns the number of
// classes, or -2 - e when incomplete element e has no complete twin.  Complete elements with one
// signature may still fall into several classes (blocks that differ in the last bits, or a mesh on
// which the signature says too little): the lowest such element lends its inverse, which keeps the
// preconditioner symmetric positive definite whatever the blocks are.
inline int64_t asm_group_partitioned(int64_t nel, const uint8_t* complete, const uint64_t* hash, const uint64_t* sig,
                                     std::vector<int32_t>* cls, std::vector<int32_t>* rep) {
  std::vector<int32_t> ids;
  std::vector<uint64_t> hc;
  for (int64_t e = 0; e < nel; ++e)
    if (complete[e]) {
      ids.push_back((int32_t)e);
      hc.push_back(hash[2 * e]);
      hc.push_back(hash[2 * e + 1]);
    }
  std::vector<int32_t> cc, rc;
  const int64_t nclass = asm_group_by_hash((int64_t)ids.size(), hc.data(), &cc, &rc);
  cls->assign((size_t)nel, -1);
  rep->assign((size_t)nclass, 0);
  for (int64_t k = 0; k < nclass; ++k) (*rep)[(size_t)k] = ids[(size_t)rc[(size_t)k]];
  struct Key {
    uint64_t a, b;
    bool operator==(const Key& o) const { return a == o.a && b == o.b; }
  };
  struct KeyHash {
    size_t operator()(const Key& k) const { return (size_t)(k.a ^ (k.b * 0x9E3779B97F4A7C15ull)); }
  };
  std::unordered_map<Key, int32_t, KeyHash> by_sig;
  for (size_t i = 0; i < ids.size(); ++i) {
    const int64_t e = ids[i];
    (*cls)[(size_t)e] = cc[i];
    const Key k{sig[2 * e], sig[2 * e + 1]};
    auto it = by_sig.find(k);
    if (it == by_sig.end()) by_sig.emplace(k, cc[i]);
  }
  for (int64_t e = 0; e < nel; ++e)
    if (!complete[e]) {
      auto it = by_sig.find(Key{sig[2 * e], sig[2 * e + 1]});
      if (it == by_sig.end()) return -2 - e;
      (*cls)[(size_t)e] = it->second;
    }
  return nclass;
}

// In-place inverse of the dense n x n matrix a (row-major) by Gauss-Jordan elimination with
// partial pivoting.  Returns 0, or -1 when a pivot vanishes (singular to working precision).
inline int asm_dense_inverse(int n, double* a) {
  std::vector<int> piv((size_t)n);
  double amax = 0.0;
  for (int64_t i = 0; i < (int64_t)n * n; ++i) amax = std::max(amax, fabs(a[i]));
  if (!(amax > 0.0) || !std::isfinite(amax)) return -1;
  for (int k = 0; k < n; ++k) {
    int p = k;
    double best = fabs(a[(int64_t)k * n + k]);
    for (int i = k + 1; i < n; ++i) {
      const double v = fabs(a[(int64_t)i * n + k]);
      if (v > best) {
        best = v;
        p = i;
      }
    }
    if (!(best > amax * 1e-300) || !std::isfinite(best)) return -1;
    piv[(size_t)k] = p;
    if (p != k)
      for (int j = 0; j < n; ++j) std::swap(a[(int64_t)k * n + j], a[(int64_t)p * n + j]);
    double* rk = a + (int64_t)k * n;
    const double d = 1.0 / rk[k];
    rk[k] = 1.0;
    for (int j = 0; j < n; ++j) rk[j] *= d;
    for (int i = 0; i < n; ++i) {
      if (i == k) continue;
      double* ri = a + (int64_t)i * n;
      const double f = ri[k];
      if (f == 0.0) continue;
      ri[k] = 0.0;
      for (int j = 0; j < n; ++j) ri[j] -= f * rk[j];
    }
  }
  for (int k = n - 1; k >= 0; --k) {          // undo the row swaps on the columns
    const int p = piv[(size_t)k];
    if (p != k)
      for (int i = 0; i < n; ++i) std::swap(a[(int64_t)i * n + k], a[(int64_t)i * n + p]);
  }
  return 0;
}

// a <- (a + a^T)/2 when a is symmetric up to rounding (the preconditioner must be exactly symmetric
// for CG); a visibly unsymmetric inverse (an unsymmetric operator under GMRES / BiCGStab) is left
// alone.  Returns whether it symmetrised.
inline bool asm_symmetrise(int n, double* a) {
  double amax = 0.0, dmax = 0.0;
  for (int i = 0; i < n; ++i)
    for (int j = 0; j < n; ++j) {
      amax = std::max(amax, fabs(a[(int64_t)i * n + j]));
      if (j > i) dmax = std::max(dmax, fabs(a[(int64_t)i * n + j] - a[(int64_t)j * n + i]));
    }
  if (dmax > 1e-8 * amax) return false;
  for (int i = 0; i < n; ++i)
    for (int j = i + 1; j < n; ++j) {
      const double v = 0.5 * (a[(int64_t)i * n + j] + a[(int64_t)j * n + i]);
      a[(int64_t)i * n + j] = a[(int64_t)j * n + i] = v;
    }
  return true;
}

// The product kernel reads row k of the stored matrix for column k of the inverse: an inverse that
// was not symmetrised is stored transposed.
inline void asm_transpose(int n, double* a) {
  for (int i = 0; i < n; ++i)
    for (int j = i + 1; j < n; ++j) std::swap(a[(int64_t)i * n + j], a[(int64_t)j * n + i]);
}

// Tiles of `te` elements of one class each (the last tile of a class padded with -1), classes
// in order, elements ascending inside a class.
inline int64_t asm_make_tiles(int64_t nel, const int32_t* cls, int64_t nclass, int te,
                              std::vector<int32_t>* tile_elems, std::vector<int32_t>* tile_class) {
  std::vector<std::vector<int32_t>> members((size_t)nclass);
  for (int64_t e = 0; e < nel; ++e) members[(size_t)cls[e]].push_back((int32_t)e);
  tile_elems->clear();
  tile_class->clear();
  for (int64_t c = 0; c < nclass; ++c) {
    const std::vector<int32_t>& m = members[(size_t)c];
    for (size_t i0 = 0; i0 < m.size(); i0 += (size_t)te) {
      for (int j = 0; j < te; ++j) tile_elems->push_back(i0 + j < m.size() ? m[i0 + j] : -1);
      tile_class->push_back((int32_t)c);
    }
  }
  return (int64_t)tile_class->size();
}

// ---- Q1 coarse space on the element corners (second level of the preconditioner) -----------------
// Coarse vertices = the fine nodes that are a corner of some element, numbered by ascending node.
struct AsmCoarseTables {
  int64_t nv = 0;
  int ncolors = 0;
  std::vector<int32_t> vert_node;    // [nv] fine node of coarse vertex v
  std::vector<int32_t> elem_verts;   // [nel][ncorner]
  std::vector<int64_t> v2e_ptr;      // [nv+1]
  std::vector<int32_t> v2e;          // [.][2] (element, corner) of vertex v, ascending element
  std::vector<int64_t> adj_ptr;      // [nv+1] vertices sharing an element with v (v included), ascending
  std::vector<int32_t> adj;
  std::vector<int32_t> color;        // [nv] no two vertices adjacent to a common vertex share a colour
};

inline void asm_coarse_tables(int64_t nel, int64_t nnodes, int nn, int ncorner, const int32_t* elem_nodes,
                              const int* corner_local, AsmCoarseTables* t) {
  std::vector<int32_t> vid((size_t)nnodes, -1);
  for (int64_t e = 0; e < nel; ++e)
    for (int c = 0; c < ncorner; ++c) vid[(size_t)elem_nodes[e * nn + corner_local[c]]] = 0;
  t->vert_node.clear();
  for (int64_t n = 0; n < nnodes; ++n)
    if (vid[(size_t)n] == 0) {
      vid[(size_t)n] = (int32_t)t->vert_node.size();
      t->vert_node.push_back((int32_t)n);
    }
  const int64_t nv = t->nv = (int64_t)t->vert_node.size();
  t->elem_verts.assign((size_t)(nel * ncorner), 0);
  t->v2e_ptr.assign((size_t)nv + 1, 0);
  for (int64_t e = 0; e < nel; ++e)
    for (int c = 0; c < ncorner; ++c) {
      const int32_t v = vid[(size_t)elem_nodes[e * nn + corner_local[c]]];
      t->elem_verts[(size_t)(e * ncorner + c)] = v;
      t->v2e_ptr[(size_t)v + 1]++;
    }
  for (int64_t v = 0; v < nv; ++v) t->v2e_ptr[(size_t)v + 1] += t->v2e_ptr[(size_t)v];
  t->v2e.assign((size_t)(2 * t->v2e_ptr[(size_t)nv]), 0);
  std::vector<int64_t> fill(t->v2e_ptr.begin(), t->v2e_ptr.end() - 1);
  for (int64_t e = 0; e < nel; ++e)              // ascending element inside every list
    for (int c = 0; c < ncorner; ++c) {
      const int32_t v = t->elem_verts[(size_t)(e * ncorner + c)];
      const int64_t k = fill[(size_t)v]++;
      t->v2e[(size_t)(2 * k)] = (int32_t)e;
      t->v2e[(size_t)(2 * k + 1)] = c;
    }
  t->adj_ptr.assign((size_t)nv + 1, 0);
  t->adj.clear();
  std::vector<int32_t> tmp;
  for (int64_t v = 0; v < nv; ++v) {
    tmp.clear();
    for (int64_t k = t->v2e_ptr[(size_t)v]; k < t->v2e_ptr[(size_t)v + 1]; ++k) {
      const int64_t e = t->v2e[(size_t)(2 * k)];
      for (int c = 0; c < ncorner; ++c) tmp.push_back(t->elem_verts[(size_t)(e * ncorner + c)]);
    }
    std::sort(tmp.begin(), tmp.end());
    tmp.erase(std::unique(tmp.begin(), tmp.end()), tmp.end());
    t->adj.insert(t->adj.end(), tmp.begin(), tmp.end());
    t->adj_ptr[(size_t)v + 1] = (int64_t)t->adj.size();
  }
  // greedy distance-2 colouring: a colour may be used by v only if no vertex within two adjacency
  // steps has it, so every vertex sees at most one vertex of a colour among its neighbours
  t->color.assign((size_t)nv, -1);
  t->ncolors = 0;
  std::vector<int64_t> stamp;
  for (int64_t v = 0; v < nv; ++v) {
    stamp.assign((size_t)t->ncolors + 1, -1);
    for (int64_t k = t->adj_ptr[(size_t)v]; k < t->adj_ptr[(size_t)v + 1]; ++k) {
      const int64_t w = t->adj[(size_t)k];
      for (int64_t k2 = t->adj_ptr[(size_t)w]; k2 < t->adj_ptr[(size_t)w + 1]; ++k2) {
        const int32_t cw = t->color[(size_t)t->adj[(size_t)k2]];
        if (cw >= 0) stamp[(size_t)cw] = v;
      }
    }
    int c = 0;
    while (c < t->ncolors && stamp[(size_t)c] == v) ++c;
    t->color[(size_t)v] = c;
    if (c == t->ncolors) t->ncolors++;
  }
}

// corner_local[c] = local node sitting on corner c; phi[a*ncorner + c] = value at local node a of
// the multilinear shape function of corner c (the Q1 prolongation weights)
inline int asm_coarse_phi(const ElemTables& t, int* corner_local, std::vector<double>* phi) {
  const Rule1D& nodal = t.rule[RULE_NODAL];
  phi->assign((size_t)t.nn * t.ncorner, 0.0);
  for (int c = 0; c < t.ncorner; ++c) {
    corner_local[c] = -1;
    for (int a = 0; a < t.nn; ++a) {
      bool here = true;
      double w = 1.0;
      for (int d = 0; d < t.dim; ++d) {
        here = here && t.node_t[a][d] == t.corn_t[c][d] * (t.ngl - 1);
        w *= nodal.hc[t.node_t[a][d]][t.corn_t[c][d]];
      }
      if (fabs(w) < 1e-14) w = 0.0;             // the other corners' functions vanish on a face
      (*phi)[(size_t)a * t.ncorner + c] = w;
      if (here) corner_local[c] = a;
    }
    if (corner_local[c] < 0) return -1;
  }
  return 0;
}

// position of column w in the (ascending) list of row v, or -1
inline int64_t asm_coarse_find(const int64_t* ptr, const int32_t* col, int64_t v, int32_t w) {
  const int32_t* lo = col + ptr[v];
  const int32_t* hi = col + ptr[v + 1];
  const int32_t* it = std::lower_bound(lo, hi, w);
  return (it != hi && *it == w) ? (int64_t)(it - col) : -1;
}

// block (v,w) <- (block (v,w) + block (w,v)^T)/2 of a block CSR with a symmetric pattern
inline void asm_coarse_symmetrise(int64_t nv, int B, const int64_t* ptr, const int32_t* col, double* val) {
  for (int64_t v = 0; v < nv; ++v)
    for (int64_t k = ptr[v]; k < ptr[v + 1]; ++k) {
      const int32_t w = col[k];
      if (w < v) continue;
      const int64_t k2 = w == v ? k : asm_coarse_find(ptr, col, w, (int32_t)v);
      if (k2 < 0) continue;
      for (int i = 0; i < B; ++i)
        for (int j = 0; j < B; ++j) {
          if (k2 == k && j < i) continue;
          double& a = val[(k * B + i) * B + j];
          double& b = val[(k2 * B + j) * B + i];
          const double m = 0.5 * (a + b);
          a = b = m;
        }
    }
}

// dinv[v*B+i] = 1/Ac[(v,i),(v,i)], or 0 for an inactive coarse dof (empty column of P: every fine dof
// it would reach is constrained)
inline void asm_coarse_jacobi(int64_t nv, int B, const int64_t* ptr, const int32_t* col, const double* val,
                              double* dinv) {
  double dmax = 0.0;
  for (int64_t v = 0; v < nv; ++v) {
    const int64_t k = asm_coarse_find(ptr, col, v, (int32_t)v);
    for (int i = 0; i < B; ++i) {
      const double d = k >= 0 ? val[(k * B + i) * B + i] : 0.0;
      dinv[v * B + i] = d;
      dmax = std::max(dmax, d);
    }
  }
  for (int64_t i = 0; i < nv * B; ++i) dinv[i] = dinv[i] > 1e-12 * dmax ? 1.0 / dinv[i] : 0.0;
}

// largest eigenvalue of D^-1 A (A block CSR over the adjacency lists, B x B blocks row-major, D its
// diagonal) by power iteration from a fixed start; dinv[i] = 0 marks an inactive coarse dof
inline double asm_power_lmax(int64_t nv, int B, const int64_t* ptr, const int32_t* col, const double* val,
                             const double* dinv, int iters) {
  const int64_t n = nv * B;
  std::vector<double> x((size_t)n), y((size_t)n);
  for (int64_t i = 0; i < n; ++i) x[(size_t)i] = dinv[i] != 0.0 ? 1.0 + 0.37 * (double)((i * 7919) % 13) : 0.0;
  double lam = 0.0;
  for (int it = 0; it < iters; ++it) {
    double nx = 0.0;
    for (int64_t i = 0; i < n; ++i) nx += x[(size_t)i] * x[(size_t)i];
    nx = sqrt(nx);
    if (!(nx > 0.0)) return 0.0;
    for (int64_t i = 0; i < n; ++i) x[(size_t)i] /= nx;
    for (int64_t v = 0; v < nv; ++v)
      for (int i = 0; i < B; ++i) {
        double s = 0.0;
        for (int64_t k = ptr[v]; k < ptr[v + 1]; ++k)
          for (int j = 0; j < B; ++j) s += val[(k * B + i) * B + j] * x[(size_t)(col[k] * (int64_t)B + j)];
        y[(size_t)(v * B + i)] = dinv[v * B + i] * s;
      }
    double ny = 0.0;
    for (int64_t i = 0; i < n; ++i) ny += y[(size_t)i] * y[(size_t)i];
    lam = sqrt(ny);
    x.swap(y);
  }
  return lam;
}

// coefficients of k Chebyshev steps on [lmin, lmax] (zero start):
//   d_0 = c2[0] D^-1 r;  step s = 1..k: x += d, r -= A d, d = c1[s] d + c2[s] D^-1 r
inline void asm_cheb_coefficients(int k, double lmin, double lmax, std::vector<double>* c1, std::vector<double>* c2) {
  const double theta = 0.5 * (lmax + lmin), delta = 0.5 * (lmax - lmin);
  const double sigma = theta / delta;
  double rho = 1.0 / sigma;
  c1->assign((size_t)k + 1, 0.0);
  c2->assign((size_t)k + 1, 0.0);
  (*c2)[0] = 1.0 / theta;
  for (int s = 1; s <= k; ++s) {
    const double rho_new = 1.0 / (2.0 * sigma - rho);
    (*c1)[(size_t)s] = rho_new * rho;
    (*c2)[(size_t)s] = 2.0 * rho_new / delta;
    rho = rho_new;
  }
}

}  // namespace pyn
