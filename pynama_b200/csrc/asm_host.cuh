// asm_host.cuh -- host arithmetic of the element-block additive Schwarz preconditioner
// (asm_pc.cu): hashing of element blocks, grouping of elements with equal blocks, dense
// inversion, tiling of the elements by class.  Plain C++ so that tests/hostcheck runs it on CPU.
#pragma once
#include <stdint.h>
#include <math.h>
#include <algorithm>
#include <unordered_map>
#include <vector>
#include "common.cuh"

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

// a <- (a + a^T)/2: the preconditioner must be exactly symmetric for CG
inline void asm_symmetrise(int n, double* a) {
  for (int i = 0; i < n; ++i)
    for (int j = i + 1; j < n; ++j) {
      const double v = 0.5 * (a[(int64_t)i * n + j] + a[(int64_t)j * n + i]);
      a[(int64_t)i * n + j] = a[(int64_t)j * n + i] = v;
    }
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

}  // namespace pyn
