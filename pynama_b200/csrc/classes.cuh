// classes.cuh -- congruence classes of element shapes (host code, no device parts).
//
// getElemKLEMatrices / getElemKLEOperators (spectral.py:92-228) depend on a cell only through
// the offsets of its corners from corner 0, and the reference evaluates them again for every
// cell of a uniform box (mat_fs.py:152-155, 196-198).  Two elements are congruent here when
// those offsets, rounded to multiples of tol * (largest offset in the mesh), are equal; the
// representative of a class is its lowest element.
#pragma once
#include <stdint.h>

#include <algorithm>
#include <cmath>
#include <unordered_map>
#include <vector>

namespace pyn {

// corners: [nel][2^dim][dim].  Returns the number of classes and fills cls[nel] (class of the
// element, numbered by first appearance) and rep[nclass].  tol <= 0, fewer than two elements
// or a degenerate / non-finite mesh: one class per element.
inline int64_t congruence_classes(int dim, int64_t nel, const double* corners, double tol,
                                  std::vector<int32_t>* cls, std::vector<int32_t>* rep) {
  const int nv = (1 << dim) * dim;
  cls->resize((size_t)nel);
  rep->clear();
  auto identity = [&]() {
    rep->resize((size_t)nel);
    for (int64_t e = 0; e < nel; ++e) (*cls)[(size_t)e] = (*rep)[(size_t)e] = (int32_t)e;
    return nel;
  };
  if (!(tol > 0.0) || nel < 2) return identity();
  double scale = 0.0;
  bool finite = true;
  for (int64_t e = 0; e < nel; ++e) {
    const double* X = corners + e * nv;
    for (int k = dim; k < nv; ++k) {
      const double d = std::fabs(X[k] - X[k % dim]);
      finite = finite && std::isfinite(d);
      scale = std::max(scale, d);
    }
  }
  if (!finite || !(scale > 0.0)) return identity();
  const double inv_q = 1.0 / (tol * scale);
  std::vector<int64_t> keys;                       // [nclass][nv] quantised offsets of the representatives
  std::unordered_multimap<uint64_t, int32_t> table;
  table.reserve(1024);
  std::vector<int64_t> key((size_t)nv);
  for (int64_t e = 0; e < nel; ++e) {
    const double* X = corners + e * nv;
    uint64_t h = 0x9E3779B97F4A7C15ull;
    for (int k = 0; k < nv; ++k) {
      key[k] = std::llround((X[k] - X[k % dim]) * inv_q);
      h = (h ^ (uint64_t)key[k]) * 0xFF51AFD7ED558CCDull;
      h ^= h >> 29;
    }
    int32_t found = -1;
    auto range = table.equal_range(h);
    for (auto it = range.first; it != range.second; ++it)
      if (std::equal(key.begin(), key.end(), keys.begin() + (size_t)it->second * nv)) {
        found = it->second;
        break;
      }
    if (found < 0) {
      found = (int32_t)rep->size();
      rep->push_back((int32_t)e);
      keys.insert(keys.end(), key.begin(), key.end());
      table.emplace(h, found);
    }
    (*cls)[(size_t)e] = found;
  }
  return (int64_t)rep->size();
}

}  // namespace pyn
