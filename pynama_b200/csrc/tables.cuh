// tables.cuh -- reference-element tables (host side) and their device image.
//
// Replaces Spectral.__init__/setUpSpectralMats2D/3D (spectral.py:17-90),
// Element.interpFun1D (element.py:13-44), gaussPoints/lobattoPoints
// (utilities.py:43-92).  Only 1-D tables are kept; kernels rebuild the tensor
// products (factors multiplied in direction order, as the reference does).
#pragma once
#include "common.cuh"

namespace pyn {

constexpr int kMaxNgl = 8;           // nodes per edge
constexpr int kMaxPts1d = 8;         // quadrature points per direction
constexpr int kMaxNN = 512;          // nodes per element (8^3)

enum RuleId { RULE_FULL = 0, RULE_RED = 1, RULE_NODAL = 2, RULE_COUNT = 3 };

// One quadrature rule, 1-D data + tensor orderings (device-friendly POD).
struct Rule1D {
  int npts1d;                       // points per direction
  int npts;                         // npts1d^dim
  double w[kMaxPts1d];              // 1-D weights
  double h[kMaxPts1d][kMaxNgl];     // element basis at the points
  double dh[kMaxPts1d][kMaxNgl];
  double hc[kMaxPts1d][2];          // corner (linear) basis at the points
  double dhc[kMaxPts1d][2];
  uint8_t gp_t[kMaxNN][3];          // tensor index of local point g
};

struct ElemTables {
  int dim, ngl, nn, ncorner;
  int dim_w, dim_s;
  uint8_t node_t[kMaxNN][3];        // tensor index of local node a
  uint8_t corn_t[8][3];             // tensor index of local corner c
  Rule1D rule[RULE_COUNT];
};

// host helpers
int rule_1d(int kind, int n, double* x, double* w);
int lagrange_1d(int nnodes, const double* nodes, int npts, const double* pts, double* h, double* dh);
int local_order(int dim, int n, int32_t* out);
int build_elem_tables(int dim, int ngl, ElemTables* t);

}  // namespace pyn
