// elem.cuh -- launchers of the element kernels and of the assembly phases.
#pragma once
#include "mesh.cuh"

namespace pyn {

// Record of one node pair in the element block store of the KLE kernel:
// two halves of whole 32-byte sectors, [K_e block (dim x dim) | pad] [Rw_e block (dim x dim_w) |
// Rd_e block (dim) | pad]: 12 + 12 doubles in 3-D (9+3 | 9+3), 4 + 4 in 2-D (4 | 2+2), so the
// gather can sum K and then Rw, Rd with half the registers
template <int DIM>
struct KleRecord {
  static constexpr int DW = DIM == 2 ? 1 : 3;
  static constexpr int HALF = (DIM * DIM + 3) / 4 * 4;
  static constexpr int OFF_K = 0, OFF_RW = HALF, OFF_RD = HALF + DIM * DW;
  static constexpr int RS = 2 * HALF;
  static_assert(DIM * DW + DIM <= HALF, "Rw and Rd must fit the second half");
};
inline int kle_record_stride(int dim) { return dim == 2 ? KleRecord<2>::RS : KleRecord<3>::RS; }
inline int kle_record_off_rw(int dim) { return dim == 2 ? KleRecord<2>::OFF_RW : KleRecord<3>::OFF_RW; }
inline int kle_record_off_rd(int dim) { return dim == 2 ? KleRecord<2>::OFF_RD : KleRecord<3>::OFF_RD; }

// element block stores are indexed by the element's position in the batch: element
// e0 + i, or rep[i] when a list of representative elements is given (rep != NULL)
// ebs: [ne][nn*nn] records of kle_record_stride(dim) doubles
int launch_elem_kle(pyn_mesh* mesh, int64_t e0, int64_t ne, const int32_t* rep, double* ebs);
int launch_elem_ops(pyn_mesh* mesh, int64_t e0, int64_t ne, const int32_t* rep, double* ebsSrT,
                    double* ebsDiv, double* ebsCurl, double* ebsW);

// symbolic phase
int build_n2e(pyn_mesh* mesh);
int build_graph(pyn_mesh* mesh, pyn::NodeGraph* g, const std::vector<int32_t>& pair_b, int ppa);
int ensure_stencil(pyn_mesh* mesh);

// One matrix to assemble from an element block store.
struct MatSpec {
  int br, bc;
  const NodeGraph* graph;          // storage graph
  const double* ebs;               // [nclass][npairs_src] records of ebs_stride doubles, block at ebs_off
  int ebs_stride = 0, ebs_off = 0; // 0: records are the blocks themselves (stride br*bc)
  int64_t ebs_pairs;               // pairs per element in the store
  int ebs_ppa = 0;                 // > 0: plane-major store [e][a][plane][m], pair = a*ebs_ppa + m; 0: block-major
  const int32_t* pair_map;         // graph pair id -> store pair id (or -1), NULL = identity
  double sign;
  bool use_rowcls, use_colcls;     // velocity dof classes on rows / columns
  uint32_t pairmask;               // allowed (rowcls*4+colcls) combinations
  uint32_t diagmask;               // classes whose diagonal gets diag_add
  double diag_add[4];
  const double* row_scale;         // per row NODE scale applied after the sum, or NULL
  bool ref_dense;
};
int assemble_matrix(pyn_mesh* mesh, const MatSpec& spec, pyn_mat** out);
// contribution list of a graph as (index in the element block store, local pair id)
int class_sources(pyn_mesh* mesh, const NodeGraph* g, const int2** out);
// K, Krhs, Rw, Rd (specs and results in this order) from one pass over the dense graph
int assemble_kle_fused(pyn_mesh* mesh, const MatSpec* specs4, pyn_mat** out4);
// per-node lumped weights: W[r] = sum over (e,a) of ebsW[class(e)][a], ascending e
int assemble_node_weights(pyn_mesh* mesh, const double* ebsW, double* W, double* Winv);
int make_pattern(pyn_mesh* mesh, const NodeGraph* graph, int br, int bc, bool use_rowcls,
                 bool use_colcls, uint32_t pairmask, uint32_t diagmask, pyn_mat* m);

}  // namespace pyn
