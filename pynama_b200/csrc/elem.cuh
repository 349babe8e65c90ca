// elem.cuh -- launchers of the element kernels and of the assembly phases.
#pragma once
#include "mesh.cuh"

namespace pyn {

// element block stores are indexed by the element's position in the batch [e0, e0+ne)
int launch_elem_kle(pyn_mesh* mesh, int64_t e0, int64_t ne, double* ebsK, double* ebsRw, double* ebsRd);
int launch_elem_ops(pyn_mesh* mesh, int64_t e0, int64_t ne, double* ebsSrT, double* ebsDiv,
                    double* ebsCurl, double* ebsW);

// symbolic phase
int build_n2e(pyn_mesh* mesh);
int build_graph(pyn_mesh* mesh, pyn::NodeGraph* g, const std::vector<int32_t>& pair_b, int ppa);
int ensure_stencil(pyn_mesh* mesh);

// One matrix to assemble from an element block store.
struct MatSpec {
  int br, bc;
  const NodeGraph* graph;          // storage graph
  const double* ebs;               // [nel][npairs_src][br*bc]
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
// per-node lumped weights: W[r] = sum over (e,a) of ebsW[e][a], ascending e
int assemble_node_weights(pyn_mesh* mesh, const double* ebsW, double* W, double* Winv);
int make_pattern(pyn_mesh* mesh, const NodeGraph* graph, int br, int bc, bool use_rowcls,
                 bool use_colcls, uint32_t pairmask, uint32_t diagmask, pyn_mat* m);

}  // namespace pyn
