// mesh.cuh -- mesh tables on the device, node graphs, block matrices.
#pragma once
#include "tables.cuh"
struct pyn_halo;
struct pyn_matfree;

namespace pyn {

// Node graph built from the element->node table and a per-element list of
// local node pairs (dense: all nn^2 pairs; stencil: pairs on a common grid
// line).  Slot = one (row node, col node) pair; every slot owns a contiguous
// segment of contributions (element, local pair id), ascending element id --
// the sorted COO of the assembly.
struct NodeGraph {
  int ppa = 0;                    // pairs per local row node
  int64_t nrows = 0;              // nodes
  int64_t nslots = 0;
  int64_t ncontrib = 0;
  int max_row = 0;                // longest row (slots)
  DevBuf<int32_t> pair_b;         // [nn*ppa] local column node of pair q = a*ppa+m
  DevBuf<int64_t> rowptr;         // [nrows+1]
  DevBuf<int32_t> colidx;         // [nslots]
  DevBuf<int64_t> cptr;           // [nslots+1]
  DevBuf<int2> csrc;              // [ncontrib] (element, local pair id)
  // the same list with the element replaced by its congruence class (= its index in the
  // element block store), rebuilt when the classes change; unused while nclass == nel
  mutable DevBuf<int2> csrc_cls;
  mutable int64_t cls_version = -1;
};

}  // namespace pyn

struct pyn_mesh {
  pyn_ctx* ctx = nullptr;
  int dim = 0, ngl = 0, nn = 0, dim_w = 0, dim_s = 0;
  int64_t nel = 0, nnodes = 0;
  pyn::ElemTables htab;                    // host copy
  pyn::DevBuf<pyn::ElemTables> dtab;       // device copy (1 element)
  pyn::DevBuf<double> corners;             // [nel][2^dim][dim]
  pyn::DevBuf<int32_t> elem_nodes;         // [nel][nn]
  pyn::DevBuf<uint8_t> dofcls;             // [nnodes*dim]
  pyn::DevBuf<uint8_t> nodecls;            // [nnodes] the dof classes of a node, 2 bits per dof
  bool has_cls = false;
  // congruence classes of element shapes: elements whose corner offsets agree to
  // congruence_tol share one set of element blocks (the reference recomputes identical
  // blocks for every congruent cell).  nclass == nel: nothing merged,
  // elem_class / class_rep unallocated.
  int64_t nclass = 0;
  int64_t class_version = 0;               // bumped by every change of the classes
  double congruence_tol = 0.0;
  pyn::DevBuf<int32_t> elem_class;         // [nel] class of the element
  pyn::DevBuf<int32_t> class_rep;          // [nclass] lowest element of the class
  // node -> (element, local index), ascending element
  // row-partitioned runs (pyn_mesh_set_global_tables): the element -> node table of the WHOLE mesh in
  // global node ids (replicated on every rank) and the global ids of the local elements; the coarse
  // level of pc asm builds identical global coarse tables on every rank from them
  std::vector<int32_t> g_elem_nodes;       // [g_nel][nn]
  std::vector<int64_t> elem_gid;           // [nel] global element of local element e
  int64_t g_nel = 0, g_nnodes = 0;
  pyn::DevBuf<int64_t> n2e_ptr;            // [nnodes+1]
  pyn::DevBuf<int2> n2e;                   // [nel*nn]
  int max_valence = 0;
  pyn::NodeGraph dense;                    // reference pattern (node graph (x) block)
  pyn::NodeGraph stencil;                  // grid-line pairs only (operators, compact)
  bool has_stencil = false;
  pyn::DevBuf<int32_t> stencil_of_dense;   // [nn*nn] stencil pair id of dense pair or -1
  double phase_ms[8] = {0, 0, 0, 0, 0, 0, 0, 0};
  pyn_matfree* mf = nullptr;               // matrix-free K data, built on first use
  ~pyn_mesh();                             // matfree.cu
};

// Node-block sparse matrix, "row-plane" BSR: the blocks of block-row r are
// stored as br*bc planes of nb_r doubles each (plane q = i*bc+j), so that a
// warp walking the blocks of a row reads every plane coalesced.  Rows assembled
// from a node graph are padded to a multiple of 4 blocks (zero blocks repeating
// the row's last column, base_slot -1), so every plane starts on a 32-byte sector:
// unaligned 256-byte runs cost B200 about half of its write bandwidth and 6 % extra
// read sectors (profiles/README.md).
struct pyn_mat {
  pyn_ctx* ctx = nullptr;
  pyn_mesh* mesh = nullptr;        // for the reference (dense) graph + dof classes
  int br = 0, bc = 0;
  int64_t nbrows = 0, nbcols = 0;  // nodes
  int64_t nblocks = 0;             // stored blocks, padding included
  int64_t nblocks_real = 0;        // without the padding blocks
  pyn::DevBuf<int64_t> rowptr;     // [nbrows+1] block offsets
  pyn::DevBuf<int32_t> colidx;     // [nblocks]
  pyn::DevBuf<double> val;         // [nblocks*br*bc] row-plane layout
  pyn::DevBuf<int64_t> base_slot;  // [nblocks] slot in the graph it was filtered from
  const pyn::NodeGraph* graph = nullptr;
  // reference-pattern rule: entry (i,j) of block (r,c) exists iff
  // pairmask bit (rowcls*4+colcls) is set, or r==c, i==j and diagmask bit rowcls is set.
  bool use_rowcls = false, use_colcls = false;
  uint32_t pairmask = 0xFFFF;
  uint32_t diagmask = 0;
  bool ref_dense = true;           // reference pattern lives on the dense graph
  // free_pattern: a matrix that does not come from a mesh graph (mesh == NULL; the immersed-
  // boundary H, S, A): its reference AIJ pattern is its own blocks, only the diagonal entries
  // of a block when diag_only (blocks w*I)
  bool free_pattern = false, diag_only = false;
  // row-partitioned run: rows [0, halo->n_own) are this rank's; x gets its ghost entries
  // refreshed before every product (comm.cu)
  pyn_halo* halo = nullptr;
  int64_t n_interior = -1;         // leading owned rows without a ghost column; -1: not counted
  int64_t own_blocks_real = -1;    // blocks (padding not counted) of the owned rows: what a product reads
  // products through the matrix-free element kernels instead of the stored blocks (matfree.cuh);
  // owned by the mesh
  pyn_matfree* mf = nullptr;
};

#define PYN_DISPATCH_BLOCK(br, bc, CALL)                        \
  do {                                                          \
    if ((br) == 3 && (bc) == 3) { CALL(3, 3); }                 \
    else if ((br) == 3 && (bc) == 1) { CALL(3, 1); }            \
    else if ((br) == 6 && (bc) == 3) { CALL(6, 3); }            \
    else if ((br) == 3 && (bc) == 6) { CALL(3, 6); }            \
    else if ((br) == 2 && (bc) == 2) { CALL(2, 2); }            \
    else if ((br) == 2 && (bc) == 1) { CALL(2, 1); }            \
    else if ((br) == 3 && (bc) == 2) { CALL(3, 2); }            \
    else if ((br) == 2 && (bc) == 3) { CALL(2, 3); }            \
    else if ((br) == 1 && (bc) == 2) { CALL(1, 2); }            \
    else if ((br) == 1 && (bc) == 1) { CALL(1, 1); }            \
    else {                                                      \
      pyn::set_error("unsupported block shape %dx%d", (br), (bc));   \
      return PYN_E_UNSUPPORTED;                                 \
    }                                                           \
  } while (0)

