// asm_pc.cu -- element-block additive Schwarz preconditioner for the Krylov solve of the KLE
// system (PETSc -pc_type asm with one subdomain per spectral element; the reference takes its
// preconditioner from the options database, kle_solver.py:49-64).
//
//   z = sum_e R_e^T W_e inv(K_e) W_e R_e r,   K_e = R_e K R_e^T,   W = diag(1/sqrt(valence))
//
// Set-up (once per operator): every element block K_e is read out of the assembled row-plane
// BSR and hashed on the device; elements with equal hashes share one class; the dense block of
// one representative per class is extracted, inverted on the host (27 blocks of 192^2 on the
// uniform 3-D box at ngl = 4) and uploaded.
// Apply (every Krylov iteration), two launches:
//   asm_mult_kernel  one CTA per tile of 16 elements of a class: gather W r into shared memory,
//                    Y = inv(K_c) X as a register-tiled fp64 product (the inverse streams from
//                    L2), element-local results to ybuf
//   asm_sum_kernel   per node: sum of the element-local results in ascending element order
//                    (fixed order: run-to-run reproducible, no atomics), optionally r.z and z.z
#include <cooperative_groups.h>

#include "asm_pc.cuh"
#include "asm_kernels.cuh"
#include "spmv.cuh"
#include "comm.cuh"

namespace pyn {
namespace cg = cooperative_groups;

// Row-partitioned set-up: a rank that has to refuse (no twin for an incomplete element, too many
// classes) must not leave the others waiting in the next collective -- all ranks fail together.
static int asm_all_ranks_ok(pyn_ctx* ctx, bool dist, bool ok_here) {
  if (!dist) return ok_here ? PYN_OK : PYN_E_UNSUPPORTED;
  DevBuf<double> flag;
  PYN_CHECK(flag.alloc(ctx, 1));
  const double mine = ok_here ? 0.0 : 1.0;
  PYN_CHECK(flag.upload(&mine, 1));
  PYN_CHECK(comm_allreduce_sum(ctx, flag.p, 1));
  double all = 0.0;
  PYN_CHECK(flag.download(&all, 1));
  if (all == 0.0) return PYN_OK;
  if (ok_here) set_error("pc asm: the set-up was refused on another rank (%d of %d ranks)", (int)all, ctx->nranks);
  return PYN_E_UNSUPPORTED;
}

template <int B>
static int asm_setup_impl(pyn_ctx* ctx, pyn_mat* A, AsmPc* pc) {
  pyn_mesh* mesh = A->mesh;
  const int nn = mesh->nn, ld = nn * B;
  const int64_t nel = mesh->nel;
  pc->mesh = mesh;
  pc->B = B;
  pc->ld = ld;
  const bool dist = A->halo != nullptr && ctx->nranks > 1;
  pc->halo = dist ? A->halo : nullptr;
  pc->n_own = dist ? A->halo->n_own : mesh->nnodes;
  // 0. valences (global ones: a partitioned run takes the ghost nodes' counts from their owners)
  DevBuf<double> valence;
  PYN_CHECK(valence.alloc(ctx, mesh->nnodes));
  asm_valence_kernel<<<grid_for(ctx, mesh->nnodes, 256), 256, 0, ctx->stream>>>(mesh->nnodes, mesh->n2e_ptr.p,
                                                                                valence.p);
  PYN_LAUNCHED(ctx);
  if (dist) PYN_CHECK(halo_exchange(A->halo, valence.p, 1));
  // 1. hash the block of every complete element (all its rows are local)
  std::vector<uint8_t> complete((size_t)nel, 1);
  std::vector<int32_t> clist;
  if (dist) {
    std::vector<int32_t> en((size_t)(nel * nn));
    PYN_CHECK(mesh->elem_nodes.download(en.data(), nel * nn));
    for (int64_t e = 0; e < nel; ++e)
      for (int a = 0; a < nn; ++a)
        if (en[(size_t)(e * nn + a)] >= pc->n_own) complete[(size_t)e] = 0;
  }
  for (int64_t e = 0; e < nel; ++e)
    if (complete[(size_t)e]) clist.push_back((int32_t)e);
  pc->ncomplete = (int64_t)clist.size();
  DevBuf<int32_t> dlist;
  PYN_CHECK(dlist.alloc(ctx, pc->ncomplete > 0 ? pc->ncomplete : 1));
  PYN_CHECK(dlist.upload(clist.data(), pc->ncomplete));
  DevBuf<unsigned long long> hash, sig;
  PYN_CHECK(hash.alloc(ctx, 2 * nel));
  const int grid = grid_for(ctx, pc->ncomplete * nn * 32, 256);
  // exact first; if that leaves more classes than the inverse store may hold, once more with the values
  // rounded to multiples of 1e-7 of the largest entry of K, so that blocks equal up to rounding share one
  // inverse (any symmetric positive definite inverse of a nearby block serves a preconditioner)
  std::vector<unsigned long long> hh((size_t)(2 * nel));
  for (int attempt = 0; attempt < 2; ++attempt) {
    double quantum = 0.0;
    if (attempt == 1) {
      DevBuf<unsigned long long> mx;
      PYN_CHECK(mx.alloc(ctx, 1));
      PYN_CHECK(mx.zero());
      asm_absmax_kernel<<<grid_for(ctx, A->val.n, 256), 256, 0, ctx->stream>>>(A->val.n, A->val.p, mx.p);
      PYN_LAUNCHED(ctx);
      unsigned long long mbits = 0ull;
      PYN_CHECK(mx.download(&mbits, 1));
      double amax;
      memcpy(&amax, &mbits, sizeof(double));
      quantum = kAsmHashQuantum * amax;
      if (!(quantum > 0.0)) break;
    }
    PYN_CHECK(hash.zero());
    asm_blocks_kernel<B, 0><<<grid, 256, 0, ctx->stream>>>(pc->ncomplete, dlist.p, mesh->elem_nodes.p, nn,
                                                           A->rowptr.p, A->colidx.p, A->val.p, hash.p, nullptr,
                                                           quantum);
    PYN_LAUNCHED(ctx);
    PYN_CHECK(hash.download(hh.data(), 2 * nel));
    pc->hash_quantum = quantum;
    if (pc->ncomplete == 0) break;
    std::vector<uint64_t> hc;
    for (int32_t e : clist) {
      hc.push_back(hh[(size_t)(2 * e)]);
      hc.push_back(hh[(size_t)(2 * e + 1)]);
    }
    std::vector<int32_t> c0, r0;
    const int64_t n0 = asm_group_by_hash((int64_t)clist.size(), hc.data(), &c0, &r0);
    if (n0 * (int64_t)ld * ld * (int64_t)sizeof(double) <= kAsmMaxInverseBytes) break;
  }
  // 2. classes of equal blocks, their representatives
  std::vector<int32_t> cls, rep;
  static_assert(sizeof(unsigned long long) == sizeof(uint64_t), "hash word");
  bool ok = true;
  if (pc->ncomplete == 0) {
    set_error("pc asm: no element of this rank has all its rows here");
    ok = false;
  } else if (pc->ncomplete == nel) {
    pc->nclass = asm_group_by_hash(nel, reinterpret_cast<const uint64_t*>(hh.data()), &cls, &rep);
  } else if (mesh->nclass != 1) {
    set_error("pc asm on a row-partitioned operator needs a mesh of congruent cells (one congruence class; this "
              "mesh has %lld); use bjacobi", (long long)mesh->nclass);
    ok = false;
  } else {
    // elements with ghost nodes: class of a complete element with the same signature
    PYN_CHECK(sig.alloc(ctx, 2 * nel));
    asm_signature_kernel<<<grid_for(ctx, nel, 256), 256, 0, ctx->stream>>>(
        nel, mesh->elem_nodes.p, nn, mesh->has_cls ? mesh->nodecls.p : nullptr, valence.p, sig.p);
    PYN_LAUNCHED(ctx);
    std::vector<unsigned long long> ss((size_t)(2 * nel));
    PYN_CHECK(sig.download(ss.data(), 2 * nel));
    pc->nclass = asm_group_partitioned(nel, complete.data(), reinterpret_cast<const uint64_t*>(hh.data()),
                                       reinterpret_cast<const uint64_t*>(ss.data()), &cls, &rep);
    if (pc->nclass < 0) {
      set_error("pc asm: element %lld touches another rank's rows and no element of this rank with all rows local "
                "has the same signature (blocks of elements too thin); use fewer ranks or bjacobi",
                (long long)(-2 - pc->nclass));
      ok = false;
    }
  }
  const int64_t inv_bytes = ok ? pc->nclass * (int64_t)ld * ld * (int64_t)sizeof(double) : 0;
  if (ok && inv_bytes > kAsmMaxInverseBytes) {
    set_error("pc asm: %lld distinct element blocks of %d^2 (%lld MB of inverses): the elements of this mesh "
              "are not congruent enough for the element-block Schwarz preconditioner; use bjacobi",
              (long long)pc->nclass, ld, (long long)(inv_bytes >> 20));
    ok = false;
  }
  PYN_CHECK(asm_all_ranks_ok(ctx, dist, ok));
  // 3. dense blocks of the representatives -> host -> inverse -> device
  DevBuf<int32_t> drep;
  PYN_CHECK(drep.alloc(ctx, pc->nclass));
  PYN_CHECK(drep.upload(rep.data(), pc->nclass));
  PYN_CHECK(pc->inv.alloc(ctx, pc->nclass * (int64_t)ld * ld));
  asm_blocks_kernel<B, 1><<<grid_for(ctx, pc->nclass * nn * 32, 256), 256, 0, ctx->stream>>>(
      pc->nclass, drep.p, mesh->elem_nodes.p, nn, A->rowptr.p, A->colidx.p, A->val.p, nullptr, pc->inv.p, 0.0);
  PYN_LAUNCHED(ctx);
  std::vector<double> blocks((size_t)(pc->nclass * (int64_t)ld * ld));
  PYN_CHECK(pc->inv.download(blocks.data(), (int64_t)blocks.size()));
  for (int64_t c = 0; c < pc->nclass; ++c) {
    double* a = blocks.data() + c * (int64_t)ld * ld;
    PYN_REQUIRE(asm_dense_inverse(ld, a) == 0, "pc asm: the block of element %d is singular", (int)rep[(size_t)c]);
    if (!asm_symmetrise(ld, a)) asm_transpose(ld, a);
  }
  PYN_CHECK(pc->inv.upload(blocks.data(), (int64_t)blocks.size()));
  // 4. tiles of elements of one class
  std::vector<int32_t> tile_elems, tile_class;
  pc->ntiles = asm_make_tiles(nel, cls.data(), pc->nclass, kAsmTile, &tile_elems, &tile_class);
  PYN_CHECK(pc->tile_elems.alloc(ctx, (int64_t)tile_elems.size()));
  PYN_CHECK(pc->tile_elems.upload(tile_elems.data(), (int64_t)tile_elems.size()));
  PYN_CHECK(pc->tile_class.alloc(ctx, pc->ntiles));
  PYN_CHECK(pc->tile_class.upload(tile_class.data(), pc->ntiles));
  // 5. weights, element-local result buffer
  PYN_CHECK(pc->wnode.alloc(ctx, mesh->nnodes));
  asm_weights_kernel<<<grid_for(ctx, mesh->nnodes, 256), 256, 0, ctx->stream>>>(mesh->nnodes, valence.p,
                                                                                pc->wnode.p);
  PYN_LAUNCHED(ctx);
  PYN_CHECK(pc->ybuf.alloc(ctx, nel * ld));
  pc->smem = sizeof(double) * ((size_t)ld * kAsmTile + (size_t)kAsmKC * kAsmRP);
  PYN_REQUIRE(pc->smem <= ctx->smem_optin, "pc asm: %zu B of shared memory per CTA needed, %zu available",
              pc->smem, ctx->smem_optin);
  PYN_CUDA(cudaFuncSetAttribute(asm_mult_kernel<B>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)pc->smem));
  PYN_CUDA(cudaStreamSynchronize(ctx->stream));     // the host vectors above leave scope
  return PYN_OK;
}

// ---- fused coarse level ---------------------------------------------------------------------------
// rc = P^T (keep r), then `steps` Chebyshev-Jacobi steps on Ac xc = rc, in one cooperative launch: one
// warp per coarse vertex, lanes over the restriction list / over the blocks of the vertex' row of Ac
// (27 on a box: one coalesced 1 944-byte run), fixed shuffle tree, grid.sync() between the steps.
// Same arithmetic as asm_restrict_kernel + asm_cheb_kernel up to the order of the row sums; r2b
// measured those at 18 + 6 x 23 us per apply (one THREAD per vertex walking its row).
struct AsmCoarseArgs {
  int64_t nv;
  int steps;
  int mode;       // 0: restriction + Chebyshev; 1: restriction only (rc = this rank's part, summed over
                  // the ranks before mode 2); 2: Chebyshev only, from rc
  const int64_t* rl_ptr;
  const int32_t* rl_node;
  const double* rl_w;
  const double* keep;
  const double* r;
  const int64_t* ptr;
  const int32_t* col;
  const double* val;
  const double* dinv;
  const double* c1;
  const double* c2;
  double* rc;
  double* xc;
  double* d0;
  double* d1;
  const double* done;
};

template <int B>
__global__ void __launch_bounds__(256)
asm_coarse_fused_kernel(AsmCoarseArgs a) {
  if (a.done && *a.done != 0.0) return;            // uniform over the grid: nobody reaches a barrier
  cg::grid_group grid = cg::this_grid();
  const int lane = threadIdx.x & 31;
  const int64_t warp = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
  const double c2_0 = a.c2[0];
  for (int64_t v = warp; v < a.nv; v += nwarps) {
    double t;
    if (a.mode != 2) {
      double acc[B];
#pragma unroll
      for (int i = 0; i < B; ++i) acc[i] = 0.0;
      for (int64_t k = a.rl_ptr[v] + lane; k < a.rl_ptr[v + 1]; k += 32) {
        const int64_t n = a.rl_node[k];
        const double w = a.rl_w[k];
#pragma unroll
        for (int i = 0; i < B; ++i) acc[i] += w * a.keep[n * B + i] * a.r[n * B + i];
      }
#pragma unroll
      for (int i = 0; i < B; ++i)
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) acc[i] += __shfl_xor_sync(0xffffffffu, acc[i], o);
      t = acc[0];
#pragma unroll
      for (int i = 1; i < B; ++i) t = lane == i ? acc[i] : t;
    } else {
      t = lane < B ? a.rc[v * B + lane] : 0.0;
    }
    if (lane < B) {
      if (a.mode != 2) a.rc[v * B + lane] = t;
      if (a.mode != 1) {
        a.xc[v * B + lane] = 0.0;
        a.d0[v * B + lane] = c2_0 * a.dinv[v * B + lane] * t;
      }
    }
  }
  if (a.mode == 1) return;                         // uniform over the grid
  double* dold = a.d0;
  double* dnew = a.d1;
  for (int s = 1; s <= a.steps; ++s) {
    grid.sync();
    const double c1 = a.c1[s], c2 = a.c2[s];
    const bool last = s == a.steps;
    for (int64_t v = warp; v < a.nv; v += nwarps) {
      double ad[B];
#pragma unroll
      for (int i = 0; i < B; ++i) ad[i] = 0.0;
      for (int64_t k = a.ptr[v] + lane; k < a.ptr[v + 1]; k += 32) {
        const double* blk = a.val + k * B * B;
        const double* d = dold + (int64_t)a.col[k] * B;
        double dv[B];
#pragma unroll
        for (int j = 0; j < B; ++j) dv[j] = __ldcg(d + j);          // written by other SMs before the barrier
#pragma unroll
        for (int i = 0; i < B; ++i)
#pragma unroll
          for (int j = 0; j < B; ++j) ad[i] += blk[i * B + j] * dv[j];
      }
#pragma unroll
      for (int i = 0; i < B; ++i)
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) ad[i] += __shfl_xor_sync(0xffffffffu, ad[i], o);
      if (lane < B) {
        double t = ad[0];
#pragma unroll
        for (int i = 1; i < B; ++i) t = lane == i ? ad[i] : t;
        const int64_t q = v * B + lane;
        const double dvv = __ldcg(dold + q);
        a.xc[q] += dvv;
        if (!last) {
          const double rn = a.rc[q] - t;
          a.rc[q] = rn;
          dnew[q] = c1 * dvv + c2 * a.dinv[q] * rn;
        }
      }
    }
    double* tmp = dold;
    dold = dnew;
    dnew = tmp;
  }
}

// mode / steps as in AsmCoarseArgs; r: the fine vector restricted (modes 0, 1)
template <int B>
static int asm_coarse_launch(pyn_ctx* ctx, AsmPc* pc, int mode, int steps, const double* r, const double* done) {
  AsmCoarseArgs ca;
  ca.nv = pc->nv;
  ca.steps = steps;
  ca.mode = mode;
  ca.rl_ptr = pc->rl_ptr.p;
  ca.rl_node = pc->rl_node.p;
  ca.rl_w = pc->rl_w.p;
  ca.keep = pc->keep.p;
  ca.r = r;
  ca.ptr = pc->ac_ptr.p;
  ca.col = pc->ac_col.p;
  ca.val = pc->ac_val.p;
  ca.dinv = pc->ac_dinv.p;
  ca.c1 = pc->cheb_c.p;
  ca.c2 = pc->cheb_c.p + (pc->coarse_steps + 1);
  ca.rc = pc->rc.p;
  ca.xc = pc->xc.p;
  ca.d0 = pc->d0.p;
  ca.d1 = pc->d1.p;
  ca.done = done;
  void* args[] = {&ca};
  PYN_CUDA(cudaLaunchCooperativeKernel((const void*)asm_coarse_fused_kernel<B>, dim3(pc->coarse_grid), dim3(256),
                                       args, 0, ctx->stream));
  PYN_LAUNCHED(ctx);
  return PYN_OK;
}

// rc = P^T (keep r) over ALL ranks through the restriction lists (set-up probing; the apply fuses it)
template <int B>
static int asm_restrict_lists(pyn_ctx* ctx, AsmPc* pc, const double* r) {
  PYN_CHECK(asm_coarse_launch<B>(ctx, pc, 1, 0, r, nullptr));
  if (pc->halo) PYN_CHECK(comm_allreduce_sum(ctx, pc->rc.p, (int)(pc->nv * B)));
  return PYN_OK;
}

template <int B>
static AsmProlong asm_prolong_args(const AsmPc* pc, const double* xc) {
  AsmProlong cp;
  cp.xc = xc;
  cp.elem_verts = pc->elem_verts.p;
  cp.phi = pc->phi.p;
  cp.keep = pc->keep.p;
  cp.ncorner = pc->ncorner;
  return cp;
}

template <int B>
static int asm_restrict(pyn_ctx* ctx, AsmPc* pc, const double* r, double* rc, const double* done) {
  pyn_mesh* mesh = pc->mesh;
  asm_restrict_kernel<B><<<grid_for(ctx, pc->nv * 32, 256), 256, 0, ctx->stream>>>(
      pc->nv, pc->v2e_ptr.p, pc->v2e.p, mesh->elem_nodes.p, mesh->nn, pc->ncorner, pc->phi.p, mesh->n2e_ptr.p,
      mesh->n2e.p, pc->keep.p, r, rc, done);
  PYN_LAUNCHED(ctx);
  return PYN_OK;
}

// Second level: tables, the free-dof mask, the Galerkin operator by coloured probing (one product
// with K per colour and component: same-coloured coarse functions never meet in a coarse row),
// its Jacobi diagonal, the Chebyshev interval.
template <int B>
static int asm_coarse_setup(pyn_ctx* ctx, pyn_mat* A, AsmPc* pc) {
  pyn_mesh* mesh = pc->mesh;
  const int nn = mesh->nn, ncorner = mesh->htab.ncorner;
  const int64_t nel = mesh->nel, nnodes = mesh->nnodes;
  pc->ncorner = ncorner;
  int corner_local[8];
  std::vector<double> phi;
  PYN_REQUIRE(asm_coarse_phi(mesh->htab, corner_local, &phi) == 0, "pc asm: element corner nodes not found");
  std::vector<int32_t> en((size_t)(nel * nn));
  PYN_CHECK(mesh->elem_nodes.download(en.data(), nel * nn));
  const bool dist = pc->halo != nullptr;
  PYN_REQUIRE(!dist || mesh->g_nel > 0,
              "pc asm: the coarse level on a row-partitioned operator needs pyn_mesh_set_global_tables");
  // the coarse space lives on the element corners of the WHOLE mesh: on several ranks the tables are
  // built from the global element table, identically everywhere (vertex ids, adjacency, colours)
  AsmCoarseTables T;
  if (dist)
    asm_coarse_tables(mesh->g_nel, mesh->g_nnodes, nn, ncorner, mesh->g_elem_nodes.data(), corner_local, &T);
  else
    asm_coarse_tables(nel, nnodes, nn, ncorner, en.data(), corner_local, &T);
  const int64_t nv = pc->nv = T.nv;
  pc->ncolors = T.ncolors;
  PYN_CHECK(pc->phi.alloc(ctx, (int64_t)phi.size()));
  PYN_CHECK(pc->phi.upload(phi.data(), (int64_t)phi.size()));
  std::vector<int32_t> ev_local((size_t)(nel * ncorner));              // coarse vertex of every local element corner
  for (int64_t e = 0; e < nel; ++e) {
    const int64_t ge = dist ? mesh->elem_gid[(size_t)e] : e;
    for (int c = 0; c < ncorner; ++c) ev_local[(size_t)(e * ncorner + c)] = T.elem_verts[(size_t)(ge * ncorner + c)];
  }
  PYN_CHECK(pc->elem_verts.alloc(ctx, nel * ncorner));
  PYN_CHECK(pc->elem_verts.upload(ev_local.data(), nel * ncorner));
  static_assert(sizeof(int2) == 2 * sizeof(int32_t), "int2 layout");
  if (!dist) {                                  // (element, corner) lists of the separate-launch path
    PYN_CHECK(pc->v2e_ptr.alloc(ctx, nv + 1));
    PYN_CHECK(pc->v2e_ptr.upload(T.v2e_ptr.data(), nv + 1));
    const int64_t nv2e = T.v2e_ptr[(size_t)nv];
    PYN_CHECK(pc->v2e.alloc(ctx, nv2e));
    PYN_CHECK(pc->v2e.upload(reinterpret_cast<const int2*>(T.v2e.data()), nv2e));
  }
  // restriction lists: an OWNED fine node contributes, through its first local element, to the
  // vertices of that element with a non-zero corner function (the weights do not depend on the
  // element: the mesh is conforming); local elements and local nodes ascending inside every list
  {
    std::vector<int64_t> first((size_t)nnodes, -1);
    for (int64_t k = nel * nn - 1; k >= 0; --k) first[(size_t)en[(size_t)k]] = k;     // lowest (e, a) wins
    std::vector<int64_t> rl_ptr((size_t)nv + 1, 0);
    for (int pass = 0; pass < 2; ++pass) {
      std::vector<int64_t> fill(rl_ptr.begin(), rl_ptr.end() - 1);
      std::vector<int32_t> rl_node(pass ? (size_t)rl_ptr[(size_t)nv] + 1 : 1);
      std::vector<double> rl_w(pass ? (size_t)rl_ptr[(size_t)nv] + 1 : 1);
      for (int64_t e = 0; e < nel; ++e)
        for (int a = 0; a < nn; ++a) {
          const int64_t n = en[(size_t)(e * nn + a)];
          if (n >= pc->n_own || first[(size_t)n] != e * nn + a) continue;
          for (int c = 0; c < ncorner; ++c) {
            const double w = phi[(size_t)(a * ncorner + c)];
            if (w == 0.0) continue;
            const int64_t v = ev_local[(size_t)(e * ncorner + c)];
            if (!pass) {
              rl_ptr[(size_t)v + 1]++;
            } else {
              rl_node[(size_t)fill[(size_t)v]] = (int32_t)n;
              rl_w[(size_t)fill[(size_t)v]++] = w;
            }
          }
        }
      if (!pass) {
        for (int64_t v = 0; v < nv; ++v) rl_ptr[(size_t)v + 1] += rl_ptr[(size_t)v];
      } else {
        PYN_CHECK(pc->rl_ptr.alloc(ctx, nv + 1));
        PYN_CHECK(pc->rl_ptr.upload(rl_ptr.data(), nv + 1));
        PYN_CHECK(pc->rl_node.alloc(ctx, (int64_t)rl_node.size()));
        PYN_CHECK(pc->rl_node.upload(rl_node.data(), (int64_t)rl_node.size()));
        PYN_CHECK(pc->rl_w.alloc(ctx, (int64_t)rl_w.size()));
        PYN_CHECK(pc->rl_w.upload(rl_w.data(), (int64_t)rl_w.size()));
      }
    }
    int per_sm = 0;
    PYN_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, asm_coarse_fused_kernel<B>, 256, 0));
    const int64_t want = (nv * 32 + 255) / 256;
    // (r2d: one CTA per SM instead of the full co-resident grid made the apply ~90 us slower: a warp
    // then walks four vertices one after the other; the barriers are the smaller cost)
    const int64_t cap = (int64_t)per_sm * ctx->num_sms;
    pc->coarse_grid = (int)std::max<int64_t>(1, std::min(want, cap));
    // placeholders until the coefficients are known (the probing below restricts with steps = 0)
    std::vector<double> cc0((size_t)(2 * (pc->coarse_steps + 1)), 0.0);
    PYN_CHECK(pc->cheb_c.alloc(ctx, (int64_t)cc0.size()));
    PYN_CHECK(pc->cheb_c.upload(cc0.data(), (int64_t)cc0.size()));
  }
  PYN_CHECK(pc->keep.alloc(ctx, nnodes * B));
  asm_free_dofs_kernel<B><<<grid_for(ctx, nnodes * 32, 256), 256, 0, ctx->stream>>>(nnodes, A->rowptr.p, A->colidx.p,
                                                                                    A->val.p, pc->keep.p);
  PYN_LAUNCHED(ctx);
  PYN_CHECK(pc->rc.alloc(ctx, nv * B));
  PYN_CHECK(pc->xc.alloc(ctx, nv * B));
  PYN_CHECK(pc->d0.alloc(ctx, nv * B));
  PYN_CHECK(pc->d1.alloc(ctx, nv * B));
  // probing
  const int64_t nnzb = T.adj_ptr[(size_t)nv];
  std::vector<double> val((size_t)(nnzb * B * B), 0.0), xh((size_t)(nv * B)), th((size_t)(nv * B));
  DevBuf<double> xf, yf;
  PYN_CHECK(xf.alloc(ctx, nnodes * B));
  PYN_CHECK(yf.alloc(ctx, nnodes * B));
  const int gridn = grid_for(ctx, nnodes, 256);
  for (int color = 0; color < T.ncolors; ++color)
    for (int j = 0; j < B; ++j) {
      for (int64_t v = 0; v < nv; ++v)
        for (int i = 0; i < B; ++i) xh[(size_t)(v * B + i)] = (T.color[(size_t)v] == color && i == j) ? 1.0 : 0.0;
      PYN_CHECK(pc->xc.upload(xh.data(), nv * B));
      asm_prolong_kernel<B><<<gridn, 256, 0, ctx->stream>>>(nnodes, mesh->n2e_ptr.p, mesh->n2e.p,
                                                            asm_prolong_args<B>(pc, pc->xc.p), xf.p);
      PYN_LAUNCHED(ctx);
      PYN_CHECK(spmv(A, 1.0, xf.p, 0.0, nullptr, yf.p));          // (row-partitioned: refreshes the ghosts of xf)
      PYN_CHECK(asm_restrict_lists<B>(ctx, pc, yf.p));
      PYN_CHECK(pc->rc.download(th.data(), nv * B));
      for (int64_t v = 0; v < nv; ++v)
        for (int64_t k = T.adj_ptr[(size_t)v]; k < T.adj_ptr[(size_t)v + 1]; ++k)
          if (T.color[(size_t)T.adj[(size_t)k]] == color)
            for (int i = 0; i < B; ++i) val[(size_t)((k * B + i) * B + j)] = th[(size_t)(v * B + i)];
    }
  asm_coarse_symmetrise(nv, B, T.adj_ptr.data(), T.adj.data(), val.data());
  std::vector<double> dinv((size_t)(nv * B), 0.0);
  asm_coarse_jacobi(nv, B, T.adj_ptr.data(), T.adj.data(), val.data(), dinv.data());
  pc->lmax = 1.1 * asm_power_lmax(nv, B, T.adj_ptr.data(), T.adj.data(), val.data(), dinv.data(), 40);
  PYN_REQUIRE(pc->lmax > 0.0 && std::isfinite(pc->lmax), "pc asm: the coarse operator is empty or not finite");
  asm_cheb_coefficients(pc->coarse_steps, 0.1 * pc->lmax, pc->lmax, &pc->c1, &pc->c2);
  PYN_CHECK(pc->ac_ptr.alloc(ctx, nv + 1));
  PYN_CHECK(pc->ac_ptr.upload(T.adj_ptr.data(), nv + 1));
  PYN_CHECK(pc->ac_col.alloc(ctx, nnzb));
  PYN_CHECK(pc->ac_col.upload(T.adj.data(), nnzb));
  PYN_CHECK(pc->ac_val.alloc(ctx, nnzb * B * B));
  PYN_CHECK(pc->ac_val.upload(val.data(), nnzb * B * B));
  PYN_CHECK(pc->ac_dinv.alloc(ctx, nv * B));
  PYN_CHECK(pc->ac_dinv.upload(dinv.data(), nv * B));
  {
    std::vector<double> cc;
    cc.insert(cc.end(), pc->c1.begin(), pc->c1.end());
    cc.insert(cc.end(), pc->c2.begin(), pc->c2.end());
    PYN_CHECK(pc->cheb_c.upload(cc.data(), (int64_t)cc.size()));
  }
  PYN_CUDA(cudaStreamSynchronize(ctx->stream));
  return PYN_OK;
}

int asm_setup(pyn_ctx* ctx, pyn_mat* A, int coarse_steps, AsmPc** out) {
  *out = nullptr;
  PYN_REQUIRE(A->mesh && !A->free_pattern, "pc asm needs an operator assembled from a mesh");
  const bool dist = A->halo && ctx->nranks > 1;
  PYN_REQUIRE(A->br == A->bc && A->nbrows == A->mesh->nnodes, "pc asm: operator is not square over the mesh nodes");
  AsmPc* pc = new AsmPc();
  pc->coarse_steps = coarse_steps > 0 ? coarse_steps : 0;
  int rc;
  if (A->br == 3) {
    rc = asm_setup_impl<3>(ctx, A, pc);
    if (rc == PYN_OK && pc->coarse_steps > 0) rc = asm_coarse_setup<3>(ctx, A, pc);
  } else if (A->br == 2) {
    rc = asm_setup_impl<2>(ctx, A, pc);
    if (rc == PYN_OK && pc->coarse_steps > 0) rc = asm_coarse_setup<2>(ctx, A, pc);
  } else {
    set_error("pc asm: unsupported block size %d", A->br);
    rc = PYN_E_UNSUPPORTED;
  }
  if (rc != PYN_OK) {
    delete pc;
    return rc;
  }
  *out = pc;
  return PYN_OK;
}

template <int B>
static int asm_apply_impl(pyn_ctx* ctx, AsmPc* pc, double* r, double* z, double* dots, const double* done) {
  pyn_mesh* mesh = pc->mesh;
  if (pc->halo) PYN_CHECK(halo_exchange(pc->halo, r, B));     // the elements of the halo layer read ghost entries
  const int64_t cap = (int64_t)ctx->num_sms * 2;
  const int gridm = (int)(pc->ntiles < cap ? pc->ntiles : cap);
  asm_mult_kernel<B><<<gridm > 0 ? gridm : 1, 256, pc->smem, ctx->stream>>>(
      pc->ntiles, pc->tile_elems.p, pc->tile_class.p, mesh->elem_nodes.p, mesh->nn, pc->inv.p, pc->wnode.p, r,
      pc->ybuf.p, done);
  PYN_LAUNCHED(ctx);
  const double* xc = nullptr;
  const bool unfused = g_asm_coarse_unfused;     // pyn_tune("asm_coarse_unfused"): the seven-launch path (r2a)
  if (pc->coarse_steps > 0 && pc->halo) {
    // row-partitioned: the coarse problem is replicated.  This rank's part of P^T r, summed over the
    // ranks (nv * B doubles), then every rank takes the same Chebyshev steps on the whole coarse problem
    PYN_CHECK(asm_coarse_launch<B>(ctx, pc, 1, 0, r, done));
    PYN_CHECK(comm_allreduce_sum(ctx, pc->rc.p, (int)(pc->nv * B)));
    PYN_CHECK(asm_coarse_launch<B>(ctx, pc, 2, pc->coarse_steps, r, done));
    xc = pc->xc.p;
  } else if (pc->coarse_steps > 0 && !unfused) {     // restriction + all Chebyshev steps in ONE cooperative launch
    PYN_CHECK(asm_coarse_launch<B>(ctx, pc, 0, pc->coarse_steps, r, done));
    xc = pc->xc.p;
  } else if (pc->coarse_steps > 0) {           // xc ~ Ac^-1 P^T r by coarse_steps Chebyshev steps
    PYN_CHECK(asm_restrict<B>(ctx, pc, r, pc->rc.p, done));
    const int gridc = grid_for(ctx, pc->nv, 256);
    double* dold = pc->d0.p;
    double* dnew = pc->d1.p;
    for (int s = 0; s <= pc->coarse_steps; ++s) {
      asm_cheb_kernel<B><<<gridc, 256, 0, ctx->stream>>>(pc->nv, pc->ac_ptr.p, pc->ac_col.p, pc->ac_val.p,
                                                         pc->ac_dinv.p, s == 0 ? 1 : 0, pc->c1[(size_t)s],
                                                         pc->c2[(size_t)s], pc->rc.p, pc->xc.p, dold, dnew, done);
      PYN_LAUNCHED(ctx);
      std::swap(dold, dnew);
    }
    xc = pc->xc.p;
  }
  asm_sum_kernel<B><<<grid_for(ctx, pc->n_own, 256), 256, 0, ctx->stream>>>(
      pc->n_own, mesh->n2e_ptr.p, mesh->n2e.p, pc->ybuf.p, pc->ld, r, z, ctx->red_partials, ctx->red_ticket,
      dots, done, asm_prolong_args<B>(pc, xc));
  PYN_LAUNCHED(ctx);
  return PYN_OK;
}

int asm_apply(pyn_ctx* ctx, AsmPc* pc, double* r, double* z, double* dots, const double* done) {
  if (pc->B == 3) return asm_apply_impl<3>(ctx, pc, r, z, dots, done);
  return asm_apply_impl<2>(ctx, pc, r, z, dots, done);
}

}  // namespace pyn
