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
#include "asm_pc.cuh"
#include "asm_kernels.cuh"

namespace pyn {

template <int B>
static int asm_setup_impl(pyn_ctx* ctx, const pyn_mat* A, AsmPc* pc) {
  pyn_mesh* mesh = A->mesh;
  const int nn = mesh->nn, ld = nn * B;
  const int64_t nel = mesh->nel;
  pc->mesh = mesh;
  pc->B = B;
  pc->ld = ld;
  // 1. hash every element block
  DevBuf<unsigned long long> hash;
  PYN_CHECK(hash.alloc(ctx, 2 * nel));
  PYN_CHECK(hash.zero());
  const int grid = grid_for(ctx, nel * nn * 32, 256);
  asm_blocks_kernel<B, 0><<<grid, 256, 0, ctx->stream>>>(nel, nullptr, mesh->elem_nodes.p, nn, A->rowptr.p,
                                                         A->colidx.p, A->val.p, hash.p, nullptr);
  PYN_LAUNCHED(ctx);
  std::vector<unsigned long long> hh((size_t)(2 * nel));
  PYN_CHECK(hash.download(hh.data(), 2 * nel));
  // 2. classes of equal blocks, their representatives
  std::vector<int32_t> cls, rep;
  static_assert(sizeof(unsigned long long) == sizeof(uint64_t), "hash word");
  pc->nclass = asm_group_by_hash(nel, reinterpret_cast<const uint64_t*>(hh.data()), &cls, &rep);
  const int64_t inv_bytes = pc->nclass * (int64_t)ld * ld * (int64_t)sizeof(double);
  PYN_REQUIRE(inv_bytes <= kAsmMaxInverseBytes,
              "pc asm: %lld distinct element blocks of %d^2 (%lld MB of inverses): the elements of this mesh "
              "are not congruent enough for the element-block Schwarz preconditioner; use bjacobi",
              (long long)pc->nclass, ld, (long long)(inv_bytes >> 20));
  // 3. dense blocks of the representatives -> host -> inverse -> device
  DevBuf<int32_t> drep;
  PYN_CHECK(drep.alloc(ctx, pc->nclass));
  PYN_CHECK(drep.upload(rep.data(), pc->nclass));
  PYN_CHECK(pc->inv.alloc(ctx, pc->nclass * (int64_t)ld * ld));
  asm_blocks_kernel<B, 1><<<grid_for(ctx, pc->nclass * nn * 32, 256), 256, 0, ctx->stream>>>(
      pc->nclass, drep.p, mesh->elem_nodes.p, nn, A->rowptr.p, A->colidx.p, A->val.p, nullptr, pc->inv.p);
  PYN_LAUNCHED(ctx);
  std::vector<double> blocks((size_t)(pc->nclass * (int64_t)ld * ld));
  PYN_CHECK(pc->inv.download(blocks.data(), (int64_t)blocks.size()));
  for (int64_t c = 0; c < pc->nclass; ++c) {
    double* a = blocks.data() + c * (int64_t)ld * ld;
    PYN_REQUIRE(asm_dense_inverse(ld, a) == 0, "pc asm: the block of element %d is singular", (int)rep[(size_t)c]);
    asm_symmetrise(ld, a);
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
  asm_weights_kernel<<<grid_for(ctx, mesh->nnodes, 256), 256, 0, ctx->stream>>>(mesh->nnodes, mesh->n2e_ptr.p,
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

int asm_setup(pyn_ctx* ctx, const pyn_mat* A, AsmPc** out) {
  *out = nullptr;
  PYN_REQUIRE(A->mesh && !A->free_pattern, "pc asm needs an operator assembled from a mesh");
  PYN_REQUIRE(A->br == A->bc && A->nbrows == A->mesh->nnodes, "pc asm: operator is not square over the mesh nodes");
  if (A->halo && ctx->nranks > 1) {
    set_error("pc asm is not available on row-partitioned operators yet (single GPU only); use bjacobi");
    return PYN_E_UNSUPPORTED;
  }
  AsmPc* pc = new AsmPc();
  int rc;
  if (A->br == 3) rc = asm_setup_impl<3>(ctx, A, pc);
  else if (A->br == 2) rc = asm_setup_impl<2>(ctx, A, pc);
  else {
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
static int asm_apply_impl(pyn_ctx* ctx, AsmPc* pc, const double* r, double* z, double* dots, const double* done) {
  pyn_mesh* mesh = pc->mesh;
  const int64_t cap = (int64_t)ctx->num_sms * 2;
  const int gridm = (int)(pc->ntiles < cap ? pc->ntiles : cap);
  asm_mult_kernel<B><<<gridm > 0 ? gridm : 1, 256, pc->smem, ctx->stream>>>(
      pc->ntiles, pc->tile_elems.p, pc->tile_class.p, mesh->elem_nodes.p, mesh->nn, pc->inv.p, pc->wnode.p, r,
      pc->ybuf.p, done);
  PYN_LAUNCHED(ctx);
  asm_sum_kernel<B><<<grid_for(ctx, mesh->nnodes, 256), 256, 0, ctx->stream>>>(
      mesh->nnodes, mesh->n2e_ptr.p, mesh->n2e.p, pc->ybuf.p, pc->ld, r, z, ctx->red_partials, ctx->red_ticket,
      dots, done);
  PYN_LAUNCHED(ctx);
  return PYN_OK;
}

int asm_apply(pyn_ctx* ctx, AsmPc* pc, const double* r, double* z, double* dots, const double* done) {
  if (pc->B == 3) return asm_apply_impl<3>(ctx, pc, r, z, dots, done);
  return asm_apply_impl<2>(ctx, pc, r, z, dots, done);
}

}  // namespace pyn
