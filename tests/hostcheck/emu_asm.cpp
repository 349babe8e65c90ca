// emu_asm.cpp -- TEST-ONLY: the kernels of the element-block Schwarz preconditioner
// (pynama_b200/csrc/asm_kernels.cuh) run on CPU threads through cuda_emu.h, driven the way
// asm_pc.cu drives them on the device (hash -> group -> extract -> invert -> tiles -> weights;
// then product + node sum).  Built with g++ (not nvcc) into libemu_asm.so by
// __graft_entry__.build(); loaded only by tests/test_asm_emulated.py.
#include "cuda_emu.h"

#include "../../pynama_b200/csrc/asm_kernels.cuh"

using namespace pyn;

namespace {

template <int B>
int run(int nn, long long nel, long long nnodes, const int32_t* elem_nodes, const int64_t* rowptr,
        const int32_t* colidx, const double* val, const int64_t* n2e_ptr, const int2* n2e, const double* r,
        int grid, int napply, double* z, double* dots, long long* nclass_out, int32_t* cls_out, double* inv_out,
        int dim, int ngl, int coarse_steps, long long* coarse_info, double* ac_dense, long long n_own,
        const double* valence_global, const uint8_t* nodecls, double quantum_rel) {
  const int ld = nn * B;
  // set-up, as asm_setup_impl
  std::vector<double> valence((size_t)nnodes, -1.0);
  cuda_emu::launch(grid, 256, 0, [&]() { asm_valence_kernel(nnodes, n2e_ptr, valence.data()); });
  if (valence_global)                           // stands for halo_exchange(valence): ghost counts from the owners
    for (long long n = n_own; n < nnodes; ++n) valence[(size_t)n] = valence_global[n];
  std::vector<uint8_t> complete((size_t)nel, 1);
  std::vector<int32_t> clist;
  for (long long e = 0; e < nel; ++e) {
    for (int a = 0; a < nn; ++a)
      if (elem_nodes[e * nn + a] >= n_own) complete[(size_t)e] = 0;
    if (complete[(size_t)e]) clist.push_back((int32_t)e);
  }
  if (clist.empty()) return -8;
  double quantum = 0.0;
  if (quantum_rel > 0.0) {                      // the set-up's second grouping attempt
    unsigned long long mbits = 0ull;
    const int64_t nval = rowptr[nnodes] * B * B;
    cuda_emu::launch(grid, 256, 0, [&]() { asm_absmax_kernel(nval, val, &mbits); });
    double amax;
    memcpy(&amax, &mbits, sizeof(double));
    quantum = quantum_rel * amax;
  }
  std::vector<unsigned long long> hash((size_t)(2 * nel), 0ull);
  cuda_emu::launch(grid, 256, 0, [&]() {
    asm_blocks_kernel<B, 0>((int64_t)clist.size(), clist.data(), elem_nodes, nn, rowptr, colidx, val, hash.data(),
                            nullptr, quantum);
  });
  std::vector<int32_t> cls, rep;
  int64_t nclass;
  if ((long long)clist.size() == nel) {
    nclass = asm_group_by_hash(nel, reinterpret_cast<const uint64_t*>(hash.data()), &cls, &rep);
  } else {
    std::vector<unsigned long long> sig((size_t)(2 * nel), 0ull);
    cuda_emu::launch(grid, 256, 0, [&]() {
      asm_signature_kernel(nel, elem_nodes, nn, nodecls, valence.data(), sig.data());
    });
    nclass = asm_group_partitioned(nel, complete.data(), reinterpret_cast<const uint64_t*>(hash.data()),
                                   reinterpret_cast<const uint64_t*>(sig.data()), &cls, &rep);
    if (nclass < 0) return -10;
  }
  *nclass_out = nclass;
  for (long long e = 0; e < nel; ++e) cls_out[e] = cls[(size_t)e];
  std::vector<double> inv((size_t)(nclass * (int64_t)ld * ld), -777.0);
  cuda_emu::launch(grid, 256, 0, [&]() {
    asm_blocks_kernel<B, 1>(nclass, rep.data(), elem_nodes, nn, rowptr, colidx, val, nullptr, inv.data(), 0.0);
  });
  for (int64_t c = 0; c < nclass; ++c) {
    double* a = inv.data() + c * (int64_t)ld * ld;
    if (asm_dense_inverse(ld, a) != 0) return -2;
    if (!asm_symmetrise(ld, a)) asm_transpose(ld, a);
  }
  if (inv_out) memcpy(inv_out, inv.data(), sizeof(double) * inv.size());
  std::vector<int32_t> tile_elems, tile_class;
  const int64_t ntiles = asm_make_tiles(nel, cls.data(), nclass, kAsmTile, &tile_elems, &tile_class);
  std::vector<double> wnode((size_t)nnodes, -1.0);
  cuda_emu::launch(grid, 256, 0, [&]() { asm_weights_kernel(nnodes, valence.data(), wnode.data()); });
  std::vector<double> ybuf((size_t)(nel * ld), -777.0);
  const size_t smem = sizeof(double) * ((size_t)ld * kAsmTile + (size_t)kAsmKC * kAsmRP);
  // second level, as asm_coarse_setup (the product with K: plain loops over the row-plane BSR)
  AsmCoarseTables T;
  std::vector<double> phi, keep, acval, dinv, c1, c2, rc, xc, d0, d1;
  int ncorner = 0;
  std::vector<int2> v2e;
  if (coarse_steps > 0) {
    ElemTables tab;
    if (build_elem_tables(dim, ngl, &tab) != 0 || tab.nn != nn) return -5;
    ncorner = tab.ncorner;
    int corner_local[8];
    if (asm_coarse_phi(tab, corner_local, &phi) != 0) return -6;
    asm_coarse_tables(nel, nnodes, nn, ncorner, elem_nodes, corner_local, &T);
    const int64_t nv = T.nv;
    v2e.resize((size_t)T.v2e_ptr[(size_t)nv]);
    memcpy(v2e.data(), T.v2e.data(), sizeof(int2) * v2e.size());
    keep.assign((size_t)(nnodes * B), -1.0);
    cuda_emu::launch(grid, 256, 0, [&]() { asm_free_dofs_kernel<B>(nnodes, rowptr, colidx, val, keep.data()); });
    rc.assign((size_t)(nv * B), 0.0);
    xc = d0 = d1 = rc;
    const int64_t nnzb = T.adj_ptr[(size_t)nv];
    acval.assign((size_t)(nnzb * B * B), 0.0);
    std::vector<double> xf((size_t)(nnodes * B)), yf((size_t)(nnodes * B)), xh((size_t)(nv * B));
    AsmProlong cp{nullptr, T.elem_verts.data(), phi.data(), keep.data(), ncorner};
    for (int color = 0; color < T.ncolors; ++color)
      for (int j = 0; j < B; ++j) {
        for (int64_t v = 0; v < nv; ++v)
          for (int i = 0; i < B; ++i) xh[(size_t)(v * B + i)] = (T.color[(size_t)v] == color && i == j) ? 1.0 : 0.0;
        cp.xc = xh.data();
        cuda_emu::launch(grid, 256, 0, [&]() { asm_prolong_kernel<B>(nnodes, n2e_ptr, n2e, cp, xf.data()); });
        for (int64_t n = 0; n < nnodes; ++n) {           // yf = K xf
          const int64_t p0 = rowptr[n];
          const int64_t nb = rowptr[n + 1] - p0;
          for (int i = 0; i < B; ++i) {
            double acc = 0.0;
            for (int64_t u = 0; u < nb; ++u)
              for (int jj = 0; jj < B; ++jj)
                acc += val[p0 * B * B + (int64_t)(i * B + jj) * nb + u] * xf[(size_t)(colidx[p0 + u] * (int64_t)B + jj)];
            yf[(size_t)(n * B + i)] = acc;
          }
        }
        cuda_emu::launch(grid, 256, 0, [&]() {
          asm_restrict_kernel<B>(nv, T.v2e_ptr.data(), v2e.data(), elem_nodes, nn, ncorner, phi.data(), n2e_ptr, n2e,
                                 keep.data(), yf.data(), rc.data(), nullptr);
        });
        for (int64_t v = 0; v < nv; ++v)
          for (int64_t k = T.adj_ptr[(size_t)v]; k < T.adj_ptr[(size_t)v + 1]; ++k)
            if (T.color[(size_t)T.adj[(size_t)k]] == color)
              for (int i = 0; i < B; ++i) acval[(size_t)((k * B + i) * B + j)] = rc[(size_t)(v * B + i)];
      }
    asm_coarse_symmetrise(nv, B, T.adj_ptr.data(), T.adj.data(), acval.data());
    dinv.assign((size_t)(nv * B), 0.0);
    asm_coarse_jacobi(nv, B, T.adj_ptr.data(), T.adj.data(), acval.data(), dinv.data());
    const double lmax = 1.1 * asm_power_lmax(nv, B, T.adj_ptr.data(), T.adj.data(), acval.data(), dinv.data(), 40);
    if (!(lmax > 0.0)) return -7;
    asm_cheb_coefficients(coarse_steps, 0.1 * lmax, lmax, &c1, &c2);
    coarse_info[0] = nv;
    coarse_info[1] = T.ncolors;
    if (ac_dense) {                                     // [nv*B][nv*B] for the test's comparison with P^T K P
      const int64_t nc = nv * B;
      for (int64_t v = 0; v < nv; ++v)
        for (int64_t k = T.adj_ptr[(size_t)v]; k < T.adj_ptr[(size_t)v + 1]; ++k)
          for (int i = 0; i < B; ++i)
            for (int j = 0; j < B; ++j)
              ac_dense[(v * B + i) * nc + T.adj[(size_t)k] * (int64_t)B + j] = acval[(size_t)((k * B + i) * B + j)];
    }
  }
  // apply, as asm_apply_impl; napply > 1 checks that nothing is left behind between applies
  std::vector<double> partials((size_t)grid * 8, 0.0);
  unsigned int ticket = 0;
  const double done0 = 0.0;
  for (int it = 0; it < napply; ++it) {
    cuda_emu::launch(grid, 256, smem, [&]() {
      asm_mult_kernel<B>(ntiles, tile_elems.data(), tile_class.data(), elem_nodes, nn, inv.data(), wnode.data(), r,
                         ybuf.data(), &done0);
    });
    AsmProlong cp{nullptr, T.elem_verts.data(), phi.data(), keep.data(), ncorner};
    if (coarse_steps > 0) {
      const int64_t nv = T.nv;
      cuda_emu::launch(grid, 256, 0, [&]() {
        asm_restrict_kernel<B>(nv, T.v2e_ptr.data(), v2e.data(), elem_nodes, nn, ncorner, phi.data(), n2e_ptr, n2e,
                               keep.data(), r, rc.data(), &done0);
      });
      double* dold = d0.data();
      double* dnew = d1.data();
      for (int s = 0; s <= coarse_steps; ++s) {
        cuda_emu::launch(grid, 256, 0, [&]() {
          asm_cheb_kernel<B>(nv, T.adj_ptr.data(), T.adj.data(), acval.data(), dinv.data(), s == 0 ? 1 : 0,
                             c1[(size_t)s], c2[(size_t)s], rc.data(), xc.data(), dold, dnew, &done0);
        });
        std::swap(dold, dnew);
      }
      cp.xc = xc.data();
    }
    cuda_emu::launch(grid, 256, 0, [&]() {
      asm_sum_kernel<B>(n_own, n2e_ptr, n2e, ybuf.data(), ld, r, z, partials.data(), &ticket, dots, &done0, cp);
    });
    if (ticket != 0) return -3;                 // the last CTA re-arms the ticket
  }
  // a raised `done` flag must leave z and the dots alone
  const double done1 = 1.0;
  std::vector<double> zkeep(z, z + n_own * B);
  double dkeep[2] = {dots[0], dots[1]};
  cuda_emu::launch(grid, 256, smem, [&]() {
    asm_mult_kernel<B>(ntiles, tile_elems.data(), tile_class.data(), elem_nodes, nn, inv.data(), wnode.data(), z,
                       ybuf.data(), &done1);
  });
  cuda_emu::launch(grid, 256, 0, [&]() {
    asm_sum_kernel<B>(n_own, n2e_ptr, n2e, ybuf.data(), ld, r, z, partials.data(), &ticket, dots, &done1,
                      AsmProlong{nullptr, nullptr, nullptr, nullptr, 0});
  });
  for (long long i = 0; i < n_own * B; ++i)
    if (z[i] != zkeep[(size_t)i]) return -4;
  if (dots[0] != dkeep[0] || dots[1] != dkeep[1]) return -4;
  return 0;
}

}  // namespace

extern "C" {
// n2e: (element, local index) pairs, int32 x 2 per entry.  inv_out: NULL or [nclass][ld][ld].
// n_own < nnodes: one rank of a row-partitioned run (nodes [n_own, nnodes) are ghosts whose rows are not
// here; valence_global[nnodes] gives their global valences, nodecls[nnodes] the packed dof classes; r holds
// refreshed ghost entries; z and the dots cover the owned nodes).
// coarse_steps > 0: two-level; coarse_info[2] = {coarse vertices, colours}; ac_dense: NULL or the
// probed coarse operator as a dense [nv*B][nv*B] (zero-initialised by the caller).
// Returns 0, or < 0 on a failed internal check.
int emu_asm_apply(int B, int nn, long long nel, long long nnodes, const int* elem_nodes, const long long* rowptr,
                  const int* colidx, const double* val, const long long* n2e_ptr, const int* n2e, const double* r,
                  int grid, int napply, double* z, double* dots, long long* nclass_out, int* cls_out,
                  double* inv_out, int dim, int ngl, int coarse_steps, long long* coarse_info, double* ac_dense,
                  long long n_own, const double* valence_global, const unsigned char* nodecls, double quantum_rel) {
  static_assert(sizeof(long long) == sizeof(int64_t) && sizeof(int2) == 2 * sizeof(int), "ABI");
  const int64_t* rp = reinterpret_cast<const int64_t*>(rowptr);
  const int64_t* np = reinterpret_cast<const int64_t*>(n2e_ptr);
  const int2* ne = reinterpret_cast<const int2*>(n2e);
  if (B == 3)
    return run<3>(nn, nel, nnodes, elem_nodes, rp, colidx, val, np, ne, r, grid, napply, z, dots, nclass_out, cls_out,
                  inv_out, dim, ngl, coarse_steps, coarse_info, ac_dense, n_own, valence_global, nodecls, quantum_rel);
  if (B == 2)
    return run<2>(nn, nel, nnodes, elem_nodes, rp, colidx, val, np, ne, r, grid, napply, z, dots, nclass_out, cls_out,
                  inv_out, dim, ngl, coarse_steps, coarse_info, ac_dense, n_own, valence_global, nodecls, quantum_rel);
  return -1;
}
}
