// emu_asm.cpp -- TEST-ONLY: the kernels of the element-block Schwarz preconditioner
// (pynama_b200/csrc/asm_kernels.cuh) run on CPU threads through cuda_emu.h, driven the way
// asm_pc.cu drives them on the device (hash -> group -> extract -> invert -> tiles -> weights;
// then product + node sum).  Built with g++ (not nvcc) into libemu_asm.so by
// __graft_entry__.build(); loaded only by tests/test_asm_emulated.py.
#include "cuda_emu.h"

#include "../../pynama_b200/csrc/asm_kernels.cuh"

namespace pyn {
void set_error(const char*, ...) {}
}  // namespace pyn

using namespace pyn;

namespace {

template <int B>
int run(int nn, long long nel, long long nnodes, const int32_t* elem_nodes, const int64_t* rowptr,
        const int32_t* colidx, const double* val, const int64_t* n2e_ptr, const int2* n2e, const double* r,
        int grid, int napply, double* z, double* dots, long long* nclass_out, int32_t* cls_out, double* inv_out) {
  const int ld = nn * B;
  // set-up, as asm_setup_impl
  std::vector<unsigned long long> hash((size_t)(2 * nel), 0ull);
  cuda_emu::launch(grid, 256, 0, [&]() {
    asm_blocks_kernel<B, 0>(nel, nullptr, elem_nodes, nn, rowptr, colidx, val, hash.data(), nullptr);
  });
  std::vector<int32_t> cls, rep;
  const int64_t nclass = asm_group_by_hash(nel, reinterpret_cast<const uint64_t*>(hash.data()), &cls, &rep);
  *nclass_out = nclass;
  for (long long e = 0; e < nel; ++e) cls_out[e] = cls[(size_t)e];
  std::vector<double> inv((size_t)(nclass * (int64_t)ld * ld), -777.0);
  cuda_emu::launch(grid, 256, 0, [&]() {
    asm_blocks_kernel<B, 1>(nclass, rep.data(), elem_nodes, nn, rowptr, colidx, val, nullptr, inv.data());
  });
  for (int64_t c = 0; c < nclass; ++c) {
    double* a = inv.data() + c * (int64_t)ld * ld;
    if (asm_dense_inverse(ld, a) != 0) return -2;
    asm_symmetrise(ld, a);
  }
  if (inv_out) memcpy(inv_out, inv.data(), sizeof(double) * inv.size());
  std::vector<int32_t> tile_elems, tile_class;
  const int64_t ntiles = asm_make_tiles(nel, cls.data(), nclass, kAsmTile, &tile_elems, &tile_class);
  std::vector<double> wnode((size_t)nnodes, -1.0);
  cuda_emu::launch(grid, 256, 0, [&]() { asm_weights_kernel(nnodes, n2e_ptr, wnode.data()); });
  std::vector<double> ybuf((size_t)(nel * ld), -777.0);
  const size_t smem = sizeof(double) * ((size_t)ld * kAsmTile + (size_t)kAsmKC * kAsmRP);
  // apply, as asm_apply_impl; napply > 1 checks that nothing is left behind between applies
  std::vector<double> partials((size_t)grid * 8, 0.0);
  unsigned int ticket = 0;
  const double done0 = 0.0;
  for (int it = 0; it < napply; ++it) {
    cuda_emu::launch(grid, 256, smem, [&]() {
      asm_mult_kernel<B>(ntiles, tile_elems.data(), tile_class.data(), elem_nodes, nn, inv.data(), wnode.data(), r,
                         ybuf.data(), &done0);
    });
    cuda_emu::launch(grid, 256, 0, [&]() {
      asm_sum_kernel<B>(nnodes, n2e_ptr, n2e, ybuf.data(), ld, r, z, partials.data(), &ticket, dots, &done0);
    });
    if (ticket != 0) return -3;                 // the last CTA re-arms the ticket
  }
  // a raised `done` flag must leave z and the dots alone
  const double done1 = 1.0;
  std::vector<double> zkeep(z, z + nnodes * B);
  double dkeep[2] = {dots[0], dots[1]};
  cuda_emu::launch(grid, 256, smem, [&]() {
    asm_mult_kernel<B>(ntiles, tile_elems.data(), tile_class.data(), elem_nodes, nn, inv.data(), wnode.data(), z,
                       ybuf.data(), &done1);
  });
  cuda_emu::launch(grid, 256, 0, [&]() {
    asm_sum_kernel<B>(nnodes, n2e_ptr, n2e, ybuf.data(), ld, r, z, partials.data(), &ticket, dots, &done1);
  });
  for (long long i = 0; i < nnodes * B; ++i)
    if (z[i] != zkeep[(size_t)i]) return -4;
  if (dots[0] != dkeep[0] || dots[1] != dkeep[1]) return -4;
  return 0;
}

}  // namespace

extern "C" {
// n2e: (element, local index) pairs, int32 x 2 per entry.  inv_out: NULL or [nclass][ld][ld].
// Returns 0, or < 0 on a failed internal check.
int emu_asm_apply(int B, int nn, long long nel, long long nnodes, const int* elem_nodes, const long long* rowptr,
                  const int* colidx, const double* val, const long long* n2e_ptr, const int* n2e, const double* r,
                  int grid, int napply, double* z, double* dots, long long* nclass_out, int* cls_out,
                  double* inv_out) {
  static_assert(sizeof(long long) == sizeof(int64_t) && sizeof(int2) == 2 * sizeof(int), "ABI");
  const int64_t* rp = reinterpret_cast<const int64_t*>(rowptr);
  const int64_t* np = reinterpret_cast<const int64_t*>(n2e_ptr);
  const int2* ne = reinterpret_cast<const int2*>(n2e);
  if (B == 3)
    return run<3>(nn, nel, nnodes, elem_nodes, rp, colidx, val, np, ne, r, grid, napply, z, dots, nclass_out, cls_out,
                  inv_out);
  if (B == 2)
    return run<2>(nn, nel, nnodes, elem_nodes, rp, colidx, val, np, ne, r, grid, napply, z, dots, nclass_out, cls_out,
                  inv_out);
  return -1;
}
}
