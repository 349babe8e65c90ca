// comm.cuh -- one-process-per-GPU plumbing: NCCL communicator (loaded with dlopen, so the
// single-GPU library has no NCCL dependency), halo exchange of ghost dofs, scalar allreduce.
//
// Replaces what PETSc/MPI does for the reference under mpiexec: the VecScatter ghost update
// inside MatMult (c1) and the MPI_Allreduce of every VecDot/VecNorm in the KSP (c2)
// (SURVEY.md 2.2, 5).  There is no assembly exchange: rows are assembled by their owner from
// redundantly evaluated halo elements (pynama_b200/parallel.py).
#pragma once
#include "common.cuh"

struct pyn_halo {
  pyn_ctx* ctx = nullptr;
  int64_t n_own = 0, n_loc = 0;          // nodes
  int nneigh = 0;
  std::vector<int> ranks;
  std::vector<int64_t> send_ptr;         // [nneigh+1] node offsets into send_nodes
  std::vector<int64_t> recv_off, recv_cnt;  // node offsets into the ghost section / counts
  pyn::DevBuf<int32_t> send_nodes;       // local ids of owned nodes, grouped by neighbour
  pyn::DevBuf<double> sendbuf;           // [total_send * bs], grown on demand
  int64_t sendbuf_bs = 0;
};

namespace pyn {
// in-place sum over ranks of n doubles on the device, on the context's stream (no-op for 1 rank)
int comm_allreduce_sum(pyn_ctx* ctx, double* dev, int n);
// ghost entries of v (n_loc*bs doubles) <- owners' values
int halo_exchange(pyn_halo* h, double* v, int bs);
}  // namespace pyn
