// comm.cuh -- one-process-per-GPU plumbing: NCCL communicator (loaded with dlopen, so the
// single-GPU library has no NCCL dependency), halo exchange of ghost dofs, scalar allreduce.
//
// Replaces what PETSc/MPI does for the reference under mpiexec: the VecScatter ghost update
// inside MatMult (c1) and the MPI_Allreduce of every VecDot/VecNorm in the KSP (c2)
// (SURVEY.md 2.2, 5).  There is no assembly exchange: rows are assembled by their owner from
// redundantly evaluated halo elements (pynama_b200/parallel.py).
#pragma once
#include "common.cuh"

namespace pyn {
constexpr int kMaxRanks = 8;            // GPUs of one box
constexpr long long kSpinMax = 40000000LL;   // bounded spin (~ seconds): a lost peer must not hang the GPU

// ---- peer-memory collectives (NVLink/NVSwitch, CUDA IPC) -----------------------------------
// Every rank owns a small "window" in its HBM that all peers map.  A reduction is: the last
// CTA of the producing kernel stores this rank's partial into EVERY peer's window (remote
// stores over NVLink), fences, then releases a sequence-numbered flag; consumers spin on the
// flags in their OWN memory and add the partials in rank order, so every rank gets the same
// bits.  Slots are double-buffered by sequence parity.  No NCCL call, no extra launch: the
// collective rides inside the compute kernels.
struct RedWindow {                      // device layout of one rank's reduction window
  double val[2][2][kMaxRanks][4];       // [kind][parity][source rank][value]
  unsigned long long flag[2][2][kMaxRanks];
};
struct P2PRed {                         // by-value kernel argument
  RedWindow* win[kMaxRanks];            // win[t] = rank t's window (win[rank] = own)
  int rank, nranks;                     // nranks <= 1: disabled
  unsigned long long seq;               // sequence number of this collective (> 0)
};

#ifdef __CUDACC__
__device__ __forceinline__ void st_release_sys(unsigned long long* p, unsigned long long v) {
  asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ unsigned long long ld_acquire_sys(const unsigned long long* p) {
  unsigned long long v;
  asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}
// Called by ONE warp.  Lane t < nranks stores this rank's NV partials into rank t's window
// and releases the flag.
template <int NV>
__device__ __forceinline__ void p2p_push(const P2PRed& c, int kind, const double* v) {
  const int lane = threadIdx.x & 31;
  const int par = (int)(c.seq & 1ull);
  if (lane < c.nranks) {
    RedWindow* w = c.win[lane];
#pragma unroll
    for (int k = 0; k < NV; ++k) w->val[kind][par][c.rank][k] = v[k];
    __threadfence_system();
    st_release_sys(&w->flag[kind][par][c.rank], c.seq);
  }
}
// Called by ONE warp.  Waits for every rank's partial of collective c.seq and returns the
// sums (rank order) in out[0..NV) on all lanes; returns false on spin timeout.
template <int NV>
__device__ __forceinline__ bool p2p_wait_sum(const P2PRed& c, int kind, double* out) {
  const int lane = threadIdx.x & 31;
  const int par = (int)(c.seq & 1ull);
  RedWindow* w = c.win[c.rank];
  bool ok = true;
  if (lane < c.nranks) {
    long long spins = 0;
    while (ld_acquire_sys(&w->flag[kind][par][lane]) < c.seq)
      if (++spins > kSpinMax) {
        ok = false;
        break;
      }
  }
  ok = __all_sync(0xffffffffu, ok);
#pragma unroll
  for (int k = 0; k < NV; ++k) {
    double s = 0.0;
    for (int r = 0; r < c.nranks; ++r) s += __ldcv(&w->val[kind][par][r][k]);
    out[k] = s;
  }
  return ok;
}
// Grid reduction whose result goes to every rank: CTAs deposit NV partials, the last CTA to
// arrive folds them in block order and pushes this rank's totals into all peers' windows.
// v[] must hold the CTA totals in warp 0 (after block_reduce); call with all threads.
template <int NV>
__device__ __forceinline__ void grid_reduce_push(double (&v)[NV], double* partials, unsigned int* ticket,
                                                 const P2PRed& red, int kind) {
  __shared__ bool is_last_p;
  __shared__ double tot_p[NV];
  if (threadIdx.x == 0) {
#pragma unroll
    for (int k = 0; k < NV; ++k) partials[(size_t)blockIdx.x * NV + k] = v[k];
    __threadfence();
    is_last_p = (atomicAdd(ticket, 1u) == gridDim.x - 1);
  }
  __syncthreads();
  if (!is_last_p) return;
  __threadfence();
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  if (wid < NV) {
    double t = 0.0;
    for (unsigned int b = lane; b < gridDim.x; b += 32) t += partials[(size_t)b * NV + wid];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) t += __shfl_xor_sync(0xffffffffu, t, o);
    if (lane == 0) tot_p[wid] = t;
  }
  __syncthreads();
  if (wid == 0) {
    double t[NV];
#pragma unroll
    for (int k = 0; k < NV; ++k) t[k] = tot_p[k];
    p2p_push<NV>(red, kind, t);
  }
  if (threadIdx.x == 0) *ticket = 0u;
}
#endif

// ---- halo exchange fused into the Krylov kernels (peer-memory path) ------------------------
// Producer side: the vector kernel that computes the next product's input stores the values
// of the nodes its neighbours hold as ghosts straight into their windows; its last CTA
// releases the flags.  Consumer side: the product kernel waits for the flags at CTA start
// and reads ghost columns from its own window.  No pack / wait / unpack launches.
constexpr int kHaloHdr = 256;   // bytes: flag[2][kMaxRanks] u64, padded
struct HaloPush {
  const int64_t* nslot_ptr;          // [n_own+1] send slots of every owned node (most have none)
  const int32_t* nslot;              // slot ids k (position in the send list, grouped by neighbour)
  const int64_t* send_ptr;           // [nneigh+1]
  const unsigned long long* peer_win;  // [nneigh] neighbours' window addresses
  const int64_t* peer_off;           // [nneigh] offset of my run in the neighbour's ghost section
  const int64_t* peer_nghost;        // [nneigh] the neighbour's ghost count
  int nneigh;                        // 0: disabled
  int bs, max_bs, my_rank;
  unsigned long long seq;
};
struct HaloWait {
  unsigned char* win;                // own window
  const int32_t* ranks;              // [nneigh]
  int nneigh;                        // 0: disabled
  int bs, max_bs;
  int64_t n_own, nghost;
  unsigned long long seq;
  int* timeout_flag;
};
#ifdef __CUDACC__
__device__ __forceinline__ unsigned long long* halo_flag(unsigned char* win, int par, int src) {
  return reinterpret_cast<unsigned long long*>(win) + par * kMaxRanks + src;
}
__device__ __forceinline__ double* halo_buf(unsigned char* win, int par, int64_t nghost, int max_bs) {
  return reinterpret_cast<double*>(win + kHaloHdr) + (int64_t)par * nghost * max_bs;
}
// values vals[0..bs) of owned node n -> every neighbour that holds n as a ghost
__device__ __forceinline__ void halo_push_node(const HaloPush& h, int64_t n, const double* vals) {
  const int par = (int)(h.seq & 1ull);
  for (int64_t q = h.nslot_ptr[n]; q < h.nslot_ptr[n + 1]; ++q) {
    const int64_t k = h.nslot[q];
    int nb = 0;
    while (k >= h.send_ptr[nb + 1]) ++nb;
    double* dst = halo_buf(reinterpret_cast<unsigned char*>(h.peer_win[nb]), par, h.peer_nghost[nb], h.max_bs) +
                  (h.peer_off[nb] + (k - h.send_ptr[nb])) * h.bs;
    for (int c = 0; c < h.bs; ++c) dst[c] = vals[c];
  }
}
// by the last CTA of the producer, after every CTA fenced its stores (threadfence_system
// before the ticket): one thread per neighbour
__device__ __forceinline__ void halo_release_flags(const HaloPush& h) {
  __threadfence_system();
  if ((int)threadIdx.x < h.nneigh)
    st_release_sys(halo_flag(reinterpret_cast<unsigned char*>(h.peer_win[threadIdx.x]), (int)(h.seq & 1ull), h.my_rank),
                   h.seq);
}
// all threads of the CTA call; false on spin timeout
__device__ __forceinline__ bool halo_wait_cta(const HaloWait& h) {
  __shared__ int halo_ok_sm;
  if (threadIdx.x == 0) halo_ok_sm = 1;
  __syncthreads();
  if ((int)threadIdx.x < h.nneigh) {
    long long spins = 0;
    while (ld_acquire_sys(halo_flag(h.win, (int)(h.seq & 1ull), h.ranks[threadIdx.x])) < h.seq)
      if (++spins > kSpinMax) {
        halo_ok_sm = 0;
        *h.timeout_flag = 1;
        break;
      }
  }
  __syncthreads();
  return halo_ok_sm != 0;
}
// all lanes of a warp call; the lanes below nneigh spin on one flag each.  A timeout raises the flag
// the host reads and lets the warp go on (the solve ends with reason -100).
__device__ __forceinline__ void halo_wait_warp(const HaloWait& h) {
  const int lane = threadIdx.x & 31;
  for (int j = lane; j < h.nneigh; j += 32) {
    long long spins = 0;
    while (ld_acquire_sys(halo_flag(h.win, (int)(h.seq & 1ull), h.ranks[j])) < h.seq)
      if (++spins > kSpinMax) {
        *h.timeout_flag = 1;
        break;
      }
  }
  __syncwarp();
}
#endif
}  // namespace pyn

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
  // peer-memory path: this rank's receive window (flags + 2 x n_ghost*max_bs doubles) and the
  // neighbours' windows mapped through CUDA IPC; senders store straight into the receiver
  bool p2p = false;
  int max_bs = 0;
  unsigned char* win = nullptr;                 // own window
  std::vector<unsigned char*> peer_win;         // [nneigh]
  std::vector<int64_t> peer_off;                // [nneigh] my run's node offset in neighbour i's ghost section
  std::vector<int64_t> peer_nghost;             // [nneigh] neighbour i's ghost count (window stride)
  pyn::DevBuf<int64_t> d_send_ptr, d_peer_off, d_peer_nghost, d_recv_off, d_recv_cnt;
  pyn::DevBuf<int32_t> d_ranks;
  pyn::DevBuf<unsigned long long> d_peer_win;   // device array of the neighbours' window addresses
  pyn::DevBuf<int64_t> d_nslot_ptr;             // [n_own+1] inverse of the send list: owned node -> its send slots
  pyn::DevBuf<int32_t> d_nslot;
  unsigned long long seq = 0;
  ~pyn_halo();
};

namespace pyn {
// in-place sum over ranks of n doubles on the device, on the context's stream (no-op for 1 rank)
int comm_allreduce_sum(pyn_ctx* ctx, double* dev, int n);
// ghost entries of v (n_loc*bs doubles) <- owners' values
int halo_exchange(pyn_halo* h, double* v, int bs);
// reduction descriptor for the next fused collective (advances the sequence); disabled
// (nranks = 1) unless the peer windows are attached
P2PRed p2p_next(pyn_ctx* ctx);
bool p2p_enabled(const pyn_ctx* ctx);
// descriptors of the NEXT fused halo exchange of h (advances its sequence); bs = dofs per node
void halo_fused_next(pyn_halo* h, int bs, HaloPush* push, HaloWait* wait);
}  // namespace pyn
