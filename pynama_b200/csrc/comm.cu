// comm.cu -- NCCL (dlopen'ed) communicator, halo exchange, scalar allreduce.
#include "comm.cuh"

#include <dlfcn.h>

namespace pyn {

// Minimal NCCL ABI (stable across 2.x): declared here so no NCCL header/library is needed at
// build or load time of the single-GPU path.
typedef struct { char internal[128]; } nccl_uid_t;
typedef void* nccl_comm_t;
enum { NCCL_SUCCESS = 0 };
enum { NCCL_FLOAT64 = 8 };   // ncclDataType_t: ncclFloat64 / ncclDouble
enum { NCCL_SUM = 0 };       // ncclRedOp_t

struct NcclApi {
  void* lib = nullptr;
  int (*GetUniqueId)(nccl_uid_t*) = nullptr;
  int (*CommInitRank)(nccl_comm_t*, int, nccl_uid_t, int) = nullptr;
  int (*CommDestroy)(nccl_comm_t) = nullptr;
  int (*AllReduce)(const void*, void*, size_t, int, int, nccl_comm_t, cudaStream_t) = nullptr;
  int (*Send)(const void*, size_t, int, int, nccl_comm_t, cudaStream_t) = nullptr;
  int (*Recv)(void*, size_t, int, int, nccl_comm_t, cudaStream_t) = nullptr;
  int (*GroupStart)() = nullptr;
  int (*GroupEnd)() = nullptr;
  const char* (*GetErrorString)(int) = nullptr;
  int (*GetVersion)(int*) = nullptr;
};

static NcclApi g_nccl;

static int load_nccl(const char* path) {
  if (g_nccl.lib) return PYN_OK;
  const char* cands[3] = {path && path[0] ? path : nullptr, "libnccl.so.2", "libnccl.so"};
  void* lib = nullptr;
  for (int i = 0; i < 3 && !lib; ++i)
    if (cands[i]) lib = dlopen(cands[i], RTLD_NOW | RTLD_LOCAL);
  if (!lib) {
    set_error("cannot load NCCL (%s): %s", path ? path : "libnccl.so.2", dlerror());
    return PYN_E_NCCL;
  }
#define SYM(field, name)                                        \
  *(void**)(&g_nccl.field) = dlsym(lib, name);                  \
  if (!g_nccl.field) {                                          \
    set_error("NCCL symbol %s not found", name);                \
    return PYN_E_NCCL;                                          \
  }
  SYM(GetUniqueId, "ncclGetUniqueId");
  SYM(CommInitRank, "ncclCommInitRank");
  SYM(CommDestroy, "ncclCommDestroy");
  SYM(AllReduce, "ncclAllReduce");
  SYM(Send, "ncclSend");
  SYM(Recv, "ncclRecv");
  SYM(GroupStart, "ncclGroupStart");
  SYM(GroupEnd, "ncclGroupEnd");
  SYM(GetErrorString, "ncclGetErrorString");
  SYM(GetVersion, "ncclGetVersion");
#undef SYM
  g_nccl.lib = lib;
  return PYN_OK;
}

#define PYN_NCCL(call)                                                                        \
  do {                                                                                        \
    int _e = (call);                                                                          \
    if (_e != NCCL_SUCCESS) {                                                                 \
      pyn::set_error("%s:%d: %s -> %s", __FILE__, __LINE__, #call, g_nccl.GetErrorString(_e)); \
      return PYN_E_NCCL;                                                                      \
    }                                                                                         \
  } while (0)

int comm_allreduce_sum(pyn_ctx* ctx, double* dev, int n) {
  if (ctx->nranks <= 1) return PYN_OK;
  PYN_NCCL(g_nccl.AllReduce(dev, dev, (size_t)n, NCCL_FLOAT64, NCCL_SUM, (nccl_comm_t)ctx->nccl_comm,
                            ctx->stream));
  ctx->collectives++;
  return PYN_OK;
}

__global__ void halo_pack_kernel(const double* __restrict__ v, const int32_t* __restrict__ nodes,
                                 int64_t n, int bs, double* __restrict__ out) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n * bs;
       i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t k = i / bs;
    const int c = (int)(i - k * bs);
    out[i] = v[(int64_t)nodes[k] * bs + c];
  }
}

// ---- peer-memory halo: senders store straight into the receiver's window -------------------

__global__ void halo_push_kernel(const double* __restrict__ v, const int32_t* __restrict__ send_nodes,
                                 const int64_t* __restrict__ send_ptr, int nneigh,
                                 const unsigned long long* __restrict__ peer_win,
                                 const int64_t* __restrict__ peer_off, const int64_t* __restrict__ peer_nghost,
                                 int bs, int max_bs, int my_rank, unsigned long long seq, unsigned int* ticket) {
  const int par = (int)(seq & 1ull);
  const int64_t nsend = send_ptr[nneigh];
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < nsend * bs;
       i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t k = i / bs;
    const int c = (int)(i - k * bs);
    int nb = 0;
    while (k >= send_ptr[nb + 1]) ++nb;                    // <= 26 neighbours
    double* dst = halo_buf(reinterpret_cast<unsigned char*>(peer_win[nb]), par, peer_nghost[nb], max_bs);
    dst[(peer_off[nb] + (k - send_ptr[nb])) * bs + c] = v[(int64_t)send_nodes[k] * bs + c];
  }
  // all stores of this CTA visible system-wide, then the last CTA releases the flags
  __threadfence_system();
  __shared__ bool is_last;
  __syncthreads();
  if (threadIdx.x == 0) is_last = (atomicAdd(ticket, 1u) == gridDim.x - 1);
  __syncthreads();
  if (!is_last) return;
  __threadfence_system();
  if (threadIdx.x < nneigh)
    st_release_sys(halo_flag(reinterpret_cast<unsigned char*>(peer_win[threadIdx.x]), par, my_rank), seq);
  if (threadIdx.x == 0) *ticket = 0u;
}

// CTAs [i*cpn, (i+1)*cpn) wait for neighbour i's flag, then copy its run into the ghost section
__global__ void halo_wait_kernel(double* __restrict__ v, unsigned char* win, int64_t n_own, int64_t nghost,
                                 const int32_t* __restrict__ ranks, const int64_t* __restrict__ recv_off,
                                 const int64_t* __restrict__ recv_cnt, int cpn, int bs, int max_bs,
                                 unsigned long long seq, int* timeout_flag) {
  const int par = (int)(seq & 1ull);
  const int nb = blockIdx.x / cpn, sub = blockIdx.x % cpn;
  __shared__ bool ok;
  if (threadIdx.x == 0) {
    long long spins = 0;
    ok = true;
    while (ld_acquire_sys(halo_flag(win, par, ranks[nb])) < seq)
      if (++spins > kSpinMax) {
        ok = false;
        *timeout_flag = 1;
        break;
      }
  }
  __syncthreads();
  if (!ok) return;
  const double* src = halo_buf(win, par, nghost, max_bs) + recv_off[nb] * bs;
  double* dst = v + (n_own + recv_off[nb]) * bs;
  const int64_t n = recv_cnt[nb] * bs;
  for (int64_t i = (int64_t)sub * blockDim.x + threadIdx.x; i < n; i += (int64_t)cpn * blockDim.x)
    dst[i] = __ldcv(src + i);
}

static int halo_exchange_p2p(pyn_halo* h, double* v, int bs) {
  pyn_ctx* ctx = h->ctx;
  PYN_REQUIRE(bs <= h->max_bs, "halo_exchange: block size %d exceeds the window (%d)", bs, h->max_bs);
  const unsigned long long seq = ++h->seq;
  const int64_t nsend = h->send_ptr[h->nneigh];
  const int gridp = grid_for(ctx, nsend * bs > 0 ? nsend * bs : 1, 256);
  halo_push_kernel<<<gridp, 256, 0, ctx->stream>>>(v, h->send_nodes.p, h->d_send_ptr.p, h->nneigh, h->d_peer_win.p,
                                                   h->d_peer_off.p, h->d_peer_nghost.p, bs, h->max_bs, ctx->rank,
                                                   seq, ctx->red_ticket);
  PYN_LAUNCHED(ctx);
  const int cpn = 4;
  halo_wait_kernel<<<h->nneigh * cpn, 256, 0, ctx->stream>>>(v, h->win, h->n_own, h->n_loc - h->n_own, h->d_ranks.p,
                                                             h->d_recv_off.p, h->d_recv_cnt.p, cpn, bs, h->max_bs,
                                                             seq, ctx->p2p_timeout);
  PYN_LAUNCHED(ctx);
  ctx->collectives++;
  return PYN_OK;
}

void halo_fused_next(pyn_halo* h, int bs, HaloPush* push, HaloWait* wait) {
  pyn_ctx* ctx = h->ctx;
  const unsigned long long seq = ++h->seq;
  push->nslot_ptr = h->d_nslot_ptr.p;
  push->nslot = h->d_nslot.p;
  push->send_ptr = h->d_send_ptr.p;
  push->peer_win = h->d_peer_win.p;
  push->peer_off = h->d_peer_off.p;
  push->peer_nghost = h->d_peer_nghost.p;
  push->nneigh = h->nneigh;
  push->bs = bs;
  push->max_bs = h->max_bs;
  push->my_rank = ctx->rank;
  push->seq = seq;
  wait->win = h->win;
  wait->ranks = h->d_ranks.p;
  wait->nneigh = h->nneigh;
  wait->bs = bs;
  wait->max_bs = h->max_bs;
  wait->n_own = h->n_own;
  wait->nghost = h->n_loc - h->n_own;
  wait->seq = seq;
  wait->timeout_flag = ctx->p2p_timeout;
  ctx->collectives++;
}

P2PRed p2p_next(pyn_ctx* ctx) {
  P2PRed c;
  memset(&c, 0, sizeof(c));
  c.rank = ctx->rank;
  c.nranks = 1;
  if (!ctx->p2p_on) return c;
  c.nranks = ctx->nranks;
  for (int t = 0; t < ctx->nranks; ++t) c.win[t] = reinterpret_cast<RedWindow*>(ctx->red_peer[t]);
  c.seq = ++ctx->p2p_seq;
  ctx->collectives++;
  return c;
}
bool p2p_enabled(const pyn_ctx* ctx) { return ctx->p2p_on && ctx->nranks > 1; }

int halo_exchange(pyn_halo* h, double* v, int bs) {
  pyn_ctx* ctx = h->ctx;
  if (ctx->nranks <= 1 || h->nneigh == 0) return PYN_OK;
  if (h->p2p) return halo_exchange_p2p(h, v, bs);
  const int64_t nsend = h->send_ptr[h->nneigh];
  if (h->sendbuf_bs < bs) {
    PYN_CHECK(h->sendbuf.alloc(ctx, nsend * bs > 0 ? nsend * bs : 1));
    h->sendbuf_bs = bs;
  }
  if (nsend > 0) {
    halo_pack_kernel<<<grid_for(ctx, nsend * bs, 256), 256, 0, ctx->stream>>>(v, h->send_nodes.p, nsend, bs,
                                                                              h->sendbuf.p);
    PYN_LAUNCHED(ctx);
  }
  PYN_NCCL(g_nccl.GroupStart());
  for (int i = 0; i < h->nneigh; ++i) {
    const int64_t ns = h->send_ptr[i + 1] - h->send_ptr[i];
    if (ns > 0)
      PYN_NCCL(g_nccl.Send(h->sendbuf.p + h->send_ptr[i] * bs, (size_t)(ns * bs), NCCL_FLOAT64, h->ranks[i],
                           (nccl_comm_t)ctx->nccl_comm, ctx->stream));
    if (h->recv_cnt[i] > 0)
      PYN_NCCL(g_nccl.Recv(v + (h->n_own + h->recv_off[i]) * bs, (size_t)(h->recv_cnt[i] * bs), NCCL_FLOAT64,
                           h->ranks[i], (nccl_comm_t)ctx->nccl_comm, ctx->stream));
  }
  PYN_NCCL(g_nccl.GroupEnd());
  ctx->collectives++;
  return PYN_OK;
}

}  // namespace pyn

using namespace pyn;

pyn_halo::~pyn_halo() {
  for (unsigned char* p : peer_win)
    if (p) cudaIpcCloseMemHandle(p);
  if (win) cudaFree(win);
}

extern "C" {

int pyn_comm_unique_id(const char* nccl_lib, void* id128) {
  PYN_REQUIRE(id128, "pyn_comm_unique_id: null id");
  PYN_CHECK(load_nccl(nccl_lib));
  nccl_uid_t id;
  PYN_NCCL(g_nccl.GetUniqueId(&id));
  memcpy(id128, &id, sizeof(id));
  return PYN_OK;
}

int pyn_comm_init(pyn_ctx* ctx, const char* nccl_lib, int rank, int nranks, const void* id128) {
  PYN_REQUIRE(ctx && nranks >= 1 && rank >= 0 && rank < nranks, "pyn_comm_init: bad rank %d of %d", rank,
              nranks);
  ctx->rank = rank;
  ctx->nranks = nranks;
  if (nranks == 1) return PYN_OK;
  PYN_REQUIRE(id128, "pyn_comm_init: null id");
  PYN_CHECK(load_nccl(nccl_lib));
  PYN_CUDA(cudaSetDevice(ctx->device));
  nccl_uid_t id;
  memcpy(&id, id128, sizeof(id));
  nccl_comm_t comm = nullptr;
  PYN_NCCL(g_nccl.CommInitRank(&comm, nranks, id, rank));
  ctx->nccl_comm = comm;
  // first collective builds the channels; do it here, not inside a timed region
  DevBuf<double> one;
  PYN_CHECK(one.alloc(ctx, 1));
  PYN_CHECK(one.zero());
  PYN_CHECK(comm_allreduce_sum(ctx, one.p, 1));
  PYN_CUDA(cudaStreamSynchronize(ctx->stream));
  return PYN_OK;
}

int pyn_comm_info(pyn_ctx* ctx, int* rank, int* nranks, int* nccl_version, int64_t* collectives) {
  PYN_REQUIRE(ctx, "null ctx");
  if (rank) *rank = ctx->rank;
  if (nranks) *nranks = ctx->nranks;
  if (nccl_version) {
    *nccl_version = 0;
    if (g_nccl.lib) g_nccl.GetVersion(nccl_version);
  }
  if (collectives) *collectives = ctx->collectives;
  return PYN_OK;
}

int pyn_comm_allreduce(pyn_ctx* ctx, double* vals, int n, int op) {
  PYN_REQUIRE(ctx && vals && n > 0 && n <= 64 && (op == 0 || op == 1), "pyn_comm_allreduce: bad argument");
  if (ctx->nranks <= 1) return PYN_OK;
  // op 0: sum; op 1: max (as a sum of one-hot slots: every rank writes its values into its own
  // row of an nranks x n table, the table is summed, the host takes the max)
  const int rows = op == 1 ? ctx->nranks : 1;
  std::vector<double> host((size_t)rows * n, 0.0);
  for (int i = 0; i < n; ++i) host[(size_t)(op == 1 ? ctx->rank : 0) * n + i] = vals[i];
  DevBuf<double> d;
  PYN_CHECK(d.alloc(ctx, rows * n));
  PYN_CHECK(d.upload(host.data(), rows * n));
  PYN_CHECK(comm_allreduce_sum(ctx, d.p, rows * n));
  PYN_CHECK(d.download(host.data(), rows * n));
  for (int i = 0; i < n; ++i) {
    double v = host[i];
    for (int r = 1; r < rows; ++r) v = host[(size_t)r * n + i] > v ? host[(size_t)r * n + i] : v;
    vals[i] = v;
  }
  return PYN_OK;
}

int pyn_comm_finalize(pyn_ctx* ctx) {
  if (ctx && ctx->p2p_on) {
    cudaStreamSynchronize(ctx->stream);
    for (int t = 0; t < ctx->nranks; ++t)
      if (t != ctx->rank && ctx->red_peer[t]) cudaIpcCloseMemHandle(ctx->red_peer[t]);
    ctx->p2p_on = false;
  }
  if (ctx && ctx->nccl_comm && g_nccl.lib) {
    cudaStreamSynchronize(ctx->stream);
    g_nccl.CommDestroy((nccl_comm_t)ctx->nccl_comm);
    ctx->nccl_comm = nullptr;
  }
  return PYN_OK;
}

int pyn_halo_create(pyn_ctx* ctx, int64_t n_own, int64_t n_loc, int nneigh, const int32_t* ranks,
                    const int64_t* send_ptr, const int32_t* send_nodes, const int64_t* recv_off,
                    const int64_t* recv_cnt, pyn_halo** out) {
  PYN_REQUIRE(ctx && out && n_own >= 0 && n_loc >= n_own && nneigh >= 0, "pyn_halo_create: bad argument");
  PYN_REQUIRE(nneigh == 0 || (ranks && send_ptr && recv_off && recv_cnt), "pyn_halo_create: null table");
  pyn_halo* h = new pyn_halo();
  h->ctx = ctx;
  h->n_own = n_own;
  h->n_loc = n_loc;
  h->nneigh = nneigh;
  h->send_ptr.assign(1, 0);
  for (int i = 0; i < nneigh; ++i) {
    bool ok = ranks[i] >= 0 && ranks[i] < ctx->nranks && ranks[i] != ctx->rank &&
              send_ptr[i + 1] >= send_ptr[i] && recv_off[i] >= 0 && recv_cnt[i] >= 0 &&
              recv_off[i] + recv_cnt[i] <= n_loc - n_own;
    if (!ok) {
      delete h;
      set_error("pyn_halo_create: inconsistent table for neighbour %d", i);
      return PYN_E_INVALID;
    }
    h->ranks.push_back(ranks[i]);
    h->send_ptr.push_back(send_ptr[i + 1]);
    h->recv_off.push_back(recv_off[i]);
    h->recv_cnt.push_back(recv_cnt[i]);
  }
  const int64_t nsend = nneigh ? send_ptr[nneigh] : 0;
  for (int64_t i = 0; i < nsend; ++i)
    if (send_nodes[i] < 0 || send_nodes[i] >= n_own) {
      delete h;
      set_error("pyn_halo_create: send node %lld is not an owned node", (long long)send_nodes[i]);
      return PYN_E_INVALID;
    }
  int rc = h->send_nodes.alloc(ctx, nsend > 0 ? nsend : 1);
  if (rc == PYN_OK && nsend > 0) rc = h->send_nodes.upload(send_nodes, nsend);
  // inverse of the send list (fused halo push: the thread that produces node n stores it remotely)
  std::vector<int64_t> nptr((size_t)n_own + 1, 0);
  for (int64_t i = 0; i < nsend; ++i) nptr[send_nodes[i] + 1]++;
  for (int64_t n = 0; n < n_own; ++n) nptr[n + 1] += nptr[n];
  std::vector<int32_t> nslot((size_t)(nsend > 0 ? nsend : 1), 0);
  {
    std::vector<int64_t> fill(nptr.begin(), nptr.end() - 1);
    for (int64_t i = 0; i < nsend; ++i) nslot[fill[send_nodes[i]]++] = (int32_t)i;
  }
  if (rc == PYN_OK) rc = h->d_nslot_ptr.alloc(ctx, n_own + 1);
  if (rc == PYN_OK) rc = h->d_nslot_ptr.upload(nptr.data(), n_own + 1);
  if (rc == PYN_OK) rc = h->d_nslot.alloc(ctx, (int64_t)nslot.size());
  if (rc == PYN_OK) rc = h->d_nslot.upload(nslot.data(), (int64_t)nslot.size());
  if (rc == PYN_OK && cudaStreamSynchronize(ctx->stream) != cudaSuccess) rc = PYN_E_CUDA;
  if (rc != PYN_OK) {
    delete h;
    return rc;
  }
  *out = h;
  return PYN_OK;
}

int pyn_halo_destroy(pyn_halo* h) {
  delete h;
  return PYN_OK;
}

// ---- peer windows (CUDA IPC) ---------------------------------------------------------------
int pyn_comm_window_export(pyn_ctx* ctx, void* handle64) {
  PYN_REQUIRE(ctx && handle64, "pyn_comm_window_export: null argument");
  static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle size");
  if (!ctx->red_win) {
    PYN_CUDA(cudaMalloc(&ctx->red_win, sizeof(RedWindow)));
    PYN_CUDA(cudaMemset(ctx->red_win, 0, sizeof(RedWindow)));
    PYN_CUDA(cudaMalloc((void**)&ctx->p2p_timeout, sizeof(int)));
    PYN_CUDA(cudaMemset(ctx->p2p_timeout, 0, sizeof(int)));
    PYN_CUDA(cudaDeviceSynchronize());
  }
  cudaIpcMemHandle_t hdl;
  PYN_CUDA(cudaIpcGetMemHandle(&hdl, ctx->red_win));
  memcpy(handle64, &hdl, 64);
  return PYN_OK;
}

int pyn_comm_window_attach(pyn_ctx* ctx, const void* handles) {
  PYN_REQUIRE(ctx && handles && ctx->red_win, "pyn_comm_window_attach: export first");
  PYN_REQUIRE(ctx->nranks <= kMaxRanks, "peer-memory collectives support up to %d ranks", kMaxRanks);
  for (int t = 0; t < ctx->nranks; ++t) {
    if (t == ctx->rank) {
      ctx->red_peer[t] = ctx->red_win;
      continue;
    }
    cudaIpcMemHandle_t hdl;
    memcpy(&hdl, (const char*)handles + 64 * t, 64);
    PYN_CUDA(cudaIpcOpenMemHandle(&ctx->red_peer[t], hdl, cudaIpcMemLazyEnablePeerAccess));
  }
  ctx->p2p_on = true;
  return PYN_OK;
}

int pyn_halo_window_export(pyn_halo* h, int max_bs, void* handle64) {
  PYN_REQUIRE(h && handle64 && max_bs > 0, "pyn_halo_window_export: bad argument");
  if (!h->win) {
    const int64_t nghost = h->n_loc - h->n_own;
    const size_t bytes = kHaloHdr + (size_t)2 * (nghost > 0 ? nghost : 1) * max_bs * 8;
    PYN_CUDA(cudaMalloc((void**)&h->win, bytes));
    PYN_CUDA(cudaMemset(h->win, 0, bytes));
    PYN_CUDA(cudaDeviceSynchronize());
    h->max_bs = max_bs;
    h->ctx->bytes_in_use += (int64_t)bytes;
  }
  cudaIpcMemHandle_t hdl;
  PYN_CUDA(cudaIpcGetMemHandle(&hdl, h->win));
  memcpy(handle64, &hdl, 64);
  return PYN_OK;
}

int pyn_halo_window_attach(pyn_halo* h, const void* handles, const int64_t* peer_off, const int64_t* peer_nghost) {
  PYN_REQUIRE(h && h->win && (h->nneigh == 0 || (handles && peer_off && peer_nghost)),
              "pyn_halo_window_attach: export first");
  PYN_REQUIRE(h->ctx->p2p_on, "pyn_halo_window_attach: the context's reduction windows are not attached");
  PYN_REQUIRE(h->nneigh <= 256, "too many neighbours");
  pyn_ctx* ctx = h->ctx;
  std::vector<unsigned long long> addr(h->nneigh > 0 ? h->nneigh : 1, 0ull);
  std::vector<int32_t> ranks(h->ranks.begin(), h->ranks.end());
  for (int i = 0; i < h->nneigh; ++i) {
    cudaIpcMemHandle_t hdl;
    memcpy(&hdl, (const char*)handles + 64 * i, 64);
    void* p = nullptr;
    PYN_CUDA(cudaIpcOpenMemHandle(&p, hdl, cudaIpcMemLazyEnablePeerAccess));
    h->peer_win.push_back((unsigned char*)p);
    h->peer_off.push_back(peer_off[i]);
    h->peer_nghost.push_back(peer_nghost[i]);
    addr[i] = (unsigned long long)p;
  }
  const int n1 = h->nneigh > 0 ? h->nneigh : 1;
  PYN_CHECK(h->d_send_ptr.alloc(ctx, h->nneigh + 1));
  PYN_CHECK(h->d_send_ptr.upload(h->send_ptr.data(), h->nneigh + 1));
  PYN_CHECK(h->d_peer_off.alloc(ctx, n1));
  PYN_CHECK(h->d_peer_nghost.alloc(ctx, n1));
  PYN_CHECK(h->d_recv_off.alloc(ctx, n1));
  PYN_CHECK(h->d_recv_cnt.alloc(ctx, n1));
  PYN_CHECK(h->d_ranks.alloc(ctx, n1));
  PYN_CHECK(h->d_peer_win.alloc(ctx, n1));
  PYN_CHECK(h->d_peer_off.upload(h->peer_off.data(), h->nneigh));
  PYN_CHECK(h->d_peer_nghost.upload(h->peer_nghost.data(), h->nneigh));
  PYN_CHECK(h->d_recv_off.upload(h->recv_off.data(), h->nneigh));
  PYN_CHECK(h->d_recv_cnt.upload(h->recv_cnt.data(), h->nneigh));
  PYN_CHECK(h->d_ranks.upload(ranks.data(), h->nneigh));
  PYN_CHECK(h->d_peer_win.upload(addr.data(), h->nneigh));
  PYN_CUDA(cudaStreamSynchronize(ctx->stream));
  h->p2p = true;
  return PYN_OK;
}

int pyn_comm_p2p_status(pyn_ctx* ctx, int* enabled, int* timed_out) {
  PYN_REQUIRE(ctx, "null ctx");
  if (enabled) *enabled = ctx->p2p_on ? 1 : 0;
  if (timed_out) {
    *timed_out = 0;
    if (ctx->p2p_timeout) {
      PYN_CUDA(cudaMemcpyAsync(timed_out, ctx->p2p_timeout, sizeof(int), cudaMemcpyDeviceToHost, ctx->stream));
      PYN_CUDA(cudaStreamSynchronize(ctx->stream));
    }
  }
  return PYN_OK;
}

int pyn_halo_exchange(pyn_halo* h, pyn_vec* v, int bs) {
  PYN_REQUIRE(h && v && bs > 0 && v->n == h->n_loc * bs, "pyn_halo_exchange: Vec has %lld entries, layout %lld x %d",
              v ? (long long)v->n : -1LL, h ? (long long)h->n_loc : -1LL, bs);
  return halo_exchange(h, v->d.p, bs);
}

}  // extern "C"
