// api.cu -- extern "C" entry points: context, mesh, element export, assembly.
#include "elem.cuh"
#include "classes.cuh"
#include "spmv.cuh"
#include "vec.cuh"

namespace pyn {
const char* get_error();
int vtensv(pyn_ctx* ctx, const double* vel, double* out, int64_t nnodes, int dim);
int stress_combine(pyn_ctx* ctx, double* aux, const double* vel, double mu, double rho,
                   int64_t nnodes, int dim);

struct PhaseTimer {
  pyn_ctx* ctx;
  cudaEvent_t a, b;
  PhaseTimer(pyn_ctx* c) : ctx(c) {
    cudaEventCreate(&a);
    cudaEventCreate(&b);
    cudaEventRecord(a, ctx->stream);
  }
  double stop() {
    cudaEventRecord(b, ctx->stream);
    cudaEventSynchronize(b);
    float ms = 0;
    cudaEventElapsedTime(&ms, a, b);
    return ms;
  }
  ~PhaseTimer() {
    cudaEventDestroy(a);
    cudaEventDestroy(b);
  }
};

// default matching tolerance of the congruence classes, relative to the largest element
// extent: far above the rounding of generated coordinates (1e-16), far below the 1e-12
// value parity of the assembled matrices
constexpr double kDefaultCongruenceTol = 1e-13;
constexpr int kProfPairs = 4096;
void prof_drain(pyn_ctx* ctx) {
  if (ctx->prof_pending == 0) return;
  cudaEventSynchronize(ctx->prof_ev[2 * (ctx->prof_pending - 1) + 1]);
  for (int i = 0; i < ctx->prof_pending; ++i) {
    float ms = 0;
    if (cudaEventElapsedTime(&ms, ctx->prof_ev[2 * i], ctx->prof_ev[2 * i + 1]) == cudaSuccess) {
      ctx->prof_ms[ctx->prof_cls[i]] += ms;
      ctx->prof_count[ctx->prof_cls[i]]++;
    }
  }
  ctx->prof_pending = 0;
}
void prof_begin(pyn_ctx* ctx, int cls) {
  if (!ctx->prof_on) return;
  if (ctx->prof_ev.empty()) {
    ctx->prof_ev.resize(2 * kProfPairs);
    ctx->prof_cls.resize(kProfPairs);
    for (auto& e : ctx->prof_ev) cudaEventCreate(&e);
  }
  if (ctx->prof_pending == kProfPairs) prof_drain(ctx);
  ctx->prof_cls[ctx->prof_pending] = cls;
  cudaEventRecord(ctx->prof_ev[2 * ctx->prof_pending], ctx->stream);
}
void prof_end(pyn_ctx* ctx) {
  if (!ctx->prof_on) return;
  cudaEventRecord(ctx->prof_ev[2 * ctx->prof_pending + 1], ctx->stream);
  ctx->prof_pending++;
}
}  // namespace pyn

using namespace pyn;

extern "C" {

int pyn_prof_enable(pyn_ctx* ctx, int on) {
  PYN_REQUIRE(ctx, "null ctx");
  prof_drain(ctx);
  ctx->prof_on = on != 0;
  if (on)
    for (int i = 0; i < PROF_COUNT; ++i) {
      ctx->prof_ms[i] = 0.0;
      ctx->prof_count[i] = 0;
    }
  return PYN_OK;
}
int pyn_prof_get(pyn_ctx* ctx, int cls, int64_t* count, double* total_ms) {
  PYN_REQUIRE(ctx && cls >= 0 && cls < PROF_COUNT && count && total_ms, "pyn_prof_get: bad argument");
  prof_drain(ctx);
  *count = ctx->prof_count[cls];
  *total_ms = ctx->prof_ms[cls];
  return PYN_OK;
}
int pyn_tune(const char* key, int value) {
  PYN_REQUIRE(key, "pyn_tune: null key");
  if (strcmp(key, "ksp_check_every") == 0 && value > 0) {
    g_ksp_check_every = value;
    return PYN_OK;
  }
  if (strcmp(key, "asm_coarse_unfused") == 0) {   // coarse level of pc asm as separate launches (comparison)
    g_asm_coarse_unfused = value != 0;
    return PYN_OK;
  }
  if (strcmp(key, "halo_wait_at_start") == 0) {   // fused halo: 1 wait at CTA start (default), 0 per warp before its first boundary row
    g_halo_wait_at_start = value != 0;
    return PYN_OK;
  }
  if (strcmp(key, "ksp_adaptive_chunks") == 0) {  // 0: fixed chunks of ksp_check_every iterations (round 1)
    g_ksp_adaptive_chunks = value != 0;
    return PYN_OK;
  }
  if (strcmp(key, "ksp_graphs") == 0) {            // 0: plain kernel launches in the CG loop (no CUDA graphs)
    g_ksp_graphs = value != 0;
    return PYN_OK;
  }
  if (strcmp(key, "force_idx64") == 0) {          // tests: take the 64-bit index instantiations on small meshes
    g_force_idx64 = value != 0;
    return PYN_OK;
  }
  set_error("pyn_tune: unknown key '%s'", key);
  return PYN_E_INVALID;
}
int pyn_host_alloc(int64_t bytes, void** out) {
  PYN_REQUIRE(out && bytes > 0, "pyn_host_alloc: bad argument");
  PYN_CUDA(cudaMallocHost(out, (size_t)bytes));
  return PYN_OK;
}
int pyn_host_free(void* p) {
  if (p) PYN_CUDA(cudaFreeHost(p));
  return PYN_OK;
}

int pyn_version(void) { return 100; }
const char* pyn_last_error(void) { return pyn::get_error(); }

int pyn_rule_1d(int kind, int n, double* x, double* w) { return rule_1d(kind, n, x, w); }
int pyn_lagrange_1d(int nnodes, const double* nodes, int npts, const double* pts, double* h, double* dh) {
  return lagrange_1d(nnodes, nodes, npts, pts, h, dh);
}
int pyn_local_order(int dim, int n, int32_t* out) { return local_order(dim, n, out); }

// ---- context ---------------------------------------------------------------------
int pyn_init(int device, pyn_ctx** out) {
  PYN_REQUIRE(out, "pyn_init: null out");
  int ndev = 0;
  cudaError_t e = cudaGetDeviceCount(&ndev);
  if (e != cudaSuccess || ndev == 0) {
    set_error("pyn_init: no CUDA device available (%s); this library has no CPU path",
              e != cudaSuccess ? cudaGetErrorString(e) : "0 devices");
    return PYN_E_CUDA;
  }
  PYN_REQUIRE(device >= 0 && device < ndev, "pyn_init: device %d of %d", device, ndev);
  PYN_CUDA(cudaSetDevice(device));
  pyn_ctx* ctx = new pyn_ctx();
  ctx->device = device;
  cudaDeviceProp prop;
  PYN_CUDA(cudaGetDeviceProperties(&prop, device));
  ctx->num_sms = prop.multiProcessorCount;
  ctx->smem_optin = prop.sharedMemPerBlockOptin;
  PYN_CUDA(cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking));
  {
    cudaMemPool_t pool;
    unsigned long long keep = ~0ull;   // never trim: assembly scratch is reused by the next phase
    PYN_CUDA(cudaDeviceGetDefaultMemPool(&pool, device));
    PYN_CUDA(cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &keep));
    // first use creates the pool (tens of ms): pay it here, with the CUDA context, and keep
    // 256 MB mapped for the vectors and sort scratch of the set-up phase
    void* warm = nullptr;
    PYN_CUDA(cudaMallocAsync(&warm, (size_t)256 << 20, ctx->stream));
    PYN_CUDA(cudaFreeAsync(warm, ctx->stream));
    PYN_CUDA(cudaStreamSynchronize(ctx->stream));
  }
  PYN_CUDA(cudaEventCreate(&ctx->ev0));
  PYN_CUDA(cudaEventCreate(&ctx->ev1));
  PYN_CUDA(cudaMalloc((void**)&ctx->red_partials, sizeof(double) * kRedMaxBlocks * kRedMaxVals));
  PYN_CUDA(cudaMalloc((void**)&ctx->red_ticket, sizeof(unsigned int)));
  PYN_CUDA(cudaMemset(ctx->red_ticket, 0, sizeof(unsigned int)));
  PYN_CUDA(cudaMalloc((void**)&ctx->red_result_dev, sizeof(double) * kRedMaxVals));
  PYN_CUDA(cudaMallocHost((void**)&ctx->red_result_host, sizeof(double) * kRedMaxVals));
  *out = ctx;
  return PYN_OK;
}

int pyn_finalize(pyn_ctx* ctx) {
  if (!ctx) return PYN_OK;
  cudaSetDevice(ctx->device);
  cudaStreamSynchronize(ctx->stream);
  prof_drain(ctx);
  for (auto& e : ctx->prof_ev) cudaEventDestroy(e);
  if (ctx->flush_buf) cudaFree(ctx->flush_buf);
  cudaFree(ctx->red_partials);
  cudaFree(ctx->red_ticket);
  cudaFree(ctx->red_result_dev);
  cudaFreeHost(ctx->red_result_host);
  cudaEventDestroy(ctx->ev0);
  cudaEventDestroy(ctx->ev1);
  cudaStreamDestroy(ctx->stream);
  delete ctx;
  return PYN_OK;
}

int pyn_sync(pyn_ctx* ctx) {
  PYN_REQUIRE(ctx, "null ctx");
  PYN_CUDA(cudaStreamSynchronize(ctx->stream));
  return PYN_OK;
}
int pyn_timer_start(pyn_ctx* ctx) {
  PYN_REQUIRE(ctx, "null ctx");
  PYN_CUDA(cudaEventRecord(ctx->ev0, ctx->stream));
  return PYN_OK;
}
int pyn_timer_stop(pyn_ctx* ctx, double* ms) {
  PYN_REQUIRE(ctx && ms, "null argument");
  PYN_CUDA(cudaEventRecord(ctx->ev1, ctx->stream));
  PYN_CUDA(cudaEventSynchronize(ctx->ev1));
  float f = 0;
  PYN_CUDA(cudaEventElapsedTime(&f, ctx->ev0, ctx->ev1));
  *ms = f;
  return PYN_OK;
}
int pyn_launch_count(pyn_ctx* ctx, int64_t* count) {
  PYN_REQUIRE(ctx && count, "null argument");
  *count = ctx->launches;
  return PYN_OK;
}
int pyn_flush_l2(pyn_ctx* ctx) {
  PYN_REQUIRE(ctx, "null ctx");
  if (!ctx->flush_buf) {
    ctx->flush_bytes = (size_t)256 << 20;  // 2x the 126 MB L2
    PYN_CUDA(cudaMalloc(&ctx->flush_buf, ctx->flush_bytes));
  }
  PYN_CHECK(vec_set(ctx, (double*)ctx->flush_buf, (int64_t)(ctx->flush_bytes / 8), 0.0));
  return PYN_OK;
}
int pyn_mem_used(pyn_ctx* ctx, int64_t* bytes) {
  PYN_REQUIRE(ctx && bytes, "null argument");
  *bytes = ctx->bytes_in_use;
  return PYN_OK;
}

// ---- mesh ------------------------------------------------------------------------
// Congruence classes of element shapes (classes.cuh) -> elem_class / class_rep on the device.
static int compute_classes(pyn_mesh* m, const double* corners, double tol) {
  m->congruence_tol = tol;
  m->class_version++;
  m->elem_class.release();
  m->class_rep.release();
  std::vector<int32_t> cls, rep;
  m->nclass = congruence_classes(m->dim, m->nel, corners, tol, &cls, &rep);
  if (m->nclass == m->nel) return PYN_OK;          // nothing merged: identity, no tables
  PYN_CHECK(m->elem_class.alloc(m->ctx, m->nel));
  PYN_CHECK(m->elem_class.upload(cls.data(), m->nel));
  PYN_CHECK(m->class_rep.alloc(m->ctx, m->nclass));
  PYN_CHECK(m->class_rep.upload(rep.data(), m->nclass));
  return PYN_OK;
}

int pyn_mesh_set_congruence(pyn_mesh* mesh, double tol, int64_t* nclass) {
  PYN_REQUIRE(mesh, "pyn_mesh_set_congruence: null mesh");
  const int64_t n = mesh->nel * (1 << mesh->dim) * mesh->dim;
  std::vector<double> h((size_t)n);
  PYN_CHECK(mesh->corners.download(h.data(), n));
  PYN_CHECK(compute_classes(mesh, h.data(), tol));
  if (nclass) *nclass = mesh->nclass;
  return PYN_OK;
}

int pyn_mesh_get_congruence(pyn_mesh* mesh, double* tol, int64_t* nclass) {
  PYN_REQUIRE(mesh, "pyn_mesh_get_congruence: null mesh");
  if (tol) *tol = mesh->congruence_tol;
  if (nclass) *nclass = mesh->nclass;
  return PYN_OK;
}

int pyn_mesh_create(pyn_ctx* ctx, int dim, int ngl, int64_t nel, int64_t nnodes,
                    const double* corner_xyz, const int32_t* elem_nodes, pyn_mesh** out) {
  PYN_REQUIRE(ctx && corner_xyz && elem_nodes && out, "pyn_mesh_create: null argument");
  PYN_REQUIRE(nel > 0 && nnodes > 0, "pyn_mesh_create: empty mesh");
  pyn_mesh* m = new pyn_mesh();
  m->ctx = ctx;
  int rc = build_elem_tables(dim, ngl, &m->htab);
  if (rc != PYN_OK) {
    delete m;
    return rc;
  }
  m->dim = dim;
  m->ngl = ngl;
  m->nn = m->htab.nn;
  m->dim_w = m->htab.dim_w;
  m->dim_s = m->htab.dim_s;
  m->nel = nel;
  m->nnodes = nnodes;
  const int nc = 1 << dim;
  for (int64_t i = 0; i < nel * m->nn; ++i) {
    if (elem_nodes[i] < 0 || elem_nodes[i] >= nnodes) {
      set_error("pyn_mesh_create: elem_nodes[%lld] = %d outside [0, %lld)", (long long)i,
                elem_nodes[i], (long long)nnodes);
      delete m;
      return PYN_E_INVALID;
    }
  }
  PhaseTimer t(ctx);
  rc = m->dtab.alloc(ctx, 1);
  if (rc == PYN_OK) rc = m->dtab.upload(&m->htab, 1);
  if (rc == PYN_OK) rc = m->corners.alloc(ctx, nel * nc * dim);
  if (rc == PYN_OK) rc = m->corners.upload(corner_xyz, nel * nc * dim);
  if (rc == PYN_OK) rc = m->elem_nodes.alloc(ctx, nel * m->nn);
  if (rc == PYN_OK) rc = m->elem_nodes.upload(elem_nodes, nel * m->nn);
  if (rc == PYN_OK) rc = m->dofcls.alloc(ctx, nnodes * dim);
  if (rc == PYN_OK) rc = m->dofcls.zero();
  if (rc == PYN_OK) rc = compute_classes(m, corner_xyz, kDefaultCongruenceTol);
  if (rc == PYN_OK) rc = build_n2e(m);
  if (rc == PYN_OK) {
    std::vector<int32_t> pb((size_t)m->nn * m->nn);
    for (int a = 0; a < m->nn; ++a)
      for (int b = 0; b < m->nn; ++b) pb[(size_t)a * m->nn + b] = b;
    rc = build_graph(m, &m->dense, pb, m->nn);
  }
  if (rc == PYN_OK && cudaStreamSynchronize(ctx->stream) != cudaSuccess) {
    set_error("pyn_mesh_create: %s", cudaGetErrorString(cudaGetLastError()));
    rc = PYN_E_CUDA;
  }
  if (rc != PYN_OK) {
    delete m;
    return rc;
  }
  m->phase_ms[0] = t.stop();
  *out = m;
  return PYN_OK;
}

int pyn_mesh_destroy(pyn_mesh* mesh) {
  delete mesh;
  return PYN_OK;
}

int pyn_mesh_graph_info(pyn_mesh* mesh, int64_t* npairs, int64_t* ncontrib, int* max_row) {
  PYN_REQUIRE(mesh, "null mesh");
  if (npairs) *npairs = mesh->dense.nslots;
  if (ncontrib) *ncontrib = mesh->dense.ncontrib;
  if (max_row) *max_row = mesh->dense.max_row;
  return PYN_OK;
}

int pyn_mesh_set_global_tables(pyn_mesh* mesh, int64_t n_global_nodes, int64_t nel_global,
                               const int32_t* elem_nodes_global, const int64_t* elem_gid) {
  PYN_REQUIRE(mesh && elem_nodes_global && elem_gid && nel_global >= mesh->nel && n_global_nodes > 0,
              "pyn_mesh_set_global_tables: bad argument");
  for (int64_t e = 0; e < mesh->nel; ++e)
    PYN_REQUIRE(elem_gid[e] >= 0 && elem_gid[e] < nel_global, "pyn_mesh_set_global_tables: element id %lld of %lld",
                (long long)elem_gid[e], (long long)nel_global);
  mesh->g_elem_nodes.assign(elem_nodes_global, elem_nodes_global + nel_global * mesh->nn);
  mesh->elem_gid.assign(elem_gid, elem_gid + mesh->nel);
  mesh->g_nel = nel_global;
  mesh->g_nnodes = n_global_nodes;
  return PYN_OK;
}

int pyn_mesh_set_dof_class(pyn_mesh* mesh, const uint8_t* cls) {
  PYN_REQUIRE(mesh && cls, "null argument");
  const int64_t n = mesh->nnodes * mesh->dim;
  for (int64_t i = 0; i < n; ++i) PYN_REQUIRE(cls[i] < 4, "dof class %d at %lld", cls[i], (long long)i);
  PYN_CHECK(mesh->dofcls.upload(cls, n));
  std::vector<uint8_t> pack((size_t)mesh->nnodes, 0);
  for (int64_t i = 0; i < n; ++i) pack[(size_t)(i / mesh->dim)] |= (uint8_t)(cls[i] << (2 * (i % mesh->dim)));
  PYN_CHECK(mesh->nodecls.alloc(mesh->ctx, mesh->nnodes));
  PYN_CHECK(mesh->nodecls.upload(pack.data(), mesh->nnodes));
  PYN_CUDA(cudaStreamSynchronize(mesh->ctx->stream));
  mesh->has_cls = true;
  return PYN_OK;
}

int pyn_mesh_phase_ms(pyn_mesh* mesh, int phase, double* ms) {
  PYN_REQUIRE(mesh && ms && phase >= 0 && phase < 8, "bad argument");
  *ms = mesh->phase_ms[phase];
  return PYN_OK;
}

// ---- element export (parity / debug) ------------------------------------------------
// block-major EBS -> the reference's dense row-major element matrix
static void ebs_to_dense(const double* ebs, int stride, int off, int64_t ne, int nn, int br, int bc,
                         double* dense) {
  const int64_t ld = (int64_t)nn * bc;
  for (int64_t e = 0; e < ne; ++e)
    for (int a = 0; a < nn; ++a)
      for (int b = 0; b < nn; ++b) {
        const double* blk = ebs + (((int64_t)e * nn + a) * nn + b) * stride + off;
        for (int i = 0; i < br; ++i)
          for (int j = 0; j < bc; ++j)
            dense[e * (int64_t)nn * br * ld + ((int64_t)a * br + i) * ld + (int64_t)b * bc + j] =
                blk[i * bc + j];
      }
}

int pyn_elem_kle(pyn_mesh* mesh, int64_t e0, int64_t ne, double* K_e, double* Rw_e, double* Rd_e) {
  PYN_REQUIRE(mesh && e0 >= 0 && ne >= 0 && e0 + ne <= mesh->nel, "pyn_elem_kle: element range");
  if (ne == 0) return PYN_OK;
  pyn_ctx* ctx = mesh->ctx;
  const int nn = mesh->nn, d = mesh->dim, dw = mesh->dim_w;
  const int64_t np = ne * nn * nn;
  const int rs = kle_record_stride(d);
  DevBuf<double> ebs;
  PYN_CHECK(ebs.alloc(ctx, np * rs));
  PYN_CHECK(launch_elem_kle(mesh, e0, ne, nullptr, ebs.p));
  std::vector<double> h((size_t)(np * rs));
  PYN_CHECK(ebs.download(h.data(), np * rs));
  if (K_e) ebs_to_dense(h.data(), rs, 0, ne, nn, d, d, K_e);
  if (Rw_e) ebs_to_dense(h.data(), rs, kle_record_off_rw(d), ne, nn, d, dw, Rw_e);
  if (Rd_e) ebs_to_dense(h.data(), rs, kle_record_off_rd(d), ne, nn, d, 1, Rd_e);
  PYN_CUDA(cudaStreamSynchronize(ctx->stream));
  return PYN_OK;
}

static void stencil_to_dense(const double* ebs, const std::vector<int32_t>& pair_b, int64_t ne, int nn,
                             int ppa, int br, int bc, double* dense) {
  const int64_t ld = (int64_t)nn * bc, esz = (int64_t)nn * br * ld;
  memset(dense, 0, sizeof(double) * ne * esz);
  for (int64_t e = 0; e < ne; ++e)
    for (int a = 0; a < nn; ++a)
      for (int m = 0; m < ppa; ++m) {
        const int b = pair_b[(size_t)a * ppa + m];
        const double* blk = ebs + ((e * nn + a) * ppa + m) * br * bc;
        for (int i = 0; i < br; ++i)
          for (int j = 0; j < bc; ++j)
            dense[e * esz + ((int64_t)a * br + i) * ld + (int64_t)b * bc + j] = blk[i * bc + j];
      }
}

int pyn_elem_ops(pyn_mesh* mesh, int64_t e0, int64_t ne, double* SrT_e, double* Div_e, double* Curl_e,
                 double* W_e) {
  PYN_REQUIRE(mesh && e0 >= 0 && ne >= 0 && e0 + ne <= mesh->nel, "pyn_elem_ops: element range");
  if (ne == 0) return PYN_OK;
  pyn_ctx* ctx = mesh->ctx;
  PYN_CHECK(ensure_stencil(mesh));
  const int nn = mesh->nn, d = mesh->dim, dw = mesh->dim_w, ds = mesh->dim_s;
  const int ppa = mesh->stencil.ppa;
  const int64_t np = ne * nn * ppa;
  DevBuf<double> s, dv, c, w;
  PYN_CHECK(s.alloc(ctx, np * ds * d));
  PYN_CHECK(dv.alloc(ctx, np * ds * d));
  PYN_CHECK(c.alloc(ctx, np * dw * d));
  PYN_CHECK(w.alloc(ctx, ne * nn));
  PYN_CHECK(launch_elem_ops(mesh, e0, ne, nullptr, s.p, dv.p, c.p, w.p));
  std::vector<int32_t> pb((size_t)nn * ppa);
  PYN_CHECK(mesh->stencil.pair_b.download(pb.data(), (int64_t)nn * ppa));
  std::vector<double> h;
  if (SrT_e) {
    h.resize(np * ds * d);
    PYN_CHECK(s.download(h.data(), np * ds * d));
    stencil_to_dense(h.data(), pb, ne, nn, ppa, ds, d, SrT_e);
  }
  if (Div_e) {
    h.resize(np * ds * d);
    PYN_CHECK(dv.download(h.data(), np * ds * d));
    stencil_to_dense(h.data(), pb, ne, nn, ppa, d, ds, Div_e);
  }
  if (Curl_e) {
    h.resize(np * dw * d);
    PYN_CHECK(c.download(h.data(), np * dw * d));
    stencil_to_dense(h.data(), pb, ne, nn, ppa, dw, d, Curl_e);
  }
  if (W_e) PYN_CHECK(w.download(W_e, ne * nn));
  PYN_CUDA(cudaStreamSynchronize(ctx->stream));
  return PYN_OK;
}

// ---- assembly ----------------------------------------------------------------------
static MatSpec kle_spec(pyn_mesh* mesh, int br, int bc, const double* ebs, int off, double sign, bool colcls,
                        uint32_t pairmask, uint32_t diagmask, double d1, double d2, double d3) {
  MatSpec s;
  s.br = br;
  s.bc = bc;
  s.graph = &mesh->dense;
  s.ebs = ebs;
  s.ebs_stride = kle_record_stride(mesh->dim);
  s.ebs_off = off;
  s.ebs_pairs = (int64_t)mesh->nn * mesh->nn;
  s.ebs_ppa = 0;
  s.pair_map = nullptr;
  s.sign = sign;
  s.use_rowcls = true;
  s.use_colcls = colcls;
  s.pairmask = pairmask;
  s.diagmask = diagmask;
  s.diag_add[0] = 0.0;
  s.diag_add[1] = d1;
  s.diag_add[2] = d2;
  s.diag_add[3] = d3;
  s.row_scale = nullptr;
  s.ref_dense = true;
  return s;
}

#define PM(rc, cc) (1u << ((rc)*4 + (cc)))

int pyn_assemble_kle(pyn_mesh* mesh, int mode, pyn_mat** out8) {
  PYN_REQUIRE(mesh && out8, "null argument");
  PYN_REQUIRE(mode == 0 || mode == 1, "mode must be 0 (FS) or 1 (NS)");
  PYN_REQUIRE(mesh->has_cls, "pyn_assemble_kle: set the dof classes first");
  pyn_ctx* ctx = mesh->ctx;
  const int d = mesh->dim, dw = mesh->dim_w, nn = mesh->nn;
  const int64_t np = mesh->nclass * nn * nn;       // one set of blocks per congruence class
  for (int i = 0; i < 8; ++i) out8[i] = nullptr;
  const int oK = 0, oW = kle_record_off_rw(d), oD = kle_record_off_rd(d);
  DevBuf<double> ebs;
  PYN_CHECK(ebs.alloc(ctx, np * kle_record_stride(d)));
  {
    PhaseTimer t(ctx);
    PYN_CHECK(launch_elem_kle(mesh, 0, mesh->nclass, mesh->class_rep.p, ebs.p));
    mesh->phase_ms[1] = t.stop();
  }
  PhaseTimer t(ctx);
  const uint32_t bc_all = 0xE;  // classes 1,2,3
  // K: free x free, +1 on every constrained diagonal (mat_fs.py:179-189, mat_ns.py:116-130)
  MatSpec sK = kle_spec(mesh, d, d, ebs.p, oK, 1.0, true, PM(0, 0), bc_all, 1.0, 1.0, 1.0);
  // Krhs: -K_e[free, constrained], +1 on the constrained diagonal (mat_fs.py:174-177,188-189)
  MatSpec sKrhs = kle_spec(mesh, d, d, ebs.p, oK, -1.0, true, PM(0, 1) | PM(0, 2) | PM(0, 3), bc_all, 1.0, 1.0, 1.0);
  // Rw: free rows, all vorticity columns (mat_fs.py:185-186)
  MatSpec sRw = kle_spec(mesh, d, dw, ebs.p, oW, 1.0, false, PM(0, 0), 0, 0, 0, 0);
  // Rd: filled only by buildNS (mat_ns.py:125-127); empty in FS mode (mat_fs.py:155)
  MatSpec sRd = kle_spec(mesh, d, 1, ebs.p, oD, 1.0, false, mode == 1 ? PM(0, 0) : 0u, 0, 0, 0, 0);
  static_assert(PYN_MAT_K == 0 && PYN_MAT_KRHS == 1 && PYN_MAT_RW == 2 && PYN_MAT_RD == 3, "matrix order");
  if (getenv("PYN_ASSEMBLE_PER_MATRIX") == nullptr) {
    const MatSpec four[4] = {sK, sKrhs, sRw, sRd};
    PYN_CHECK(assemble_kle_fused(mesh, four, out8));
  } else {                                         // one gather pass per matrix (debug / comparison)
    PYN_CHECK(assemble_matrix(mesh, sK, &out8[PYN_MAT_K]));
    PYN_CHECK(assemble_matrix(mesh, sKrhs, &out8[PYN_MAT_KRHS]));
    PYN_CHECK(assemble_matrix(mesh, sRw, &out8[PYN_MAT_RW]));
    PYN_CHECK(assemble_matrix(mesh, sRd, &out8[PYN_MAT_RD]));
  }
  if (mode == 1) {
    // Kfs: tang x free, free x tang, tang x tang, -1 on tang diagonal (mat_ns.py:86-97,132-133)
    MatSpec sKfs = kle_spec(mesh, d, d, ebs.p, oK, 1.0, true, PM(3, 0) | PM(0, 3) | PM(3, 3), 1u << 3, 0, 0, -1.0);
    // Krhsfs: -K_e[tang|free, normal], 1 on the normal diagonal (mat_ns.py:107-114,140-142)
    MatSpec sKrhsfs = kle_spec(mesh, d, d, ebs.p, oK, -1.0, true, PM(3, 2) | PM(0, 2), 1u << 2, 0, 1.0, 0);
    MatSpec sRwfs = kle_spec(mesh, d, dw, ebs.p, oW, 1.0, false, PM(3, 0), 0, 0, 0, 0);
    MatSpec sRdfs = kle_spec(mesh, d, 1, ebs.p, oD, 1.0, false, PM(3, 0), 0, 0, 0, 0);
    static_assert(PYN_MAT_KFS == 4 && PYN_MAT_KRHSFS == 5 && PYN_MAT_RWFS == 6 && PYN_MAT_RDFS == 7, "matrix order");
    if (getenv("PYN_ASSEMBLE_PER_MATRIX") == nullptr) {
      const MatSpec four[4] = {sKfs, sKrhsfs, sRwfs, sRdfs};
      PYN_CHECK(assemble_kle_fused(mesh, four, out8 + 4));
    } else {
      PYN_CHECK(assemble_matrix(mesh, sKfs, &out8[PYN_MAT_KFS]));
      PYN_CHECK(assemble_matrix(mesh, sKrhsfs, &out8[PYN_MAT_KRHSFS]));
      PYN_CHECK(assemble_matrix(mesh, sRwfs, &out8[PYN_MAT_RWFS]));
      PYN_CHECK(assemble_matrix(mesh, sRdfs, &out8[PYN_MAT_RDFS]));
    }
  }
  mesh->phase_ms[2] = t.stop();
  PYN_CUDA(cudaStreamSynchronize(ctx->stream));
  return PYN_OK;
}

int pyn_assemble_ops(pyn_mesh* mesh, int compact, pyn_mat** curl, pyn_mat** srt, pyn_mat** divsrt,
                     pyn_vec** weights) {
  PYN_REQUIRE(mesh && curl && srt && divsrt, "null argument");
  pyn_ctx* ctx = mesh->ctx;
  PYN_CHECK(ensure_stencil(mesh));
  const int d = mesh->dim, dw = mesh->dim_w, ds = mesh->dim_s, nn = mesh->nn;
  const int ppa = mesh->stencil.ppa;
  const int64_t np = mesh->nclass * nn * ppa;
  DevBuf<double> s, dv, c, we, winv;
  PYN_CHECK(s.alloc(ctx, np * ds * d));
  PYN_CHECK(dv.alloc(ctx, np * ds * d));
  PYN_CHECK(c.alloc(ctx, np * dw * d));
  PYN_CHECK(we.alloc(ctx, mesh->nclass * nn));
  PYN_CHECK(winv.alloc(ctx, mesh->nnodes));
  {
    PhaseTimer t(ctx);
    PYN_CHECK(launch_elem_ops(mesh, 0, mesh->nclass, mesh->class_rep.p, s.p, dv.p, c.p, we.p));
    mesh->phase_ms[3] = t.stop();
  }
  PhaseTimer t(ctx);
  pyn_vec* w = new pyn_vec();
  w->ctx = ctx;
  w->n = mesh->nnodes;
  int rc = w->d.alloc(ctx, mesh->nnodes);
  if (rc == PYN_OK) rc = assemble_node_weights(mesh, we.p, w->d.p, winv.p);
  MatSpec sp;
  sp.graph = compact ? &mesh->stencil : &mesh->dense;
  sp.ebs_pairs = (int64_t)nn * ppa;
  sp.pair_map = compact ? nullptr : mesh->stencil_of_dense.p;
  sp.sign = 1.0;
  sp.use_rowcls = sp.use_colcls = false;
  sp.pairmask = PM(0, 0);
  sp.diagmask = 0;
  for (int i = 0; i < 4; ++i) sp.diag_add[i] = 0.0;
  sp.row_scale = winv.p;
  sp.ref_dense = true;
  *curl = *srt = *divsrt = nullptr;
  if (rc == PYN_OK) {
    sp.br = dw; sp.bc = d; sp.ebs = c.p;
    rc = assemble_matrix(mesh, sp, curl);
  }
  if (rc == PYN_OK) {
    sp.br = ds; sp.bc = d; sp.ebs = s.p;
    rc = assemble_matrix(mesh, sp, srt);
  }
  if (rc == PYN_OK) {
    sp.br = d; sp.bc = ds; sp.ebs = dv.p;
    rc = assemble_matrix(mesh, sp, divsrt);
  }
  mesh->phase_ms[4] = t.stop();
  if (rc == PYN_OK && cudaStreamSynchronize(ctx->stream) != cudaSuccess) rc = PYN_E_CUDA;
  if (rc != PYN_OK) {
    delete w;
    return rc;
  }
  if (weights)
    *weights = w;
  else
    delete w;
  return PYN_OK;
}

}  // extern "C"
