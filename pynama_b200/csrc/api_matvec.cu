// api_matvec.cu -- extern "C" entry points for Mat and Vec objects.
#include <algorithm>

#include "elem.cuh"
#include "spmv.cuh"
#include "matfree.cuh"
#include "vec.cuh"
#include "comm.cuh"

namespace pyn {
int vtensv(pyn_ctx* ctx, const double* vel, double* out, int64_t nnodes, int dim);
int stress_combine(pyn_ctx* ctx, double* aux, const double* vel, double mu, double rho,
                   int64_t nnodes, int dim);

// Host image of a matrix + its reference graph, used by export / nnz.
struct HostMat {
  std::vector<int64_t> rowptr, g_rowptr;
  std::vector<int32_t> colidx, g_colidx;
  std::vector<double> val;
  std::vector<uint8_t> cls;
};

static int fetch_host(pyn_mat* m, HostMat* h, bool with_values) {
  h->rowptr.resize(m->nbrows + 1);
  h->colidx.resize(m->nblocks);
  PYN_CHECK(m->rowptr.download(h->rowptr.data(), m->nbrows + 1));
  PYN_CHECK(m->colidx.download(h->colidx.data(), m->nblocks));
  if (m->free_pattern) {                 // reference pattern = the matrix's own blocks
    if (with_values) {
      h->val.resize(m->nblocks * m->br * m->bc);
      PYN_CHECK(m->val.download(h->val.data(), m->nblocks * m->br * m->bc));
    }
    return PYN_OK;
  }
  const NodeGraph* g = &m->mesh->dense;  // reference pattern = dense node graph (x) block
  h->g_rowptr.resize(g->nrows + 1);
  h->g_colidx.resize(g->nslots);
  PYN_CHECK(g->rowptr.download(h->g_rowptr.data(), g->nrows + 1));
  PYN_CHECK(g->colidx.download(h->g_colidx.data(), g->nslots));
  if (with_values) {
    h->val.resize(m->nblocks * m->br * m->bc);
    PYN_CHECK(m->val.download(h->val.data(), m->nblocks * m->br * m->bc));
  }
  if (m->use_rowcls || m->use_colcls) {
    h->cls.resize(m->mesh->nnodes * m->mesh->dim);
    PYN_CHECK(m->mesh->dofcls.download(h->cls.data(), (int64_t)h->cls.size()));
  }
  return PYN_OK;
}

// does scalar entry (i,j) of node block (r,c) exist in the reference's AIJ pattern?
static inline bool ref_entry(const pyn_mat* m, const HostMat& h, int64_t r, int32_t c, int i, int j) {
  const int rc = m->use_rowcls ? h.cls[r * m->br + i] : 0;
  const int cc = m->use_colcls ? h.cls[(int64_t)c * m->bc + j] : 0;
  if ((m->pairmask >> (rc * 4 + cc)) & 1u) return true;
  return m->br == m->bc && r == c && i == j && ((m->diagmask >> rc) & 1u);
}

// Walk the reference pattern; fn(row, col, value) in CSR order.
template <typename F>
static void walk_reference(const pyn_mat* m, const HostMat& h, bool with_values, F fn) {
  const int br = m->br, bc = m->bc, bb = br * bc;
  if (m->free_pattern) {
    for (int64_t r = 0; r < m->nbrows; ++r) {
      const int64_t p0 = h.rowptr[r], nb = h.rowptr[r + 1] - p0;
      for (int i = 0; i < br; ++i)
        for (int64_t u = 0; u < nb; ++u)
          for (int j = 0; j < bc; ++j) {
            if (m->diag_only && i != j) continue;
            fn(r * br + i, (int64_t)h.colidx[p0 + u] * bc + j,
               with_values ? h.val[p0 * bb + (int64_t)(i * bc + j) * nb + u] : 0.0);
          }
    }
    return;
  }
  for (int64_t r = 0; r < m->nbrows; ++r) {
    const int64_t p0 = h.rowptr[r], nb = h.rowptr[r + 1] - p0;
    for (int i = 0; i < br; ++i) {
      int64_t u = 0;
      for (int64_t s = h.g_rowptr[r]; s < h.g_rowptr[r + 1]; ++s) {
        const int32_t c = h.g_colidx[s];
        while (u < nb && h.colidx[p0 + u] < c) ++u;
        const bool have = u < nb && h.colidx[p0 + u] == c;
        for (int j = 0; j < bc; ++j) {
          if (!ref_entry(m, h, r, c, i, j)) continue;
          double v = 0.0;
          if (have && with_values) v = h.val[p0 * bb + (int64_t)(i * bc + j) * nb + u];
          fn(r * br + i, (int64_t)c * bc + j, v);
        }
      }
    }
  }
}
}  // namespace pyn

using namespace pyn;

extern "C" {

// ---- Mat ---------------------------------------------------------------------------
int pyn_mat_shape(pyn_mat* m, int64_t* nrows, int64_t* ncols, int* br, int* bc) {
  PYN_REQUIRE(m, "null mat");
  if (nrows) *nrows = m->nbrows * m->br;
  if (ncols) *ncols = m->nbcols * m->bc;
  if (br) *br = m->br;
  if (bc) *bc = m->bc;
  return PYN_OK;
}

int pyn_mat_nnz(pyn_mat* m, int64_t* nnz_ref, int64_t* nblocks) {
  PYN_REQUIRE(m, "null mat");
  if (nblocks) *nblocks = m->free_pattern ? m->nblocks : m->nblocks_real;   // without padding
  if (nnz_ref) {
    HostMat h;
    PYN_CHECK(fetch_host(m, &h, false));
    int64_t n = 0;
    walk_reference(m, h, false, [&](int64_t, int64_t, double) { ++n; });
    *nnz_ref = n;
  }
  return PYN_OK;
}

int pyn_mat_export_csr(pyn_mat* m, int64_t* rowptr, int32_t* col, double* val) {
  PYN_REQUIRE(m && rowptr && col && val, "null argument");
  HostMat h;
  PYN_CHECK(fetch_host(m, &h, true));
  const int64_t nrows = m->nbrows * m->br;
  for (int64_t i = 0; i <= nrows; ++i) rowptr[i] = 0;
  int64_t n = 0;
  walk_reference(m, h, true, [&](int64_t r, int64_t c, double v) {
    col[n] = (int32_t)c;
    val[n] = v;
    rowptr[r + 1] = ++n;
  });
  for (int64_t i = 1; i <= nrows; ++i)
    if (rowptr[i] < rowptr[i - 1]) rowptr[i] = rowptr[i - 1];  // empty rows
  return PYN_OK;
}

int pyn_mat_mult(pyn_mat* m, pyn_vec* x, pyn_vec* y) {
  return pyn_mat_mult_axpby(m, 1.0, x, 0.0, nullptr, y);
}

int pyn_mat_mult_axpby(pyn_mat* m, double alpha, pyn_vec* x, double beta, pyn_vec* z, pyn_vec* y) {
  PYN_REQUIRE(m && x && y, "null argument");
  PYN_REQUIRE(x->n == m->nbcols * m->bc, "Mat.mult: x has %lld entries, matrix has %lld columns",
              (long long)x->n, (long long)(m->nbcols * m->bc));
  PYN_REQUIRE(y->n == m->nbrows * m->br, "Mat.mult: y has %lld entries, matrix has %lld rows",
              (long long)y->n, (long long)(m->nbrows * m->br));
  PYN_REQUIRE(x != y, "Mat.mult: x and y must be different vectors");
  PYN_REQUIRE(beta == 0.0 || (z && z->n == y->n), "Mat.mult: bad z");
  return spmv(m, alpha, x->d.p, beta, z ? z->d.p : nullptr, y->d.p);
}

int pyn_mat_add_new(pyn_mat* a, pyn_mat* b, pyn_mat** out) {
  PYN_REQUIRE(a && b && out, "null argument");
  return mat_add_new(a, b, out);
}

__global__ void extract_diag_kernel(const double* dblk, int b, int64_t n, double* d) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n;
       i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t node = i / b;
    const int k = (int)(i % b);
    d[i] = dblk[node * b * b + k * b + k];
  }
}

int pyn_mat_get_diagonal(pyn_mat* m, pyn_vec* d) {
  PYN_REQUIRE(m && d, "null argument");
  PYN_REQUIRE(m->br == m->bc && d->n == m->nbrows * m->br, "getDiagonal: shape");
  pyn_ctx* ctx = m->ctx;
  DevBuf<double> dblk;
  PYN_CHECK(dblk.alloc(ctx, m->nbrows * m->br * m->br));
  PYN_CHECK(diag_blocks(m, dblk.p));
  extract_diag_kernel<<<grid_for(ctx, d->n, 256), 256, 0, ctx->stream>>>(dblk.p, m->br, d->n, d->d.p);
  PYN_LAUNCHED(ctx);
  PYN_CUDA(cudaStreamSynchronize(ctx->stream));
  return PYN_OK;
}

int pyn_mat_scale_rows(pyn_mat* m, pyn_vec* l) {
  PYN_REQUIRE(m && l && l->n == m->nbrows * m->br, "diagonalScale: shape");
  return scale_rows(m, l->d.p);
}

int pyn_mat_bytes(pyn_mat* m, int64_t* stored_bytes, int64_t* spmv_bytes) {
  PYN_REQUIRE(m, "null mat");
  const int64_t bb = m->br * m->bc;
  if (stored_bytes) *stored_bytes = m->nblocks * (8 * bb + 4 + 8) + (m->nbrows + 1) * 8;
  // compulsory SpMV traffic of the stored format: values + block
  // column ids + row pointers + y written + x read once
  // (the padding blocks of the rows are stored and read but not credited)
  const int64_t nreal = m->free_pattern ? m->nblocks : m->nblocks_real;
  if (spmv_bytes) {
    if (m->halo && m->own_blocks_real >= 0)
      // row-partitioned: a product multiplies the owned rows only (the rows of ghost nodes are
      // stored but never read), x is read over owned + ghost entries
      *spmv_bytes = m->own_blocks_real * (8 * bb + 4) + m->halo->n_own * (8 * m->br + 8) + m->nbcols * 8 * m->bc;
    else
      *spmv_bytes = nreal * (8 * bb + 4) + m->nbrows * (8 * m->br + 8) + m->nbcols * 8 * m->bc;
  }
  return PYN_OK;
}

int pyn_mat_destroy(pyn_mat* m) {
  delete m;
  return PYN_OK;
}

// ---- Vec ---------------------------------------------------------------------------
int pyn_vec_create(pyn_ctx* ctx, int64_t n, pyn_vec** out) {
  PYN_REQUIRE(ctx && out && n >= 0, "pyn_vec_create: bad argument");
  pyn_vec* v = new pyn_vec();
  v->ctx = ctx;
  v->n = n;
  int rc = v->d.alloc(ctx, n);
  if (rc == PYN_OK) rc = v->d.zero();
  if (rc != PYN_OK) {
    delete v;
    return rc;
  }
  *out = v;
  return PYN_OK;
}
int pyn_vec_destroy(pyn_vec* v) {
  delete v;
  return PYN_OK;
}
int pyn_vec_set_owned(pyn_vec* v, int64_t owned) {
  PYN_REQUIRE(v && owned <= v->n, "pyn_vec_set_owned: %lld owned entries of %lld", (long long)owned,
              v ? (long long)v->n : -1LL);
  v->owned = owned;
  return PYN_OK;
}
int pyn_mat_set_halo(pyn_mat* m, pyn_halo* h) {
  PYN_REQUIRE(m, "null mat");
  PYN_REQUIRE(!h || (h->n_loc == m->nbrows && h->n_loc == m->nbcols),
              "pyn_mat_set_halo: layout has %lld local nodes, matrix is %lld x %lld", h ? (long long)h->n_loc : 0LL,
              (long long)m->nbrows, (long long)m->nbcols);
  m->halo = h;
  if (h) PYN_CHECK(pyn::spmv_build_row_order(m));
  return PYN_OK;
}
int pyn_mat_set_matfree(pyn_mat* m, int on) {
  PYN_REQUIRE(m, "null mat");
  if (!on) {
    m->mf = nullptr;
    return PYN_OK;
  }
  pyn_mesh* mesh = m->mesh;
  PYN_REQUIRE(mesh && m->ref_dense && m->br == mesh->dim && m->bc == mesh->dim && m->use_rowcls &&
                  m->use_colcls && m->pairmask == 1u && m->diagmask == 0xEu,
              "pyn_mat_set_matfree: not the KLE matrix K of pyn_assemble_kle");
  if (!mesh->mf) PYN_CHECK(pyn::matfree_create(mesh, &mesh->mf));
  m->mf = mesh->mf;
  return PYN_OK;
}
int pyn_vec_size(pyn_vec* v, int64_t* n) {
  PYN_REQUIRE(v && n, "null argument");
  *n = v->n;
  return PYN_OK;
}
int pyn_vec_set(pyn_vec* v, double a) {
  PYN_REQUIRE(v, "null vec");
  return vec_set(v->ctx, v->d.p, v->n, a);
}
int pyn_vec_copy(pyn_vec* src, pyn_vec* dst) {
  PYN_REQUIRE(src && dst && src->n == dst->n, "Vec.copy: size mismatch");
  if (src != dst && src->n > 0)
    PYN_CUDA(cudaMemcpyAsync(dst->d.p, src->d.p, sizeof(double) * src->n, cudaMemcpyDeviceToDevice,
                             src->ctx->stream));
  return PYN_OK;
}
int pyn_vec_scale(pyn_vec* v, double a) {
  PYN_REQUIRE(v, "null vec");
  return vec_scale(v->ctx, v->d.p, v->n, a);
}
int pyn_vec_axpy(pyn_vec* y, double a, pyn_vec* x) {
  PYN_REQUIRE(x && y && x->n == y->n, "Vec.axpy: size mismatch");
  return vec_axpby(y->ctx, y->d.p, a, x->d.p, 1.0, y->n);
}
int pyn_vec_aypx(pyn_vec* y, double a, pyn_vec* x) {
  PYN_REQUIRE(x && y && x->n == y->n, "Vec.aypx: size mismatch");
  return vec_axpby(y->ctx, y->d.p, 1.0, x->d.p, a, y->n);
}
int pyn_vec_axpby(pyn_vec* y, double a, pyn_vec* x, double b) {
  PYN_REQUIRE(x && y && x->n == y->n, "Vec.axpby: size mismatch");
  return vec_axpby(y->ctx, y->d.p, a, x->d.p, b, y->n);
}
int pyn_vec_waxpy(pyn_vec* w, double a, pyn_vec* x, pyn_vec* y) {
  PYN_REQUIRE(w && x && y && x->n == y->n && w->n == x->n, "Vec.waxpy: size mismatch");
  return vec_waxpy(w->ctx, w->d.p, a, x->d.p, y->d.p, w->n);
}
int pyn_vec_pointwise_mult(pyn_vec* w, pyn_vec* x, pyn_vec* y) {
  PYN_REQUIRE(w && x && y && x->n == y->n && w->n == x->n, "Vec.pointwiseMult: size mismatch");
  return vec_pointwise_mult(w->ctx, w->d.p, x->d.p, y->d.p, w->n);
}
int pyn_vec_reciprocal(pyn_vec* v) {
  PYN_REQUIRE(v, "null vec");
  return vec_reciprocal(v->ctx, v->d.p, v->n);
}
int pyn_vec_dot(pyn_vec* x, pyn_vec* y, double* out) {
  PYN_REQUIRE(x && y && out && x->n == y->n, "Vec.dot: size mismatch");
  return vec_dot(x->ctx, x->d.p, y->d.p, x->owned >= 0 ? x->owned : x->n, out, x->owned >= 0);
}
int pyn_vec_norm2(pyn_vec* x, double* out) {
  PYN_REQUIRE(x && out, "null argument");
  double s = 0;
  PYN_CHECK(vec_dot(x->ctx, x->d.p, x->d.p, x->owned >= 0 ? x->owned : x->n, &s, x->owned >= 0));
  *out = sqrt(s);
  return PYN_OK;
}
int pyn_vec_sum(pyn_vec* x, double* out) {
  PYN_REQUIRE(x && out, "null argument");
  return vec_sum(x->ctx, x->d.p, x->owned >= 0 ? x->owned : x->n, out, x->owned >= 0);
}
int pyn_vec_maxpy(pyn_vec* y, int k, const double* a, pyn_vec* const* x) {
  PYN_REQUIRE(y && k >= 0 && (k == 0 || (a && x)), "Vec.maxpy: bad argument");
  std::vector<const double*> ptrs(k);
  for (int i = 0; i < k; ++i) {
    PYN_REQUIRE(x[i] && x[i]->n == y->n, "Vec.maxpy: size mismatch");
    ptrs[i] = x[i]->d.p;
  }
  return vec_maxpy(y->ctx, y->d.p, k, a, ptrs.data(), y->n);
}

int pyn_vec_set_values(pyn_vec* v, int64_t k, const int64_t* idx, const double* vals, int add) {
  PYN_REQUIRE(v && k >= 0 && (k == 0 || (idx && vals)), "Vec.setValues: bad argument");
  if (k == 0) return PYN_OK;
  for (int64_t i = 0; i < k; ++i)
    PYN_REQUIRE(idx[i] >= 0 && idx[i] < v->n, "Vec.setValues: index %lld out of range [0,%lld)",
                (long long)idx[i], (long long)v->n);
  pyn_ctx* ctx = v->ctx;
  DevBuf<int64_t> di;
  DevBuf<double> dv;
  PYN_CHECK(di.alloc(ctx, k));
  PYN_CHECK(dv.alloc(ctx, k));
  PYN_CHECK(di.upload(idx, k));
  PYN_CHECK(dv.upload(vals, k));
  PYN_CHECK(vec_scatter(ctx, v->d.p, k, di.p, dv.p, add));
  PYN_CUDA(cudaStreamSynchronize(ctx->stream));
  return PYN_OK;
}
// ---- persistent scatter plan (k10): indices uploaded once, values per call ------------
struct pyn_scatter {
  pyn_ctx* ctx = nullptr;
  int64_t k = 0, nmax = 0;
  DevBuf<int64_t> idx;
  static constexpr int kSlots = 4;
  DevBuf<double> dvals[kSlots];
  double* hvals[kSlots] = {nullptr, nullptr, nullptr, nullptr};   // pinned staging
  cudaEvent_t done[kSlots] = {nullptr, nullptr, nullptr, nullptr};
  int next = 0;
};

int pyn_scatter_create(pyn_ctx* ctx, int64_t k, const int64_t* idx, pyn_scatter** out) {
  PYN_REQUIRE(ctx && out && k >= 0 && (k == 0 || idx), "pyn_scatter_create: bad argument");
  pyn_scatter* p = new pyn_scatter();
  p->ctx = ctx;
  p->k = k;
  for (int64_t i = 0; i < k; ++i) {
    if (idx[i] < 0) {
      delete p;
      set_error("pyn_scatter_create: negative index");
      return PYN_E_INVALID;
    }
    if (idx[i] + 1 > p->nmax) p->nmax = idx[i] + 1;
  }
  int rc = p->idx.alloc(ctx, k);
  if (rc == PYN_OK) rc = p->idx.upload(idx, k);
  for (int s = 0; s < pyn_scatter::kSlots && rc == PYN_OK; ++s) {
    rc = p->dvals[s].alloc(ctx, k);
    if (rc == PYN_OK && k > 0 && cudaMallocHost((void**)&p->hvals[s], sizeof(double) * k) != cudaSuccess) {
      set_error("pyn_scatter_create: cudaMallocHost failed");
      rc = PYN_E_NOMEM;
    }
    if (rc == PYN_OK) cudaEventCreateWithFlags(&p->done[s], cudaEventDisableTiming);
  }
  if (rc == PYN_OK && cudaStreamSynchronize(ctx->stream) != cudaSuccess) rc = PYN_E_CUDA;
  if (rc != PYN_OK) {
    pyn_scatter_destroy(p);
    return rc;
  }
  *out = p;
  return PYN_OK;
}

int pyn_scatter_apply(pyn_scatter* p, pyn_vec* v, const double* vals, int add) {
  PYN_REQUIRE(p && v && (p->k == 0 || vals), "pyn_scatter_apply: bad argument");
  PYN_REQUIRE(p->nmax <= v->n, "pyn_scatter_apply: plan index %lld beyond Vec size %lld",
              (long long)(p->nmax - 1), (long long)v->n);
  if (p->k == 0) return PYN_OK;
  pyn_ctx* ctx = p->ctx;
  const int s = p->next;
  p->next = (p->next + 1) % pyn_scatter::kSlots;
  PYN_CUDA(cudaEventSynchronize(p->done[s]));          // staging slot free again
  memcpy(p->hvals[s], vals, sizeof(double) * p->k);
  PYN_CUDA(cudaMemcpyAsync(p->dvals[s].p, p->hvals[s], sizeof(double) * p->k, cudaMemcpyHostToDevice,
                           ctx->stream));
  PYN_CHECK(vec_scatter(ctx, v->d.p, p->k, p->idx.p, p->dvals[s].p, add));
  PYN_CUDA(cudaEventRecord(p->done[s], ctx->stream));
  return PYN_OK;
}

int pyn_scatter_destroy(pyn_scatter* p) {
  if (!p) return PYN_OK;
  for (int s = 0; s < pyn_scatter::kSlots; ++s) {
    if (p->done[s]) {
      cudaEventSynchronize(p->done[s]);
      cudaEventDestroy(p->done[s]);
    }
    if (p->hvals[s]) cudaFreeHost(p->hvals[s]);
  }
  delete p;
  return PYN_OK;
}

int pyn_vec_get_values(pyn_vec* v, int64_t k, const int64_t* idx, double* vals) {
  PYN_REQUIRE(v && k >= 0 && (k == 0 || (idx && vals)), "Vec.getValues: bad argument");
  if (k == 0) return PYN_OK;
  for (int64_t i = 0; i < k; ++i)
    PYN_REQUIRE(idx[i] >= 0 && idx[i] < v->n, "Vec.getValues: index out of range");
  pyn_ctx* ctx = v->ctx;
  DevBuf<int64_t> di;
  DevBuf<double> dv;
  PYN_CHECK(di.alloc(ctx, k));
  PYN_CHECK(dv.alloc(ctx, k));
  PYN_CHECK(di.upload(idx, k));
  PYN_CHECK(vec_gather(ctx, v->d.p, k, di.p, dv.p));
  PYN_CHECK(dv.download(vals, k));
  return PYN_OK;
}
int pyn_vec_to_host(pyn_vec* v, double* out) {
  PYN_REQUIRE(v && (out || v->n == 0), "null argument");
  return v->d.download(out, v->n);
}
int pyn_vec_from_host(pyn_vec* v, const double* in) {
  PYN_REQUIRE(v && (in || v->n == 0), "null argument");
  PYN_CHECK(v->d.upload(in, v->n));
  PYN_CUDA(cudaStreamSynchronize(v->ctx->stream));
  return PYN_OK;
}
int pyn_vec_to_host_n(pyn_vec* v, double* out, int64_t n) {
  PYN_REQUIRE(v && n >= 0 && n <= v->n && (out || n == 0), "pyn_vec_to_host_n: bad argument");
  return v->d.download(out, n);
}
int pyn_vec_from_host_n(pyn_vec* v, const double* in, int64_t n) {
  PYN_REQUIRE(v && n >= 0 && n <= v->n && (in || n == 0), "pyn_vec_from_host_n: bad argument");
  PYN_CHECK(v->d.upload(in, n));
  PYN_CUDA(cudaStreamSynchronize(v->ctx->stream));
  return PYN_OK;
}
int pyn_vec_device_ptr(pyn_vec* v, void** ptr) {
  PYN_REQUIRE(v && ptr, "null argument");
  *ptr = v->d.p;
  return PYN_OK;
}

// ---- RHS pieces ----------------------------------------------------------------------
int pyn_vtensv(pyn_vec* vel, pyn_vec* out, int dim) {
  PYN_REQUIRE(vel && out && (dim == 2 || dim == 3), "vtensv: bad argument");
  const int ds = dim == 2 ? 3 : 6;
  PYN_REQUIRE(vel->n % dim == 0 && out->n == vel->n / dim * ds, "vtensv: size mismatch");
  return vtensv(vel->ctx, vel->d.p, out->d.p, vel->n / dim, dim);
}
int pyn_stress_combine(pyn_vec* aux, pyn_vec* vel, double mu, double rho, int dim) {
  PYN_REQUIRE(vel && aux && (dim == 2 || dim == 3), "stress_combine: bad argument");
  const int ds = dim == 2 ? 3 : 6;
  PYN_REQUIRE(vel->n % dim == 0 && aux->n == vel->n / dim * ds, "stress_combine: size mismatch");
  return stress_combine(vel->ctx, aux->d.p, vel->d.p, mu, rho, vel->n / dim, dim);
}

}  // extern "C"
