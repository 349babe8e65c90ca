/* Oracle, C + OpenMP: the CPU arm of bench.py at the full benchmark size.
 *
 * TEST INFRASTRUCTURE -- see oracle/__init__.py.  Only bench.py's CPU legs and tests/ load this
 * library (oracle/cport.py); nothing under pynama_b200/ does.
 *
 * The numpy/scipy oracle (oracle/assembly.py, oracle/timestep.py) cannot assemble BASELINE.json
 * configs[1] (16^3 elements, ngl = 4: 1.26e8 scalar non-zeros in K, 3.0e8 in SrT) within minutes --
 * scipy's COO -> CSR conversion is single-threaded and its SpMV uses one core.  This file restates
 * the same three pieces in C so that the host cores of the GPU box can be timed on the SAME mesh:
 *
 *   ckle_assemble  the scatter-add of per-cell dense blocks into scalar AIJ/CSR with row / column
 *                  dof masks, in ascending cell order, explicit zeros kept:
 *                    MatFS.buildFS          src/matrices/mat_fs.py:150-192  (K, Krhs, Rw)
 *                    Operators.setValues    src/matrices/mat_fs.py:238-251  (Curl, SrT, DivSrT)
 *                    Operators.assembleAll  src/matrices/mat_fs.py:253-271  (row scaling)
 *                  (element blocks come from oracle/sem.py, i.e. spectral.py:92-228)
 *   ckle_spmv      y = A x on scalar CSR (PETSc MatMult on AIJ; kle_solver.py:35-41,
 *                  base_problem.py:123-136), rows split over the threads
 *   ckle_pcg       preconditioned CG with a node-block Jacobi preconditioner, stopping on
 *                  ||r|| <= rtol ||b||: what oracle/timestep.py asks scipy.sparse.linalg.cg for
 *
 * tests/test_oracle_cport.py checks every function against the numpy oracle on small meshes
 * (patterns bit-exact, values to rounding).
 */
#include <math.h>
#include <omp.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#define CHUNK_NODES 128
#define MAX_ROW_NODES 4096

int ckle_threads(void) { return omp_get_max_threads(); }
int ckle_chunk_nodes(void) { return CHUNK_NODES; }

static int cmp_i32(const void* a, const void* b) {
  const int32_t x = *(const int32_t*)a, y = *(const int32_t*)b;
  return x < y ? -1 : x > y;
}

/* node -> (element, local index) adjacency, ascending element order.
 * ptr: nnodes+1, adj: nel*nn entries packed as e*nn + a */
void ckle_node_adjacency(int64_t nel, int nn, int64_t nnodes, const int32_t* en, int64_t* ptr, int64_t* adj) {
  memset(ptr, 0, sizeof(int64_t) * (size_t)(nnodes + 1));
  for (int64_t k = 0; k < nel * nn; ++k) ptr[en[k] + 1]++;
  for (int64_t i = 0; i < nnodes; ++i) ptr[i + 1] += ptr[i];
  int64_t* fill = (int64_t*)malloc(sizeof(int64_t) * (size_t)nnodes);
  memcpy(fill, ptr, sizeof(int64_t) * (size_t)nnodes);
  for (int64_t k = 0; k < nel * nn; ++k) adj[fill[en[k]]++] = k;
  free(fill);
}

/* One matrix.  Row dofs r*br+i, column dofs c*bc+j over the node graph (r, c share a cell).
 *   rowmask[r*br+i] != 0 : the row takes the cells' contributions, columns with colmask != 0
 *                          (colmask NULL: every column)
 *   diag[r*br+i]    != 0 : the row holds the diagonal entry (inserted as explicit 0 by the cell
 *                          loop, mat_fs.py:182-183) plus diag_add (setIndices2One, :115-120)
 *   E: [ncls][nn*br][nn*bc] dense cell blocks; ecls[e]: class of cell e; sign: +1 / -1 (Krhs)
 *   rowscale: NULL or one factor per row dof (diagonalScale, mat_fs.py:257-267)
 * pass 0 writes the row lengths into indptr[1..]; the caller prefix-sums; pass 1 fills indices/data.
 * Returns 0, or -1 when a row touches more than MAX_ROW_NODES column nodes. */
int ckle_assemble(int pass, int64_t nel, int nn, int64_t nnodes, const int32_t* en, const int32_t* ecls,
                  const int64_t* nptr, const int64_t* nadj, const double* E, int br, int bc,
                  const uint8_t* rowmask, const uint8_t* colmask, const uint8_t* diag, double diag_add,
                  double sign, const double* rowscale, int64_t* indptr, int32_t* indices, double* data) {
  int bad = 0;
  const int64_t ldE = (int64_t)nn * bc, szE = (int64_t)nn * br * ldE;
#pragma omp parallel
  {
    int32_t* cols = (int32_t*)malloc(sizeof(int32_t) * MAX_ROW_NODES * 2);
    double* acc = pass ? (double*)malloc(sizeof(double) * MAX_ROW_NODES * (size_t)(br * bc)) : NULL;
    /* chunks of CHUNK_NODES nodes dealt round-robin: ckle_spmv deals the same rows to the same
     * threads, so the pages a thread first touches here are the ones it streams later (NUMA) */
#pragma omp for schedule(static, CHUNK_NODES)
    for (int64_t r = 0; r < nnodes; ++r) {
      /* column nodes of the row: union over the cells of r, sorted */
      if ((nptr[r + 1] - nptr[r]) * nn > MAX_ROW_NODES * 2) { bad = 1; continue; }
      int m = 0;
      for (int64_t q = nptr[r]; q < nptr[r + 1]; ++q) {
        const int32_t* nodes = en + (nadj[q] / nn) * nn;
        for (int b = 0; b < nn; ++b) cols[m++] = nodes[b];
      }
      qsort(cols, (size_t)m, sizeof(int32_t), cmp_i32);
      int nc = 0;
      for (int k = 0; k < m; ++k)
        if (k == 0 || cols[k] != cols[k - 1]) cols[nc++] = cols[k];
      if (nc > MAX_ROW_NODES) { bad = 1; continue; }
      if (pass) {
        memset(acc, 0, sizeof(double) * (size_t)nc * br * bc);
        for (int64_t q = nptr[r]; q < nptr[r + 1]; ++q) {      /* ascending cell order */
          const int64_t e = nadj[q] / nn;
          const int a = (int)(nadj[q] % nn);
          const int32_t* nodes = en + e * nn;
          const double* Ee = E + (int64_t)ecls[e] * szE + (int64_t)a * br * ldE;
          for (int b = 0; b < nn; ++b) {
            const int32_t key = nodes[b];
            int lo = 0, hi = nc - 1;
            while (lo < hi) {
              const int mid = (lo + hi) >> 1;
              if (cols[mid] < key) lo = mid + 1; else hi = mid;
            }
            double* t = acc + (size_t)lo * br * bc;
            for (int i = 0; i < br; ++i)
              for (int j = 0; j < bc; ++j) t[i * bc + j] += Ee[(int64_t)i * ldE + b * bc + j];
          }
        }
      }
      for (int i = 0; i < br; ++i) {
        const int64_t row = r * br + i;
        int64_t len = 0, pos = pass ? indptr[row] : 0;
        const double sc = rowscale ? rowscale[row] : 1.0;
        if (rowmask[row]) {
          for (int k = 0; k < nc; ++k)
            for (int j = 0; j < bc; ++j) {
              const int64_t col = (int64_t)cols[k] * bc + j;
              if (colmask && !colmask[col]) continue;
              if (pass) {
                indices[pos] = (int32_t)col;
                data[pos] = sign * acc[((size_t)k * br + i) * bc + j] * sc;
                ++pos;
              }
              ++len;
            }
        }
        if (diag && diag[row]) {        /* never coincides with a masked-in entry (bc rows are masked out) */
          if (pass) {
            indices[pos] = (int32_t)row;
            data[pos] = diag_add;
            ++pos;
          }
          ++len;
        }
        if (!pass) indptr[row + 1] = len;
      }
    }
    free(cols);
    free(acc);
  }
  return bad ? -1 : 0;
}

/* y = A x  (scalar CSR).  Chunks of chunk_rows rows (= CHUNK_NODES nodes of the assembly) dealt
 * round-robin over the threads: balanced although constrained rows hold one entry and interior rows
 * hundreds, and every thread streams the pages it first touched during the assembly. */
void ckle_spmv(int64_t nrows, int chunk_rows, const int64_t* indptr, const int32_t* indices,
               const double* data, const double* x, double* y) {
  const int64_t nchunks = (nrows + chunk_rows - 1) / chunk_rows;
#pragma omp parallel for schedule(static, 1)
  for (int64_t c = 0; c < nchunks; ++c) {
    const int64_t r1 = (c + 1) * chunk_rows < nrows ? (c + 1) * chunk_rows : nrows;
    for (int64_t r = c * chunk_rows; r < r1; ++r) {
      double s = 0.0;
      for (int64_t k = indptr[r]; k < indptr[r + 1]; ++k) s += data[k] * x[indices[k]];
      y[r] = s;
    }
  }
}

static double dot_(int64_t n, const double* a, const double* b) {
  double s = 0.0;
#pragma omp parallel for reduction(+ : s) schedule(static)
  for (int64_t i = 0; i < n; ++i) s += a[i] * b[i];
  return s;
}

/* z = M^-1 r with M^-1 given as d x d blocks per node (d = 1: point Jacobi) */
static void pc_(int64_t nb, int d, const double* minv, const double* r, double* z) {
#pragma omp parallel for schedule(static)
  for (int64_t n = 0; n < nb; ++n)
    for (int i = 0; i < d; ++i) {
      double s = 0.0;
      for (int j = 0; j < d; ++j) s += minv[(n * d + i) * d + j] * r[n * d + j];
      z[n * d + i] = s;
    }
}

/* Preconditioned conjugate gradients on A x = b from the guess in x (Hestenes-Stiefel, the
 * recurrence of scipy.sparse.linalg.cg): stops when ||b - A x||_2 (recurred) <= rtol ||b||_2.
 * work: 4 n doubles.  Returns the iterations taken, negative if maxit was reached. */
int ckle_pcg(int64_t n, int chunk_rows, const int64_t* indptr, const int32_t* indices, const double* data,
             int d, const double* minv, const double* b, double* x, double rtol, int maxit, double* work,
             double* rnorm_out) {
  double *r = work, *z = work + n, *p = work + 2 * n, *q = work + 3 * n;
  const double bnorm = sqrt(dot_(n, b, b));
  const double tol = rtol * bnorm;
  ckle_spmv(n, chunk_rows, indptr, indices, data, x, q);
#pragma omp parallel for schedule(static)
  for (int64_t i = 0; i < n; ++i) r[i] = b[i] - q[i];
  double rn = sqrt(dot_(n, r, r));
  int it = 0;
  double rz_old = 0.0;
  while (rn > tol && it < maxit) {
    pc_(n / d, d, minv, r, z);
    const double rz = dot_(n, r, z);
    if (it == 0) {
      memcpy(p, z, sizeof(double) * (size_t)n);
    } else {
      const double beta = rz / rz_old;
#pragma omp parallel for schedule(static)
      for (int64_t i = 0; i < n; ++i) p[i] = z[i] + beta * p[i];
    }
    ckle_spmv(n, chunk_rows, indptr, indices, data, p, q);
    const double alpha = rz / dot_(n, p, q);
    double rr = 0.0;
#pragma omp parallel for reduction(+ : rr) schedule(static)
    for (int64_t i = 0; i < n; ++i) {
      x[i] += alpha * p[i];
      r[i] -= alpha * q[i];
      rr += r[i] * r[i];
    }
    rn = sqrt(rr);
    rz_old = rz;
    ++it;
  }
  if (rnorm_out) *rnorm_out = rn;
  return rn <= tol ? it : -it;
}

/* d x d diagonal blocks of a scalar CSR matrix (node-block Jacobi: PCPBJACOBI on the AIJ matrix);
 * blk: [n/d][d][d], entries not stored stay 0 */
void ckle_diag_blocks(int64_t n, int d, const int64_t* indptr, const int32_t* indices, const double* data,
                      double* blk) {
#pragma omp parallel for schedule(static, 1024)
  for (int64_t r = 0; r < n; ++r) {
    const int64_t c0 = (r / d) * d;
    for (int j = 0; j < d; ++j) blk[r * d + j] = 0.0;
    for (int64_t k = indptr[r]; k < indptr[r + 1]; ++k) {
      const int64_t c = indices[k];
      if (c >= c0 && c < c0 + d) blk[r * d + (c - c0)] = data[k];
    }
  }
}
