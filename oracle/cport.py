"""Oracle: Python side of the C + OpenMP port (oracle/ckle.c -> oracle/libckle.so).

TEST INFRASTRUCTURE -- see oracle/__init__.py.  Used by bench.py's CPU legs (`cpu_baseline`,
`--impl reference`) and by tests/test_oracle_cport.py; never by the product.

It times the reference's path on the host cores of the GPU box at BASELINE.json's full size, which
the numpy/scipy oracle cannot assemble within minutes.  The per-cell matrices still come from
oracle/sem.py (spectral.py:92-228); the scatter-add (mat_fs.py:150-201, 238-271), the products
(PETSc MatMult on scalar AIJ) and the Krylov loop run in C on all cores.  `CProblem` is
oracle.timestep.OracleProblem with those three pieces swapped; evalRHS, the BC scatter, the Fischer
guess and the RK5BS recurrence are inherited unchanged.  Free-slip (MatFS) only -- the benchmark
configuration; the no-slip matrices stay with the numpy oracle.
"""
import ctypes
import os
import subprocess

import numpy as np
import scipy.sparse as sp

from . import sem, timestep

HERE = os.path.dirname(os.path.abspath(__file__))
LIB = os.path.join(HERE, "libckle.so")
_lib = None

c_i64, c_int, c_dbl = ctypes.c_int64, ctypes.c_int, ctypes.c_double
vp = ctypes.c_void_p


def build():
    """make -C oracle (gcc -O3 -fopenmp); __graft_entry__.build() calls this"""
    r = subprocess.run(["make", "-C", HERE], stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    if r.returncode != 0:
        raise RuntimeError("oracle/libckle.so did not build:\n" + r.stdout[-3000:])


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB):
            build()
        L = ctypes.CDLL(LIB)
        L.ckle_threads.restype = c_int
        L.ckle_chunk_nodes.restype = c_int
        L.ckle_node_adjacency.argtypes = [c_i64, c_int, c_i64, vp, vp, vp]
        L.ckle_node_adjacency.restype = None
        L.ckle_assemble.argtypes = [c_int, c_i64, c_int, c_i64, vp, vp, vp, vp, vp, c_int, c_int, vp, vp, vp,
                                    c_dbl, c_dbl, vp, vp, vp, vp]
        L.ckle_assemble.restype = c_int
        L.ckle_spmv.argtypes = [c_i64, c_int, vp, vp, vp, vp, vp]
        L.ckle_spmv.restype = None
        L.ckle_pcg.argtypes = [c_i64, c_int, vp, vp, vp, c_int, vp, vp, vp, c_dbl, c_int, vp, vp]
        L.ckle_pcg.restype = c_int
        L.ckle_diag_blocks.argtypes = [c_i64, c_int, vp, vp, vp, vp]
        L.ckle_diag_blocks.restype = None
        _lib = L
    return _lib


def threads():
    return lib().ckle_threads()


def _p(a):
    return None if a is None else a.ctypes.data_as(vp)


class CMat:
    """scalar CSR (int64 row pointers, int32 columns) whose product runs in C over all cores"""

    def __init__(self, shape, indptr, indices, data, br):
        self.shape, self.indptr, self.indices, self.data = shape, indptr, indices, data
        self.chunk_rows = lib().ckle_chunk_nodes() * br
        self.nnz = int(indptr[-1])

    def __matmul__(self, x):
        x = np.ascontiguousarray(x, dtype=np.float64)
        assert x.size == self.shape[1]
        y = np.empty(self.shape[0])
        lib().ckle_spmv(self.shape[0], self.chunk_rows, _p(self.indptr), _p(self.indices), _p(self.data),
                        _p(x), _p(y))
        return y

    def tocsr(self):
        return sp.csr_matrix((self.data, self.indices, self.indptr), shape=self.shape)

    def diag_blocks(self, d):
        blk = np.empty((self.shape[0] // d, d, d))
        lib().ckle_diag_blocks(self.shape[0], d, _p(self.indptr), _p(self.indices), _p(self.data), _p(blk))
        return blk


def element_classes(corner_xyz, dim, reuse=True):
    """class id per cell: cells with the same corner offsets share their matrices
    (oracle.assembly.ElementCache's key); reuse=False: every cell its own class, the reference's cost"""
    X = np.asarray(corner_xyz, dtype=np.float64).reshape(-1, 2 ** dim, dim)
    if not reuse:
        return np.arange(len(X), dtype=np.int32), np.arange(len(X))
    key = np.round((X - X[:, :1]) * 2.0 ** 40).astype(np.int64).reshape(len(X), -1)
    _, rep, inv = np.unique(key, axis=0, return_index=True, return_inverse=True)
    return inv.astype(np.int32).ravel(), rep


def _assemble_one(L, nel, nn, nnodes, en, ecls, nptr, nadj, E, br, bc, rowmask, colmask=None, diag=None,
                  diag_add=0.0, sign=1.0, rowscale=None):
    nrows = nnodes * br
    E = np.ascontiguousarray(E, dtype=np.float64)
    rowmask = np.ascontiguousarray(rowmask, dtype=np.uint8)
    colmask = None if colmask is None else np.ascontiguousarray(colmask, dtype=np.uint8)
    diag = None if diag is None else np.ascontiguousarray(diag, dtype=np.uint8)
    rowscale = None if rowscale is None else np.ascontiguousarray(rowscale, dtype=np.float64)
    indptr = np.zeros(nrows + 1, dtype=np.int64)
    args = (nel, nn, nnodes, _p(en), _p(ecls), _p(nptr), _p(nadj), _p(E), br, bc, _p(rowmask), _p(colmask),
            _p(diag), float(diag_add), float(sign), _p(rowscale))
    if L.ckle_assemble(0, *args, _p(indptr), None, None) != 0:
        raise RuntimeError("ckle_assemble: a row touches more than 4096 nodes")
    np.cumsum(indptr, out=indptr)
    indices = np.empty(int(indptr[-1]), dtype=np.int32)
    data = np.empty(int(indptr[-1]), dtype=np.float64)
    L.ckle_assemble(1, *args, _p(indptr), _p(indices), _p(data))
    return CMat((nrows, nnodes * bc), indptr, indices, data, br)


def assemble_fs(tab, corner_xyz, elem_nodes, nnodes, cls, reuse=True, ops=True):
    """K, Krhs, Rw (MatFS.buildFS, mat_fs.py:150-192) and Curl, SrT, DivSrT, W (mat_fs.py:194-271)
    as CMat: same patterns (explicit zeros kept) and values as oracle.assembly on the same tables"""
    L = lib()
    dim, dim_w, dim_s = tab.dim, tab.dim_w, tab.dim_s
    en = np.ascontiguousarray(elem_nodes, dtype=np.int32)
    nel, nn = en.shape
    X = np.asarray(corner_xyz, dtype=np.float64).reshape(nel, 2 ** dim, dim)
    ecls, rep = element_classes(X, dim, reuse)
    nptr = np.empty(nnodes + 1, dtype=np.int64)
    nadj = np.empty(nel * nn, dtype=np.int64)
    L.ckle_node_adjacency(nel, nn, nnodes, _p(en), _p(nptr), _p(nadj))
    cls = np.asarray(cls, dtype=np.uint8)
    free, bc = (cls == 0), (cls == 1)
    kle = [sem.elem_kle(tab, X[e]) for e in rep]                 # spectral.py:92-160, once per class
    EK = np.stack([k[0] for k in kle])
    ERw = np.stack([k[1] for k in kle])
    common = (L, nel, nn, nnodes, en, ecls, nptr, nadj)
    out = {
        "K": _assemble_one(*common, EK, dim, dim, free, colmask=free, diag=bc, diag_add=1.0),
        "Krhs": _assemble_one(*common, EK, dim, dim, free, colmask=bc, diag=bc, diag_add=1.0, sign=-1.0),
        "Rw": _assemble_one(*common, ERw, dim, dim_w, free),
    }
    if ops:
        eo = [sem.elem_ops(tab, X[e]) for e in rep]              # spectral.py:162-228
        W = np.zeros(nnodes)
        np.add.at(W, en.ravel(), np.stack([o[3] for o in eo])[ecls].ravel())
        winv = 1.0 / W
        ones = lambda br: np.ones(nnodes * br, np.uint8)
        out["SrT"] = _assemble_one(*common, np.stack([o[0] for o in eo]), dim_s, dim, ones(dim_s),
                                   rowscale=np.repeat(winv, dim_s))
        out["DivSrT"] = _assemble_one(*common, np.stack([o[1] for o in eo]), dim, dim_s, ones(dim),
                                      rowscale=np.repeat(winv, dim))
        out["Curl"] = _assemble_one(*common, np.stack([o[2] for o in eo]), dim_w, dim, ones(dim_w),
                                    rowscale=np.repeat(winv, dim_w))
        out["W"] = W
    return out


class CProblem(timestep.OracleProblem):
    """oracle.timestep.OracleProblem (free-slip Taylor-Green) with the assembly, the products and the
    CG loop in C over all host cores; solver = cg, pc = jacobi | bjacobi"""

    def __init__(self, dim, ngl, corner_xyz, elem_nodes, coords, bnd_nodes, rho, mu, rtol=1e-10, pc="bjacobi",
                 reuse=True, fischer=0):
        self.dim, self.dim_w = dim, (1 if dim == 2 else 3)
        self.rho, self.mu, self.nu = rho, mu, mu / rho
        self.coords = np.asarray(coords, dtype=np.float64)
        nnodes = len(self.coords)
        self.bnd = np.asarray(bnd_nodes, dtype=np.int64)
        cls = np.zeros(nnodes * dim, np.uint8)
        self.bdofs = (self.bnd[:, None] * dim + np.arange(dim)[None, :]).ravel()
        cls[self.bdofs] = 1
        self.bwdofs = self.bnd if dim == 2 else self.bdofs
        self.ngl = ngl
        tab = sem.SpectralTables(ngl, dim)
        m = assemble_fs(tab, corner_xyz, elem_nodes, nnodes, cls, reuse=reuse)
        self.mats = {k: m[k] for k in ("K", "Krhs", "Rw")}
        self.ops = {k: m[k] for k in ("Curl", "SrT", "DivSrT", "W")}
        self.solver, self.rtol, self.pc = "cg", rtol, pc
        self.K = self.mats["K"]
        n = self.K.shape[0]
        if pc == "jacobi":
            self._pcd, self._minv = 1, np.ascontiguousarray(1.0 / np.einsum("nii->ni", self.K.diag_blocks(1)).ravel())
        elif pc == "bjacobi":
            self._pcd, self._minv = dim, np.ascontiguousarray(np.linalg.inv(self.K.diag_blocks(dim)))
        else:
            raise ValueError("CProblem: pc must be jacobi or bjacobi")
        self._work = np.empty(4 * n)
        a0 = timestep.tg_alpha(self.nu, 0.0, dim)
        self.vel = timestep.tg_velocity(self.coords, a0)
        self.vort = timestep.tg_vorticity(self.coords, a0)
        self.iterations = self.solves = 0
        self.fischer, self._fx, self._fax = int(fischer), [], []

    def _cg(self, b, x0):
        K = self.K
        x = np.array(x0, dtype=np.float64, copy=True)
        b = np.ascontiguousarray(b)
        its = lib().ckle_pcg(K.shape[0], K.chunk_rows, _p(K.indptr), _p(K.indices), _p(K.data), self._pcd,
                             _p(self._minv), _p(b), _p(x), float(self.rtol), 20000, _p(self._work), None)
        self.iterations += abs(its)
        return x

    # kle_solver.py:33-37 with PETSc's -ksp_fischer_guess 1,<n> as in OracleProblem.solve_kle
    def solve_kle(self, vort):
        K = self.K
        b = self.mats["Rw"] @ vort + self.mats["Krhs"] @ self.vel
        x0 = self.vel
        if self.fischer:
            x0 = sum(((fx @ b) * fx for fx in self._fx), np.zeros_like(b))
        r0 = b - K @ x0
        x = self._cg(b, x0)
        if self.fischer:
            if len(self._fx) == self.fischer:
                self._fx, self._fax = [], []
                d, ad = x.copy(), K @ x
            else:
                d, ad = x - x0, r0 - (b - K @ x)
            a0 = d @ ad
            for fx, fax in zip(self._fx, self._fax):
                c = fx @ ad
                d = d - c * fx
                ad = ad - c * fax
            a = d @ ad
            tol = max(100.0 * self.rtol, 1e-12)
            if a > 0 and a0 > 0 and a > tol * tol * abs(x @ (K @ x)):
                self._fx.append(d / np.sqrt(a))
                self._fax.append(ad / np.sqrt(a))
        self.vel = x
        self.solves += 1
        return self.vel
