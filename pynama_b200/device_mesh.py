"""Mesh tables on the device + the batched assembly entry points.

One ``DeviceMesh`` replaces the reference's per-cell calls
``Domain.computeLocalKLEMats(cell)`` / ``computeLocalOperators(cell)``
(src/domain/domain.py:262-276) and the Python scatter loops around them
(mat_fs.py:150-201, mat_ns.py:47-145) by one call per phase.
"""
import ctypes

import numpy as np

from . import _lib
from ._lib import call, ptr, vp, c_int, c_i64, c_dbl
from .petsc import Context, Mat, Vec

KLE_NAMES = ["K", "Krhs", "Rw", "Rd", "Kfs", "Krhsfs", "Rwfs", "Rdfs"]
PHASES = {"symbolic": 0, "elem_kle": 1, "kle_numeric": 2, "elem_ops": 3, "ops_numeric": 4}


class DeviceMesh:
    def __init__(self, dim, ngl, corner_xyz, elem_nodes, nnodes, layout=None):
        """layout: petsc.Layout of a row-partitioned run -- the tables are then this rank's local
        ones (owned + halo elements, local node ids) and every assembled matrix computes only
        the owned rows."""
        self.layout = layout
        self.dim, self.ngl = int(dim), int(ngl)
        self.nn = self.ngl ** self.dim
        self.dim_w = 1 if dim == 2 else 3
        self.dim_s = 3 if dim == 2 else 6
        cx = _lib.as_f64(corner_xyz).reshape(-1, 2 ** self.dim, self.dim)
        en = np.ascontiguousarray(elem_nodes, dtype=np.int32).reshape(-1, self.nn)
        if len(cx) != len(en):
            raise ValueError("corner_xyz and elem_nodes disagree on the number of elements")
        self.nel, self.nnodes = len(en), int(nnodes)
        h = vp()
        call("pyn_mesh_create", Context.get().handle, self.dim, self.ngl, self.nel, self.nnodes,
             ptr(cx), ptr(en), ctypes.byref(h))
        self._h = h

    def destroy(self):
        if getattr(self, "_h", None) is not None and _lib._lib is not None:
            _lib._lib.pyn_mesh_destroy(self._h)
        self._h = None

    def __del__(self):
        try:
            self.destroy()
        except Exception:
            pass

    def graph_info(self):
        a, b, c = c_i64(), c_i64(), c_int()
        call("pyn_mesh_graph_info", self._h, ctypes.byref(a), ctypes.byref(b), ctypes.byref(c))
        return {"node_pairs": a.value, "coo_length": b.value, "max_row": c.value}

    def phase_ms(self, name):
        ms = c_dbl()
        call("pyn_mesh_phase_ms", self._h, PHASES[name], ctypes.byref(ms))
        return ms.value

    def set_congruence(self, tol):
        """Merge elements whose corner offsets agree to tol * (largest offset in the mesh) into one
        congruence class evaluated once (default at creation: 1e-13); tol <= 0: every element is
        evaluated on its own, as Domain.computeLocalKLEMats does per cell (domain.py:262-276).
        Returns the number of classes."""
        n = c_i64()
        call("pyn_mesh_set_congruence", self._h, float(tol), ctypes.byref(n))
        return n.value

    def congruence(self):
        """-> (tolerance in force, number of congruence classes)"""
        t, n = c_dbl(), c_i64()
        call("pyn_mesh_get_congruence", self._h, ctypes.byref(t), ctypes.byref(n))
        return t.value, n.value

    def set_global_tables(self, elem_nodes_global, n_global_nodes, elem_gid):
        """row-partitioned runs: the whole mesh's element -> node table (global ids) and the global
        element of every local element (second level of -pc_type asm: replicated coarse space)"""
        en = np.ascontiguousarray(elem_nodes_global, dtype=np.int32).reshape(-1, self.nn)
        gid = np.ascontiguousarray(elem_gid, dtype=np.int64).ravel()
        if gid.size != self.nel:
            raise ValueError("elem_gid must have one entry per local element")
        call("pyn_mesh_set_global_tables", self._h, int(n_global_nodes), len(en), ptr(en), ptr(gid))

    def set_dof_class(self, cls):
        cls = np.ascontiguousarray(cls, dtype=np.uint8).ravel()
        if cls.size != self.nnodes * self.dim:
            raise ValueError("dof class array must have nnodes*dim entries")
        call("pyn_mesh_set_dof_class", self._h, ptr(cls))

    # ---- raw element blocks (parity export) ------------------------------------------------
    def elem_kle(self, e0=0, ne=None):
        ne = self.nel - e0 if ne is None else ne
        n, d, dw = self.nn, self.dim, self.dim_w
        K = np.zeros((ne, n * d, n * d))
        Rw = np.zeros((ne, n * d, n * dw))
        Rd = np.zeros((ne, n * d, n))
        call("pyn_elem_kle", self._h, e0, ne, ptr(K), ptr(Rw), ptr(Rd))
        return K, Rw, Rd

    def elem_ops(self, e0=0, ne=None):
        ne = self.nel - e0 if ne is None else ne
        n, d, dw, ds = self.nn, self.dim, self.dim_w, self.dim_s
        S = np.zeros((ne, n * ds, n * d))
        D = np.zeros((ne, n * d, n * ds))
        C = np.zeros((ne, n * dw, n * d))
        W = np.zeros((ne, n))
        call("pyn_elem_ops", self._h, e0, ne, ptr(S), ptr(D), ptr(C), ptr(W))
        return S, D, C, W

    # ---- assembly ----------------------------------------------------------------------------
    def assemble_kle(self, mode):
        """mode 'FS' | 'NS' -> dict name -> Mat (mat_fs.py:131-192 / mat_ns.py:47-166)"""
        out = (vp * 8)()
        call("pyn_assemble_kle", self._h, {"FS": 0, "NS": 1}[mode], out)
        mats = {}
        for i, name in enumerate(KLE_NAMES):
            if out[i]:
                mats[name] = Mat(vp(out[i]), name=name)
                mats[name]._keep = self          # matrices point into the mesh's graphs
                if self.layout is not None:
                    mats[name]._setLayout(self.layout)
        return mats

    def assemble_ops(self, compact=True):
        """-> (Curl, SrT, DivSrT, lumped node weights)  (mat_fs.py:194-271)"""
        c, s, d, w = vp(), vp(), vp(), vp()
        call("pyn_assemble_ops", self._h, 1 if compact else 0, ctypes.byref(c), ctypes.byref(s),
             ctypes.byref(d), ctypes.byref(w))
        mats = [Mat(c, "Curl"), Mat(s, "SrT"), Mat(d, "DivSrT")]
        for m in mats:
            m._keep = self
            if self.layout is not None:
                m._setLayout(self.layout)
        wv = Vec(handle=w, name="lumped-mass")
        if self.layout is not None:
            wv._setLayout(self.layout, 1)
        return mats[0], mats[1], mats[2], wv
