"""Device-resident stand-ins for the petsc4py objects the hot path uses.

The reference calls ``petsc4py.PETSc`` directly; this module
gives the same method names and argument meaning on top of the C ABI, so the
builders / solver / time loop read like the reference's:

* ``Vec``  -- PETSc.Vec surface used in cases/base_problem.py:111-154,
  solver/kle_solver.py:33-47, boundaries/boundary_conditions.py:252-277
* ``Mat``  -- PETSc.Mat (AIJ) surface used in matrices/mat_fs.py,
  solver/kle_solver.py:12-41
* ``KSP`` / ``PC`` -- solver/kle_solver.py:49-64
* ``Options`` -- the PETSc options database (run_case.py:9-13)

Everything computes on the GPU; numpy only carries host copies in and out.
"""
import ctypes
import logging
import os
import sys

import numpy as np

from . import _lib
from ._lib import PynamaError, call, ptr, vp, c_int, c_i64, c_dbl


# ---- communicator / options -----------------------------------------------------------
class _Comm:
    """COMM_WORLD stand-in: rank/size from the torchrun environment."""

    @property
    def rank(self):
        return int(os.environ.get("RANK", "0"))

    @property
    def size(self):
        return int(os.environ.get("WORLD_SIZE", "1"))

    def getRank(self):
        return self.rank

    def getSize(self):
        return self.size

    def tompi4py(self):
        return self

    def allgather(self, obj):
        if self.size == 1:
            return [obj]
        import torch.distributed as dist
        out = [None] * self.size
        dist.all_gather_object(out, obj)
        return out

    def barrier(self):
        if self.size > 1:
            import torch.distributed as dist
            dist.barrier()


COMM_WORLD = _Comm()
COMM_SELF = COMM_WORLD


class Options:
    """PETSc options database: ``-name value`` pairs from argv (run_case.py:9-13)
    plus values set programmatically.  One process-wide table."""
    _db = None

    def __init__(self, prefix=None):
        if Options._db is None:
            Options._db = self._parse(sys.argv[1:])

    @staticmethod
    def _parse(argv):
        db, i = {}, 0
        while i < len(argv):
            a = argv[i]
            if a.startswith("-") and len(a) > 1 and not a[1].isdigit():
                key = a.lstrip("-")
                if i + 1 < len(argv) and not (argv[i + 1].startswith("-") and
                                              not argv[i + 1][1:2].isdigit() and argv[i + 1][1:2] != "."):
                    db[key] = argv[i + 1]
                    i += 2
                    continue
                db[key] = "true"
            i += 1
        return db

    def hasName(self, name):
        return name in Options._db

    def setValue(self, name, value):
        Options._db[name.lstrip("-")] = str(value)

    def delValue(self, name):
        Options._db.pop(name.lstrip("-"), None)

    def getString(self, name, default=None):
        return Options._db.get(name, default)

    def getInt(self, name, default=None):
        return int(Options._db[name]) if name in Options._db else default

    def getReal(self, name, default=None):
        return float(Options._db[name]) if name in Options._db else default

    def getBool(self, name, default=None):
        if name not in Options._db:
            return default
        return Options._db[name].lower() in ("1", "true", "yes", "on")

    def getAll(self):
        return dict(Options._db)


# ---- context ------------------------------------------------------------------------------
class Context:
    """One library context (CUDA stream + scratch) per process, on cuda:LOCAL_RANK."""
    _inst = None

    def __init__(self, device=None):
        if device is None:
            device = int(os.environ.get("LOCAL_RANK", "0"))
        h = vp()
        call("pyn_init", device, ctypes.byref(h))
        self.handle = h
        self.device = device
        self.rank, self.size = COMM_WORLD.rank, COMM_WORLD.size
        if self.size > 1:
            self._init_comm()

    def _init_comm(self):
        """NCCL communicator over the GPUs of the box: rank 0 makes the unique id, the host-side
        rendezvous (torch.distributed, gloo, 127.0.0.1) ships it."""
        import torch.distributed as dist
        if not dist.is_initialized():
            os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
            os.environ.setdefault("MASTER_PORT", "29511")
            dist.init_process_group("gloo", rank=self.rank, world_size=self.size)
        lib = _nccl_library()
        libb = lib.encode() if lib else None
        uid = ctypes.create_string_buffer(128)
        if self.rank == 0:
            call("pyn_comm_unique_id", libb, uid)
        box = [bytes(uid.raw)]
        dist.broadcast_object_list(box, src=0)
        uid = ctypes.create_string_buffer(box[0], 128)
        call("pyn_comm_init", self.handle, libb, self.rank, self.size, uid)
        # peer-memory windows over NVLink (CUDA IPC): fused in-kernel allreduce / halo stores
        self.p2p = os.environ.get("PYN_P2P", "1") != "0" and self.size <= 8
        if self.p2p:
            h = ctypes.create_string_buffer(64)
            call("pyn_comm_window_export", self.handle, h)
            allh = COMM_WORLD.allgather(bytes(h.raw))
            call("pyn_comm_window_attach", self.handle, ctypes.create_string_buffer(b"".join(allh), 64 * self.size))

    def p2p_status(self):
        en, to = c_int(), c_int()
        call("pyn_comm_p2p_status", self.handle, ctypes.byref(en), ctypes.byref(to))
        return bool(en.value), bool(to.value)

    def allreduce(self, values, op="sum"):
        """small host-value collective over the ranks (sum | max)"""
        a = _lib.as_f64(np.atleast_1d(values)).copy()
        if self.size > 1:
            call("pyn_comm_allreduce", self.handle, ptr(a), a.size, {"sum": 0, "max": 1}[op])
        return a

    def collectives(self):
        n = c_i64()
        call("pyn_comm_info", self.handle, None, None, None, ctypes.byref(n))
        return n.value

    @classmethod
    def get(cls):
        if cls._inst is None:
            cls._inst = Context()
        return cls._inst

    def sync(self):
        call("pyn_sync", self.handle)

    def timer_start(self):
        call("pyn_timer_start", self.handle)

    def timer_stop(self):
        ms = c_dbl()
        call("pyn_timer_stop", self.handle, ctypes.byref(ms))
        return ms.value

    def launch_count(self):
        n = c_i64()
        call("pyn_launch_count", self.handle, ctypes.byref(n))
        return n.value

    def fp64_peak(self):
        """measured DFMA throughput of this device, TFLOP/s (pyn_fp64_peak)"""
        v = c_dbl()
        call("pyn_fp64_peak", self.handle, ctypes.byref(v))
        return v.value

    def flush_l2(self):
        call("pyn_flush_l2", self.handle)

    def mem_used(self):
        n = c_i64()
        call("pyn_mem_used", self.handle, ctypes.byref(n))
        return n.value

    PROF_CLASSES = {"spmv_ksp": 0, "spmv": 1, "cg_update": 2, "cg_direction": 3, "elem_kle": 4,
                    "numeric": 5}

    def prof_enable(self, on=True):
        call("pyn_prof_enable", self.handle, 1 if on else 0)

    def prof_get(self, name):
        """(launches, summed device ms) of a kernel class since prof_enable(True)"""
        n, ms = c_i64(), c_dbl()
        call("pyn_prof_get", self.handle, self.PROF_CLASSES[name], ctypes.byref(n), ctypes.byref(ms))
        return n.value, ms.value

    @staticmethod
    def pinned_empty(n):
        """page-locked host staging buffer of n doubles as a numpy array"""
        p = vp()
        call("pyn_host_alloc", int(n) * 8, ctypes.byref(p))
        buf = (ctypes.c_double * int(n)).from_address(p.value)
        a = np.frombuffer(buf, dtype=np.float64)
        return a


def _nccl_library():
    """path of the NCCL shipped with torch's wheels (newer than the system one), else None"""
    override = os.environ.get("PYN_NCCL_LIB")
    if override:
        return override
    try:
        import importlib.util
        spec = importlib.util.find_spec("nvidia.nccl")
        for base in (spec.submodule_search_locations if spec else []):
            cand = os.path.join(base, "lib", "libnccl.so.2")
            if os.path.exists(cand):
                return cand
    except Exception:
        pass
    return None


class Layout:
    """Node layout of a row-partitioned mesh on this rank (pynama_b200/parallel.py) + the device
    halo-exchange plan.  Vecs/Mats of a partitioned run carry one."""

    def __init__(self, part, n_global_nodes):
        self.part = part
        self.n_own, self.n_loc, self.n_glob = part.n_own, part.n_loc, int(n_global_nodes)
        ranks, send_ptr, send_nodes, recv_off, recv_cnt = part.halo_tables()
        h = vp()
        call("pyn_halo_create", Context.get().handle, self.n_own, self.n_loc, len(ranks), ptr(ranks),
             ptr(send_ptr), ptr(send_nodes), ptr(recv_off), ptr(recv_cnt), ctypes.byref(h))
        self._h = h
        self.neighbors = [int(r) for r in ranks]
        self.halo_nodes = int(send_ptr[-1]), int(recv_cnt.sum())
        ctx = Context.get()
        if getattr(ctx, "p2p", False):
            # every rank exports its receive window; senders store straight into it
            hb = ctypes.create_string_buffer(64)
            call("pyn_halo_window_export", h, 6, hb)
            info = COMM_WORLD.allgather((bytes(hb.raw), [int(r) for r in ranks], [int(o) for o in recv_off],
                                         int(part.n_ghost)))
            handles, peer_off, peer_ng = [], [], []
            for q in self.neighbors:
                qh, qranks, qoff, qng = info[q]
                handles.append(qh)
                peer_off.append(qoff[qranks.index(part.rank)])
                peer_ng.append(qng)
            po, pn = np.asarray(peer_off, dtype=np.int64), np.asarray(peer_ng, dtype=np.int64)
            call("pyn_halo_window_attach", h, ctypes.create_string_buffer(b"".join(handles), 64 * max(len(handles), 1)),
                 ptr(po), ptr(pn))

    def exchange(self, vec):
        call("pyn_halo_exchange", self._h, vec._h, vec._bs)


class ScatterPlan:
    """Fixed index list + per-call values: ``Vec.setValues(idx, vals)`` for an index set that
    never changes (the boundary dofs).  Indices live on the device; values go through pinned
    staging and a scatter kernel on the library's stream, without synchronising."""

    def __init__(self, indices):
        self._idx = np.ascontiguousarray(np.asarray(indices).ravel(), dtype=np.int64)
        self.k = self._idx.size
        self._h = None
        self._keep = None

    def _bind(self, vec):
        idx = self._idx
        if vec._layout is not None:          # global dof ids -> the ones this rank holds
            self._keep, idx = vec._layout.part.local_dofs(idx, vec._bs)
            idx = np.ascontiguousarray(idx, dtype=np.int64)
        h = vp()
        call("pyn_scatter_create", Context.get().handle, idx.size, ptr(idx), ctypes.byref(h))
        self._h = h

    def apply(self, vec, values, addv=False):
        vals = _lib.as_f64(values).ravel()
        if vals.size != self.k:
            raise ValueError(f"ScatterPlan.apply: {vals.size} values for {self.k} indices")
        if self._h is None:
            self._bind(vec)
        if self._keep is not None:
            vals = np.ascontiguousarray(vals[self._keep])
        call("pyn_scatter_apply", self._h, vec._h, ptr(vals), 1 if addv else 0)

    def destroy(self):
        if getattr(self, "_h", None) is not None and _lib._lib is not None:
            _lib._lib.pyn_scatter_destroy(self._h)
        self._h = None

    def __del__(self):
        try:
            self.destroy()
        except Exception:
            pass


# ---- Vec ------------------------------------------------------------------------------------
class Vec:
    class NormType:
        NORM_2 = 2
        NORM_1 = 1
        NORM_INFINITY = 3

    def __init__(self, n=None, handle=None, name=""):
        self._h = None
        self._name = name
        self._n = 0
        self._layout, self._bs = None, 1
        if handle is not None:
            self._h = handle
            m = c_i64()
            call("pyn_vec_size", handle, ctypes.byref(m))
            self._n = m.value
        elif n is not None:
            self.createSeq(n)

    # -- construction
    def createSeq(self, size, bsize=None, comm=None):
        if isinstance(size, (tuple, list)):
            size = size[0] if size[0] is not None else size[1]
        self.destroy()
        h = vp()
        call("pyn_vec_create", Context.get().handle, int(size), ctypes.byref(h))
        self._h, self._n = h, int(size)
        return self

    createMPI = createSeq
    create = createSeq

    def createWithArray(self, array, size=None, bsize=None, comm=None):
        array = _lib.as_f64(array).ravel()
        self.createSeq(len(array))
        call("pyn_vec_from_host", self._h, ptr(array))
        return self

    def _setLayout(self, layout, bs):
        """row-partitioned Vec: n_loc*bs local entries, the first n_own*bs owned"""
        self._layout, self._bs = layout, int(bs)
        if layout is not None:
            call("pyn_vec_set_owned", self._h, layout.n_own * self._bs)
        return self

    def _like(self, name=""):
        v = Vec(self._n, name=name)
        if self._layout is not None:
            v._setLayout(self._layout, self._bs)
        return v

    def _owned(self):
        return self._n if self._layout is None else self._layout.n_own * self._bs

    def duplicate(self, array=None):
        v = self._like()
        if array is not None:
            v.setArray(array)
        return v

    def copy(self, result=None):
        if result is None:
            result = self._like(self._name)
        call("pyn_vec_copy", self._h, result._h)
        return result

    def ghostUpdate(self):
        """refresh the ghost entries from their owners (VecGhostUpdate)"""
        if self._layout is not None:
            self._layout.exchange(self)

    def destroy(self):
        if getattr(self, "_h", None) is not None and _lib._lib is not None:
            _lib._lib.pyn_vec_destroy(self._h)
        self._h = None
        return self

    def __del__(self):
        try:
            self.destroy()
        except Exception:
            pass

    # -- metadata
    def setName(self, name):
        self._name = name

    def getName(self):
        return self._name

    def getSize(self):
        return self._n if self._layout is None else self._layout.n_glob * self._bs

    def getLocalSize(self):
        return self._owned()

    def getSizes(self):
        return (self.getLocalSize(), self.getSize())

    def getOwnershipRange(self):
        return (0, self._n)

    owner_range = property(getOwnershipRange)
    size = property(getSize)

    def assemble(self):
        return None

    assemblyBegin = assemblyEnd = assemble

    # -- host <-> device
    def getArray(self, readonly=False, out=None):
        """host copy of the local (owned) part, as PETSc's Vec.getArray"""
        n = self._owned()
        if out is None:
            out = np.empty(n)
        call("pyn_vec_to_host_n", self._h, ptr(out), n)
        return out

    def getLocalArray(self):
        """owned + ghost entries"""
        out = np.empty(self._n)
        call("pyn_vec_to_host", self._h, ptr(out))
        return out

    def getGlobalArray(self):
        """the whole distributed Vec in global numbering, on every rank (tests / output)"""
        if self._layout is None:
            return self.getArray()
        lay = self._layout
        parts = COMM_WORLD.allgather((lay.part.owned, self.getArray()))
        out = np.zeros(lay.n_glob * self._bs)
        for nodes, vals in parts:
            idx = (nodes[:, None] * self._bs + np.arange(self._bs)[None, :]).ravel()
            out[idx] = vals
        return out

    array = property(getArray)

    def setArray(self, array):
        array = _lib.as_f64(array).ravel()
        if len(array) == self._n:
            call("pyn_vec_from_host", self._h, ptr(array))
        elif len(array) == self._owned():
            call("pyn_vec_from_host_n", self._h, ptr(array), len(array))
        else:
            raise ValueError(f"Vec.setArray: {len(array)} values for a Vec of local size {self._owned()}")

    def __array__(self, dtype=None, copy=None):
        a = self.getArray()
        return a if dtype is None else a.astype(dtype)

    def setValues(self, indices, values, addv=None):
        idx = np.ascontiguousarray(np.asarray(indices).ravel(), dtype=np.int64)
        vals = _lib.as_f64(values).ravel()
        if vals.size == 1 and idx.size > 1:
            vals = np.full(idx.size, vals[0])
        if idx.size != vals.size:
            raise ValueError(f"Vec.setValues: {idx.size} indices but {vals.size} values")
        if self._layout is not None:         # global ids; entries of nodes held elsewhere are dropped
            keep, idx = self._layout.part.local_dofs(idx, self._bs)
            idx, vals = np.ascontiguousarray(idx, dtype=np.int64), np.ascontiguousarray(vals[keep])
        call("pyn_vec_set_values", self._h, idx.size, ptr(idx), ptr(vals), 1 if addv else 0)

    def setValue(self, index, value, addv=None):
        self.setValues([index], [value], addv)

    def getValues(self, indices):
        idx = np.ascontiguousarray(np.asarray(indices).ravel(), dtype=np.int64)
        if self._layout is not None:
            keep, lidx = self._layout.part.local_dofs(idx, self._bs)
            if len(keep) != idx.size:
                raise PynamaError("Vec.getValues: some of the requested entries live on another rank")
            idx = np.ascontiguousarray(lidx, dtype=np.int64)
        out = np.empty(idx.size)
        call("pyn_vec_get_values", self._h, idx.size, ptr(idx), ptr(out))
        return out

    def getValue(self, index):
        return float(self.getValues([index])[0])

    # -- arithmetic (all on the device)
    def set(self, alpha):
        call("pyn_vec_set", self._h, float(alpha))

    def zeroEntries(self):
        self.set(0.0)

    def scale(self, alpha):
        call("pyn_vec_scale", self._h, float(alpha))

    def axpy(self, alpha, x):
        call("pyn_vec_axpy", self._h, float(alpha), x._h)

    def aypx(self, alpha, x):
        call("pyn_vec_aypx", self._h, float(alpha), x._h)

    def waxpy(self, alpha, x, y):
        call("pyn_vec_waxpy", self._h, float(alpha), x._h, y._h)

    def axpby(self, alpha, beta, x):
        call("pyn_vec_axpby", self._h, float(alpha), x._h, float(beta))

    def maxpy(self, alphas, vecs):
        a = _lib.as_f64(alphas)
        arr = (vp * len(vecs))(*[v._h for v in vecs])
        call("pyn_vec_maxpy", self._h, len(vecs), ptr(a), arr)

    def pointwiseMult(self, x, y):
        call("pyn_vec_pointwise_mult", self._h, x._h, y._h)

    def reciprocal(self):
        call("pyn_vec_reciprocal", self._h)

    def dot(self, other):
        out = c_dbl()
        call("pyn_vec_dot", self._h, other._h, ctypes.byref(out))
        return out.value

    def norm(self, norm_type=None):
        if norm_type not in (None, 2, Vec.NormType.NORM_2):
            a = np.abs(self.getArray())                      # owned entries
            if norm_type == 1:
                v, op = float(a.sum()), "sum"
            else:
                v, op = float(a.max(initial=0.0)), "max"
            if self._layout is not None and COMM_WORLD.size > 1:
                v = float(Context.get().allreduce([v], op)[0])
            return v
        out = c_dbl()
        call("pyn_vec_norm2", self._h, ctypes.byref(out))
        return out.value

    def sum(self):
        out = c_dbl()
        call("pyn_vec_sum", self._h, ctypes.byref(out))
        return out.value

    def __mul__(self, other):
        r = self._like()
        if isinstance(other, Vec):
            r.pointwiseMult(self, other)
        else:
            call("pyn_vec_axpby", r._h, float(other), self._h, 0.0)
        return r

    __rmul__ = __mul__

    def __imul__(self, alpha):
        if isinstance(alpha, Vec):
            self.pointwiseMult(self, alpha)
        else:
            self.scale(alpha)
        return self

    def __add__(self, other):
        r = self._like()
        r.waxpy(1.0, self, other)
        return r

    def __sub__(self, other):
        r = self._like()
        r.waxpy(-1.0, other, self)
        return r

    def __iadd__(self, other):
        self.axpy(1.0, other)
        return self

    def __isub__(self, other):
        self.axpy(-1.0, other)
        return self

    def __neg__(self):
        return self * (-1.0)

    def __len__(self):
        return self._n


# ---- Mat ------------------------------------------------------------------------------------
class Mat:
    """Handle on a device node-block matrix assembled by the library."""

    def __init__(self, handle=None, name=""):
        self._h = handle
        self._name = name
        self._shape = None
        self._layout = None
        self._matfree = False

    def _setLayout(self, layout):
        """matrix of a row-partitioned mesh: owned rows only, ghost refresh before each product"""
        self._layout = layout
        call("pyn_mat_set_halo", self._h, layout._h if layout is not None else None)
        return self

    def setMatFree(self, on=True):
        """products of the KLE matrix K evaluated element by element (sum-factorised quadrature)
        instead of from the stored blocks; same result up to summation order"""
        call("pyn_mat_set_matfree", self._h, 1 if on else 0)
        self._matfree = bool(on)
        return self

    def _info(self):
        if self._shape is None:
            nr, nc, br, bc = c_i64(), c_i64(), c_int(), c_int()
            call("pyn_mat_shape", self._h, ctypes.byref(nr), ctypes.byref(nc), ctypes.byref(br),
                 ctypes.byref(bc))
            self._shape = (nr.value, nc.value, br.value, bc.value)
        return self._shape

    def setName(self, name):
        self._name = name

    def getName(self):
        return self._name

    def getSize(self):
        s = self._info()
        if self._layout is not None:
            return (self._layout.n_glob * s[2], self._layout.n_glob * s[3])
        return (s[0], s[1])

    def getLocalSize(self):
        s = self._info()
        if self._layout is not None:
            return (self._layout.n_own * s[2], self._layout.n_own * s[3])
        return (s[0], s[1])

    def getSizes(self):
        g, l = self.getSize(), self.getLocalSize()
        return ((l[0], g[0]), (l[1], g[1]))

    def getBlockSizes(self):
        s = self._info()
        return (s[2], s[3])

    def getOwnershipRange(self):
        return (0, self._info()[0])

    def createVecRight(self):
        s = self._info()
        return Vec(s[1])._setLayout(self._layout, s[3])

    def createVecLeft(self):
        s = self._info()
        return Vec(s[0])._setLayout(self._layout, s[2])

    def createVecs(self):
        return self.createVecRight(), self.createVecLeft()

    def assemble(self):
        return None

    assemblyBegin = assemblyEnd = setUp = assemble

    def mult(self, x, y):
        call("pyn_mat_mult", self._h, x._h, y._h)

    def multAdd(self, x, v, y):
        """y = v + A x"""
        call("pyn_mat_mult_axpby", self._h, 1.0, x._h, 1.0, v._h, y._h)

    def __mul__(self, x):
        if isinstance(x, Vec):
            y = self.createVecLeft()
            self.mult(x, y)
            return y
        return NotImplemented

    def __add__(self, other):
        h = vp()
        call("pyn_mat_add_new", self._h, other._h, ctypes.byref(h))
        m = Mat(h, name=f"{self._name}+{other._name}")
        m._keep = (self, other, getattr(self, "_keep", None))
        m._layout = self._layout            # the library copies the halo plan of the left operand
        return m

    def getDiagonal(self, result=None):
        if result is None:
            result = self.createVecLeft()
        call("pyn_mat_get_diagonal", self._h, result._h)
        return result

    def diagonalScale(self, L=None, R=None):
        if R is not None:
            raise PynamaError("Mat.diagonalScale(R=) is not used by the hot path and not implemented")
        if L is not None:
            call("pyn_mat_scale_rows", self._h, L._h)

    def nnz(self):
        """(nnz of the reference AIJ pattern incl. explicit zeros, stored node blocks)"""
        a, b = c_i64(), c_i64()
        call("pyn_mat_nnz", self._h, ctypes.byref(a), ctypes.byref(b))
        return a.value, b.value

    def nblocks(self):
        b = c_i64()
        call("pyn_mat_nnz", self._h, None, ctypes.byref(b))
        return b.value

    def bytes(self):
        """(bytes held on the device, compulsory bytes of one SpMV)"""
        a, b = c_i64(), c_i64()
        call("pyn_mat_bytes", self._h, ctypes.byref(a), ctypes.byref(b))
        return a.value, b.value

    def getValuesCSR(self):
        """(indptr, indices, data) of the scalar CSR in the reference's pattern."""
        nnz, _ = self.nnz()
        nrows = self._info()[0]
        rowptr = np.zeros(nrows + 1, dtype=np.int64)
        col = np.zeros(nnz, dtype=np.int32)
        val = np.zeros(nnz)
        call("pyn_mat_export_csr", self._h, ptr(rowptr), ptr(col), ptr(val))
        return rowptr, col, val

    def getRow(self, row):
        """(column indices, values) of one row of the reference pattern (host-side diagnostic)"""
        ip, ix, v = self.getValuesCSR()
        return ix[ip[row]:ip[row + 1]], v[ip[row]:ip[row + 1]]

    def getInfo(self):
        nnz, nb = self.nnz()
        stored, _ = self.bytes()
        return {"nz_used": float(nnz), "nz_allocated": float(nnz), "nz_unneeded": 0.0,
                "memory": float(stored), "block_size": float(self._info()[2]), "blocks": float(nb)}

    def isSymmetric(self, tol=0.0):
        import scipy.sparse as sp  # host-side diagnostic only (mat_fs.py:129)
        ip, ix, v = self.getValuesCSR()
        a = sp.csr_matrix((v, ix, ip), shape=self.getSize())
        d = abs(a - a.T)
        return bool(d.max() <= tol) if d.nnz else True

    def destroy(self):
        if getattr(self, "_h", None) is not None and _lib._lib is not None:
            _lib._lib.pyn_mat_destroy(self._h)
        self._h = None
        return self

    def __del__(self):
        try:
            self.destroy()
        except Exception:
            pass


# ---- KSP / PC ---------------------------------------------------------------------------------
class PC:
    def __init__(self):
        self._type = "bjacobi"

    def create(self, comm=None):
        return self

    def setType(self, t):
        self._type = str(t)

    def getType(self):
        return self._type

    def setFromOptions(self):
        t = Options().getString("pc_type", None)
        if t:
            self._type = t


class KSP:
    """PETSc.KSP surface used by KspSolver (solver/kle_solver.py:49-64).

    Types: cg (production path: the KLE operator is SPD), gmres (restart
    ``-ksp_gmres_restart``, default 30) and bcgs, all left-preconditioned with
    none / jacobi / bjacobi / asm (additive Schwarz over the spectral elements, one
    GPU).  ``-pc_type lu`` -- the reference default, a direct
    solve -- is mapped to CG + node-block Jacobi at rtol 1e-12 and logged, because
    a sparse direct factorisation is not part of the B200 path.
    """
    _TYPES = {"cg": 0, "gmres": 1, "bcgs": 2, "preonly": 3}
    _PCS = {"none": 0, "jacobi": 1, "bjacobi": 2, "pbjacobi": 2, "asm": 3}

    def __init__(self):
        self._h = None
        self._init_fields()

    def _init_fields(self):
        self._type = "cg"
        self._pc = PC()
        self._mat = None
        self.rtol, self.atol, self.max_it, self.restart = 1e-5, 1e-50, 10000, 30
        self._its, self._reason, self._rnorm = 0, 0, 0.0
        self._guess_nonzero = True
        self._fischer = (0, 0)              # (model, size); KSPSetUseFischerGuess, off by default
        self._cg_single_reduction = False   # KSPCGUseSingleReduction
        self._asm_coarse = 0                # -pc_asm_coarse: Chebyshev steps of the Schwarz coarse level
        self._log = logging.getLogger("KSP")

    def create(self, comm=None):
        if not hasattr(self, "_pc"):
            self._h = None
            self._init_fields()
        return self

    def setType(self, t):
        self._type = str(t)
        self._drop()

    def getType(self):
        return self._type

    def setPC(self, pc):
        self._pc = pc
        self._drop()

    def getPC(self):
        return self._pc

    def setOperators(self, A, P=None):
        self._mat = A
        self._drop()

    def getOperators(self):
        return self._mat, self._mat

    def setTolerances(self, rtol=None, atol=None, divtol=None, max_it=None):
        if rtol is not None:
            self.rtol = rtol
        if atol is not None:
            self.atol = atol
        if max_it is not None:
            self.max_it = max_it

    def setInitialGuessNonzero(self, flag):
        self._guess_nonzero = bool(flag)

    def setUseFischerGuess(self, model, size):
        """KSPSetUseFischerGuess: model 1 = A-conjugate projection on the last `size` solutions"""
        self._fischer = (int(model), int(size))
        self._drop()

    def setCGSingleReduction(self, flag=True):
        """KSPCGUseSingleReduction: one fused reduction and two launches per CG iteration"""
        self._cg_single_reduction = bool(flag)
        self._drop()

    def setIterationHint(self, n):
        """expected iteration count of the next solve (sizes the first chunk of launches; result unchanged)"""
        self._hint = int(n)

    def setASMCoarse(self, steps):
        """second level of -pc_type asm: Q1 coarse space, `steps` Chebyshev steps (0 = off)"""
        self._asm_coarse = int(steps)
        self._drop()

    def setFromOptions(self):
        o = Options()
        self._type = o.getString("ksp_type", self._type)
        self._asm_coarse = o.getInt("pc_asm_coarse", self._asm_coarse)
        sr = o.getBool("ksp_cg_single_reduction", None)
        if sr is not None:
            self._cg_single_reduction = sr
        self._pc.setFromOptions()
        self.rtol = o.getReal("ksp_rtol", self.rtol)
        self.atol = o.getReal("ksp_atol", self.atol)
        self.max_it = o.getInt("ksp_max_it", self.max_it)
        self.restart = o.getInt("ksp_gmres_restart", self.restart)
        g = o.getBool("ksp_initial_guess_nonzero", None)
        if g is not None:
            self._guess_nonzero = g
        fg = o.getString("ksp_fischer_guess", None)          # -ksp_fischer_guess <model>,<size>
        if fg:
            parts = [int(p) for p in str(fg).replace(" ", "").split(",") if p]
            self._fischer = (parts[0], parts[1] if len(parts) > 1 else 10)
        self._drop()

    def _drop(self):
        if getattr(self, "_h", None) is not None and _lib._lib is not None:
            _lib._lib.pyn_ksp_destroy(self._h)
        self._h = None

    def setUp(self):
        if self._h is not None:
            return
        if self._mat is None:
            raise PynamaError("KSP.setUp: no operator set")
        ktype, pctype, rtol = self._type, self._pc.getType(), self.rtol
        self._direct_mapped = False
        if pctype in ("lu", "cholesky") or ktype == "preonly":
            # the reference default (kle_solver.py:58-60, makefile:7).  Only the KLE operator K is known
            # to be symmetric positive definite; anything else (K + Kfs, the immersed-boundary H S) gets
            # GMRES.  A direct solve cannot "fail to converge", so solve() raises if the mapped one does.
            spd = getattr(self._mat, "_name", "") == "K"
            mapped = "cg" if spd else "gmres"
            self._log.warning("direct solve (-ksp_type %s -pc_type %s) requested on '%s': there is no sparse "
                              "factorisation on the GPU path; using -ksp_type %s -pc_type bjacobi -ksp_rtol "
                              "1e-12 instead", ktype, pctype, getattr(self._mat, "_name", "?"), mapped)
            ktype, pctype, rtol = mapped, "bjacobi", min(self.rtol, 1e-12)
            self._direct_mapped = True
        if ktype not in self._TYPES:
            raise PynamaError(f"unknown KSP type '{ktype}'")
        if pctype not in self._PCS:
            raise PynamaError(f"unsupported PC type '{pctype}' (none, jacobi, bjacobi/pbjacobi, asm)")
        h = vp()
        call("pyn_ksp_create", Context.get().handle, self._mat._h, self._TYPES[ktype], self._PCS[pctype],
             ctypes.byref(h))
        self._h = h
        call("pyn_ksp_set_tolerances", h, float(rtol), float(self.atol), int(self.max_it),
             int(self.restart))
        if self._fischer[1] > 0:
            call("pyn_ksp_set_fischer_guess", h, self._fischer[0], min(self._fischer[1], 32))
        if self._cg_single_reduction and ktype == "cg":
            call("pyn_ksp_set_cg_single_reduction", h, 1)
        if pctype == "asm" and self._asm_coarse > 0:
            call("pyn_ksp_set_asm_coarse", h, self._asm_coarse)

    def solve(self, b, x):
        self.setUp()
        if not self._guess_nonzero and self._fischer[1] == 0:
            x.set(0.0)
        hint = getattr(self, "_hint", -1)
        if hint >= 0:
            call("pyn_ksp_hint_iterations", self._h, hint)
            self._hint = -1
        its, reason, rn = c_int(), c_int(), c_dbl()
        call("pyn_ksp_solve", self._h, b._h, x._h, ctypes.byref(its), ctypes.byref(reason),
             ctypes.byref(rn))
        self._its, self._reason, self._rnorm = its.value, reason.value, rn.value
        if reason.value < 0:
            msg = "KSP did not converge: reason %d%s after %d iterations, residual %.3e" % (
                reason.value, " (indefinite operator: p.Ap < 0 in CG)" if reason.value == -10 else "",
                its.value, rn.value)
            self.failures = getattr(self, "failures", 0) + 1           # sticky: callers may check
            if getattr(self, "_direct_mapped", False):
                raise PynamaError(msg + " -- the requested direct solve (-pc_type lu) was mapped to this "
                                        "Krylov solve, which must converge")
            self._log.warning(msg)

    def __call__(self, b, x=None):
        if x is None:
            x = self._mat.createVecRight()
        self.solve(b, x)
        return x

    def getIterationNumber(self):
        return self._its

    def getConvergedReason(self):
        return self._reason

    def getResidualNorm(self):
        return self._rnorm

    its = property(getIterationNumber)
    reason = property(getConvergedReason)
    norm = property(getResidualNorm)

    def destroy(self):
        self._drop()
        return self

    def __del__(self):
        try:
            self._drop()
        except Exception:
            pass
