"""A numpy / scipy stand-in for the part of `petsc4py.PETSc` that the reference's hot path touches.

TEST INFRASTRUCTURE (like oracle/): it exists so that the reference's OWN, UNMODIFIED Python modules --
cases/base_problem.py, matrices/mat_fs.py, matrices/mat_ns.py, solver/kle_solver.py, domain/domain.py,
boundaries/*.py, domain/elements/*.py, functions/*.py -- can be imported and run in the build container,
which has no petsc4py, no PETSc and no MPI.  What those modules compute on this stand-in is recorded as
golden vectors (tests/golden/make_golden_refpath.py) and the oracle is pinned against them.
Nothing under pynama_b200/ imports this file.

Semantics restated from the PETSc manual pages (PETSc 3.12, the version the reference's README names):

* Mat (AIJ): `setValues(rows, cols, block, addv)` adds or inserts a dense logical block; EVERY inserted
  entry stays in the nonzero pattern, zeros included (MAT_IGNORE_ZERO_ENTRIES is off by default); values are
  accumulated in call order.  `A + B` is MatAXPY with DIFFERENT_NONZERO_PATTERN: the union pattern.
  `diagonalScale(L)` scales rows.  Preallocation (`nnz=`) is recorded, and `assemble()` notes in
  `prealloc_exceeded` the rows that received more entries than were preallocated (PETSc would raise a
  new-nonzero-allocation error there): the golden script asserts that list is empty.
* Vec: a dense array; `setValues(..., addv)`, `duplicate()` (zeroed), `copy()`, `*` is pointwise for Vec * Vec.
* IS: block and general index sets, `union` = sorted unique union (ISExpand).
* KSP / PC: the reference's default is `gmres` preconditioned by `lu` (kle_solver.py:58-60) = one exact
  solve; `solve` here is scipy's sparse LU.  Options are ignored: this stand-in is the DEFAULT configuration.
* TS: only what ts_solver.py and base_problem.py call; `step()` is a fixed-step explicit RK loop over the
  Bogacki-Shampine 5(4) tableau taken from oracle/timestep.py (PETSc's TSRK5BS, FSAL) that passes a STAGE
  work vector to the RHS callback and updates the solution vector in place, as TSStep_RK does.  The time
  integrator is therefore the stand-in's, the right-hand side it calls is the reference's.
* COMM_WORLD: one rank.
"""
import numpy as np
import scipy.sparse as sp
import scipy.sparse.linalg as spl


class _Comm:
    rank = 0
    size = 1

    def getRank(self):
        return 0

    def getSize(self):
        return 1

    def tompi4py(self):
        return self

    def allgather(self, obj):
        return [obj]

    def barrier(self):
        pass


COMM_WORLD = _Comm()
COMM_SELF = _Comm()
DECIDE = None


class InsertMode:
    INSERT_VALUES = INSERT = False
    ADD_VALUES = ADD = True


class NormType:
    NORM_1, NORM_2, NORM_FROBENIUS, NORM_INFINITY = 0, 1, 2, 3
    N1, N2, FROBENIUS, INFINITY = 0, 1, 2, 3


class Options:
    _db = {}

    def __init__(self, prefix=None):
        pass

    def _get(self, name, default, conv):
        return conv(self._db[name]) if name in self._db else default

    def getString(self, name, default=None):
        return self._get(name, default, str)

    def getInt(self, name, default=None):
        return self._get(name, default, int)

    def getReal(self, name, default=None):
        return self._get(name, default, float)

    def getBool(self, name, default=None):
        return self._get(name, default, lambda v: str(v).lower() in ("1", "true", "yes"))

    def hasName(self, name):
        return name in self._db

    def setValue(self, name, value):
        self._db[name] = value

    def delValue(self, name):
        self._db.pop(name, None)


def _local_size(size):
    """petsc4py sizes: n, (n, N) or ((n, N)) with None for 'decide'"""
    if isinstance(size, (tuple, list)):
        return int(size[0] if size[0] is not None else size[1])
    return int(size)


class IS:
    def __init__(self):
        self._bs, self._blocks = 1, np.zeros(0, dtype=np.int32)

    def createGeneral(self, indices, comm=None):
        self._bs, self._blocks = 1, np.asarray(list(indices), dtype=np.int32).ravel()
        return self

    def createBlock(self, bsize, indices, comm=None):
        self._bs, self._blocks = int(bsize), np.asarray(list(indices), dtype=np.int32).ravel()
        return self

    def getBlockSize(self):
        return self._bs

    def getBlockIndices(self):
        return self._blocks.copy()

    def getIndices(self):
        return (self._blocks[:, None] * self._bs + np.arange(self._bs, dtype=np.int32)[None, :]).ravel()

    def getSize(self):
        return self._blocks.size * self._bs

    def union(self, other):
        return IS().createGeneral(np.union1d(self.getIndices(), other.getIndices()))

    def destroy(self):
        self._destroyed = True
        return self

    def getRefCount(self):
        return 0 if getattr(self, "_destroyed", False) else 1


class Vec:
    def __init__(self):
        self._a = np.zeros(0)
        self._name = ""

    # -- creation
    def createWithArray(self, array, size=None, bsize=None, comm=None):
        self._a = np.ascontiguousarray(array, dtype=np.float64).ravel()
        return self

    def createSeq(self, size, bsize=None, comm=None):
        self._a = np.zeros(_local_size(size))
        return self

    createMPI = createSeq

    def duplicate(self, array=None):
        v = Vec()
        v._a = np.zeros_like(self._a) if array is None else np.array(array, dtype=np.float64)
        return v

    def copy(self, result=None):
        if result is None:
            result = Vec()
            result._name = self._name
            result._a = self._a.copy()
        else:
            result._a[...] = self._a
        return result

    def destroy(self):
        return self

    # -- names, sizes
    def setName(self, name):
        self._name = name

    def getName(self):
        return self._name

    def getSize(self):
        return self._a.size

    getLocalSize = getSize
    size = property(getSize)

    def getSizes(self):
        return self._a.size, self._a.size

    def getOwnershipRange(self):
        return 0, self._a.size

    owner_range = property(getOwnershipRange)

    # -- values
    def getArray(self, readonly=False):
        return self._a

    array = property(getArray)

    def __array__(self, dtype=None, copy=None):
        return self._a if dtype is None else self._a.astype(dtype)

    def __len__(self):
        return self._a.size

    def setArray(self, array):
        self._a[...] = np.asarray(array, dtype=np.float64).ravel()

    def getValues(self, indices):
        return self._a[np.asarray(indices, dtype=np.int64)].copy()

    def getValue(self, index):
        return float(self._a[index])

    def setValues(self, indices, values, addv=None):
        idx = np.asarray(indices, dtype=np.int64).ravel()
        val = np.asarray(values, dtype=np.float64).ravel()
        if val.size == 1 and idx.size != 1:
            val = np.full(idx.size, val[0])
        assert idx.size == val.size, (idx.size, val.size)
        if addv:
            np.add.at(self._a, idx, val)
        else:
            self._a[idx] = val             # later writes to the same index win, as VecSetValues(INSERT)

    def setValue(self, index, value, addv=None):
        self.setValues([index], [value], addv)

    def set(self, alpha):
        self._a[...] = alpha

    def zeroEntries(self):
        self._a[...] = 0.0

    def assemble(self):
        pass

    assemblyBegin = assemblyEnd = assemble

    # -- arithmetic
    def scale(self, alpha):
        self._a *= alpha

    def axpy(self, alpha, x):
        self._a += alpha * x._a

    def aypx(self, alpha, x):
        self._a[...] = alpha * self._a + x._a

    def waxpy(self, alpha, x, y):
        self._a[...] = alpha * x._a + y._a

    def pointwiseMult(self, x, y):
        self._a[...] = x._a * y._a

    def reciprocal(self):
        nz = self._a != 0.0
        self._a[nz] = 1.0 / self._a[nz]

    def dot(self, other):
        return float(np.dot(self._a, other._a))

    def norm(self, norm_type=None):
        """NORM_1 = 0, NORM_2 = 1, NORM_FROBENIUS = 2 (what the reference passes: the 2-norm), NORM_INFINITY = 3"""
        if norm_type in (None, 1, 2):
            return float(np.sqrt(np.dot(self._a, self._a)))
        if norm_type == 0:
            return float(np.abs(self._a).sum())
        return float(np.abs(self._a).max())

    def max(self):
        i = int(np.argmax(self._a))
        return i, float(self._a[i])

    def min(self):
        i = int(np.argmin(self._a))
        return i, float(self._a[i])

    def sum(self):
        return float(self._a.sum())

    def _binary(self, other, op):
        out = self.duplicate()
        out._a[...] = op(self._a, other._a if isinstance(other, Vec) else other)
        return out

    def __add__(self, other):
        return self._binary(other, np.add)

    __radd__ = __add__

    def __sub__(self, other):
        return self._binary(other, np.subtract)

    def __mul__(self, other):
        return self._binary(other, np.multiply)

    __rmul__ = __mul__

    def __neg__(self):
        return self._binary(-1.0, np.multiply)

    def __iadd__(self, other):
        self._a += other._a if isinstance(other, Vec) else other
        return self

    def __isub__(self, other):
        self._a -= other._a if isinstance(other, Vec) else other
        return self

    def __imul__(self, other):
        self._a *= other._a if isinstance(other, Vec) else other
        return self

    def __getitem__(self, i):
        return self._a[i]

    def __setitem__(self, i, v):
        self._a[i] = v


class Mat:
    """AIJ matrix as an insertion-ordered dict (i, j) -> value: accumulation order = call order"""

    def __init__(self):
        self._shape = (0, 0)
        self._d = {}
        self._name = ""
        self._nnz = None
        self.prealloc_exceeded = []
        self._csr = None

    def createAIJ(self, size, bsize=None, nnz=None, csr=None, comm=None):
        if isinstance(size, (tuple, list)) and isinstance(size[0], (tuple, list)):
            self._shape = (_local_size(size[0]), _local_size(size[1]))
        elif isinstance(size, (tuple, list)):
            self._shape = (int(size[0]), int(size[1]))
        else:
            self._shape = (int(size), int(size))
        if nnz is not None:
            d, o = nnz if isinstance(nnz, (tuple, list)) and len(nnz) == 2 and np.ndim(nnz[0]) > 0 else (nnz, 0)
            self._nnz = np.broadcast_to(np.asarray(d, dtype=np.int64), (self._shape[0],)) + \
                np.broadcast_to(np.asarray(o, dtype=np.int64), (self._shape[0],))
        return self

    def setUp(self):
        return self

    def setName(self, name):
        self._name = name

    def getName(self):
        return self._name

    def getSize(self):
        return self._shape

    getLocalSize = getSize
    size = property(getSize)

    def getOwnershipRange(self):
        return 0, self._shape[0]

    def setOption(self, *a, **k):
        pass

    def setValues(self, rows, cols, values, addv=None):
        rows = np.atleast_1d(np.asarray(rows, dtype=np.int64)).ravel()
        cols = np.atleast_1d(np.asarray(cols, dtype=np.int64)).ravel()
        vals = np.asarray(values, dtype=np.float64)
        if vals.size == 1 and rows.size * cols.size != 1:
            vals = np.full((rows.size, cols.size), float(vals.ravel()[0]))
        vals = vals.reshape(rows.size, cols.size)
        self._csr = None
        d = self._d
        if rows.size and (rows.min() < 0 or rows.max() >= self._shape[0]) or \
                cols.size and (cols.min() < 0 or cols.max() >= self._shape[1]):
            raise IndexError(f"Mat {self._name}: index out of range")
        for a, i in enumerate(rows.tolist()):
            for b, j in enumerate(cols.tolist()):
                key = (i, j)
                if addv:
                    d[key] = d.get(key, 0.0) + vals[a, b]
                else:
                    d[key] = vals[a, b]

    def setValue(self, row, col, value, addv=None):
        self.setValues([row], [col], [value], addv)

    def assemble(self, assembly=None):
        if self._nnz is not None:
            cnt = np.zeros(self._shape[0], dtype=np.int64)
            for (i, _j) in self._d:
                cnt[i] += 1
            self.prealloc_exceeded = np.flatnonzero(cnt > self._nnz).tolist()
        return self

    assemblyBegin = assemblyEnd = assemble

    def tocsr(self):
        """sorted-column CSR with every inserted entry, explicit zeros included"""
        if self._csr is None:
            n = len(self._d)
            r = np.fromiter((k[0] for k in self._d), dtype=np.int64, count=n)
            c = np.fromiter((k[1] for k in self._d), dtype=np.int64, count=n)
            v = np.fromiter(self._d.values(), dtype=np.float64, count=n)
            order = np.lexsort((c, r))
            indptr = np.zeros(self._shape[0] + 1, dtype=np.int64)
            np.add.at(indptr, r + 1, 1)
            self._csr = sp.csr_matrix((v[order], c[order], np.cumsum(indptr)), shape=self._shape)
        return self._csr

    def getValuesCSR(self):
        m = self.tocsr()
        return m.indptr.copy(), m.indices.copy(), m.data.copy()

    def getRow(self, row):
        m = self.tocsr()
        lo, hi = m.indptr[row], m.indptr[row + 1]
        return m.indices[lo:hi].astype(np.int32), m.data[lo:hi].copy()

    def getInfo(self, info=None):
        return {"memory": 12.0 * len(self._d), "nz_unneeded": 0.0, "nz_used": float(len(self._d)),
                "nz_allocated": float(len(self._d)), "mallocs": 0.0}

    def isSymmetric(self, tol=0):
        m = self.tocsr()
        return bool(abs(m - m.T).max() <= tol) if m.shape[0] == m.shape[1] else False

    def createVecRight(self):
        return Vec().createSeq(self._shape[1])

    def createVecLeft(self):
        return Vec().createSeq(self._shape[0])

    def createVecs(self):
        return self.createVecRight(), self.createVecLeft()

    def mult(self, x, y):
        y._a[...] = self.tocsr() @ x._a

    def multAdd(self, x, v, y):
        y._a[...] = v._a + self.tocsr() @ x._a

    def multTranspose(self, x, y):
        y._a[...] = self.tocsr().T @ x._a

    def diagonalScale(self, L=None, R=None):
        self._csr = None
        for (i, j) in self._d:
            if L is not None:
                self._d[(i, j)] *= L._a[i]
            if R is not None:
                self._d[(i, j)] *= R._a[j]

    def getDiagonal(self, result=None):
        result = result or self.createVecLeft()
        result._a[...] = [self._d.get((i, i), 0.0) for i in range(self._shape[0])]
        return result

    def scale(self, alpha):
        self._csr = None
        for k in self._d:
            self._d[k] *= alpha

    def axpy(self, alpha, X, structure=None):
        self._csr = None
        for k, v in X._d.items():
            self._d[k] = self._d.get(k, 0.0) + alpha * v

    def duplicate(self, copy=False):
        m = Mat()
        m._shape, m._name = self._shape, self._name
        m._d = dict(self._d) if copy else {k: 0.0 for k in self._d}
        return m

    def copy(self, result=None, structure=None):
        return self.duplicate(copy=True)

    def zeroEntries(self):
        self._csr = None
        for k in self._d:
            self._d[k] = 0.0

    def destroy(self):
        return self

    def __mul__(self, other):
        if isinstance(other, Vec):
            y = self.createVecLeft()
            self.mult(other, y)
            return y
        m = self.duplicate(copy=True)
        m.scale(float(other))
        return m

    __rmul__ = __mul__

    def __add__(self, other):
        m = self.duplicate(copy=True)      # MatDuplicate + MatAXPY(DIFFERENT_NONZERO_PATTERN)
        m.axpy(1.0, other)
        return m

    def __sub__(self, other):
        m = self.duplicate(copy=True)
        m.axpy(-1.0, other)
        return m


class PC:
    def __init__(self):
        self._type = None

    def create(self, comm=None):
        return self

    def setType(self, t):
        self._type = t

    def getType(self):
        return self._type

    def setFromOptions(self):
        pass

    def setUp(self):
        pass

    def destroy(self):
        return self


class KSP:
    """exact solve: the reference's default gmres + lu (kle_solver.py:56-60)"""

    def __init__(self):
        self._init()

    def _init(self):
        self._type, self._pc, self._A, self._lu, self.its = None, None, None, None, 0

    def create(self, comm=None):
        self._init()
        return self

    def setType(self, t):
        self._type = t

    def getType(self):
        return self._type

    def setPC(self, pc):
        self._pc = pc

    def getPC(self):
        if getattr(self, "_pc", None) is None:
            self._pc = PC()
        return self._pc

    def setFromOptions(self):
        pass

    def setTolerances(self, rtol=None, atol=None, divtol=None, max_it=None):
        pass

    def setOperators(self, A, P=None):
        self._A, self._lu = A, None

    def getOperators(self):
        return self._A, self._A

    def setUp(self):
        self._lu = spl.splu(self._A.tocsr().tocsc())

    def solve(self, b, x):
        if self._lu is None:
            self.setUp()
        x._a[...] = self._lu.solve(b._a)
        self.its = 1

    __call__ = solve

    def getIterationNumber(self):
        return self.its

    def getConvergedReason(self):
        return 2

    def destroy(self):
        return self


class TS:
    class ProblemType:
        LINEAR, NONLINEAR = 0, 1

    class EquationType:
        UNSPECIFIED, ODE_EXPLICIT, ODE_IMPLICIT = -1, 1, 2

    class ExactFinalTime:
        UNSPECIFIED, STEPOVER, INTERPOLATE, MATCHSTEP = 0, 1, 2, 3

    def __init__(self, comm=None):
        self.create(comm)

    def create(self, comm=None):
        self._type, self._rk = None, None
        self.time, self.step_number, self._max_time, self._max_steps = 0.0, 0, 5.0, 5000
        self._dt = float(Options().getReal("ts_dt", 1e-3))
        self._rhs, self._post, self._u, self._k_last = None, None, None, None
        return self

    def setProblemType(self, t):
        pass

    def setEquationType(self, t):
        pass

    def setType(self, t):
        self._type = t

    def getType(self):
        return self._type

    def setRKType(self, t):
        self._rk = t

    def getRKType(self):
        return self._rk

    def setTime(self, t):
        self.time = float(t)

    def getTime(self):
        return self.time

    def setMaxTime(self, t):
        self._max_time = float(t)

    def getMaxTime(self):
        return self._max_time

    def setMaxSteps(self, n):
        self._max_steps = int(n)

    def getMaxSteps(self):
        return self._max_steps

    def setExactFinalTime(self, v):
        pass

    def setTimeStep(self, dt):
        self._dt = float(dt)

    def getTimeStep(self):
        return self._dt

    def getStepNumber(self):
        return self.step_number

    def setRHSFunction(self, f, vec=None, args=None, kargs=None):
        self._rhs = f

    def setPostStep(self, f, args=None, kargs=None):
        self._post = f

    def setFromOptions(self):
        pass

    def setSolution(self, u):
        self._u, self._k_last = u, None

    def getSolution(self):
        return self._u

    def step(self):
        """one fixed step of RK5BS: stage vectors are work vectors, the solution is updated in place"""
        from oracle.timestep import RK5BS_A, RK5BS_B, RK5BS_C
        assert self._rk == "5bs", "the stand-in integrates with the 5bs tableau only"
        u, h, t = self._u, self._dt, self.time
        s = len(RK5BS_B)
        k = [None] * s
        for i in range(s):
            if i == 0 and self._k_last is not None:
                k[0] = self._k_last                     # FSAL
                continue
            y = u.copy()
            for j in range(i):
                if RK5BS_A[i][j] != 0.0:
                    y.axpy(h * RK5BS_A[i][j], k[j])
            k[i] = u.duplicate()
            self._rhs(self, t + RK5BS_C[i] * h, y, k[i])
        for j in range(s):
            if RK5BS_B[j] != 0.0:
                u.axpy(h * RK5BS_B[j], k[j])
        self._k_last = k[s - 1]
        self.time, self.step_number = t + h, self.step_number + 1
        if self._post is not None:
            self._post(self)

    def solve(self, u):
        self.setSolution(u)
        while self.step_number < self._max_steps and self.time < self._max_time - 1e-14:
            self.step()

    def destroy(self):
        return self


class DMPlex:
    """base class of the reference's DMPlexDom; the mesh queries are supplied by dm_standin.py"""

    def __init__(self):
        pass


class Viewer:
    class Mode:
        READ, WRITE, APPEND = "r", "w", "a"

    def __getattr__(self, name):
        return lambda *a, **k: self


ViewerHDF5 = Viewer


class Log:
    @staticmethod
    def begin():
        pass
