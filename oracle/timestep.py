"""Oracle: KLE solve, vorticity-transport RHS and the RK5BS step (numpy/scipy, fp64).

TEST INFRASTRUCTURE -- see oracle/__init__.py.  Restates

* KleSolver.solve           src/solver/kle_solver.py:33-37   (b = Rw*w + Krhs*v; KSP(b, v))
* BaseProblem.evalRHS       src/cases/base_problem.py:111-136
* BaseProblem.computeVtensV src/cases/base_problem.py:139-154
* setValuesToVec            src/boundaries/boundary_conditions.py:252-260 (BC scatter)
* Taylor-Green 3-D fields   src/functions/taylor_green3d.py:4-33
* Taylor-Green 2-D fields   src/functions/taylor_green.py:4-36
* TS rk/5bs                 src/solver/ts_solver.py:5-23 (tableau: SURVEY.md 8a-A13;
                            PETSc's TS itself is not in the reference tree -> the
                            fixed-step stage recurrence is restated, adaptivity UNPINNED)

The free-slip (FS) path only: boundary nodes are Dirichlet with analytic values.
``evalRHS`` keeps the reference's quirk of differentiating the step-start
vorticity for every stage (base_problem.py:115,125 use self.vort, not the
stage vector TS passes in).
"""
from math import exp, pi

import numpy as np
import scipy.sparse.linalg as spl

from . import assembly, sem

# Bogacki-Shampine 5(4): same numbers as SURVEY.md 8a-A13 (checked against the order conditions
# in tests/test_oracle_kats.py through the product's table; restated here so the oracle is standalone)
RK5BS_A = [
    [],
    [1 / 6],
    [2 / 27, 4 / 27],
    [183 / 1372, -162 / 343, 1053 / 1372],
    [68 / 297, -4 / 11, 42 / 143, 1960 / 3861],
    [597 / 22528, 81 / 352, 63099 / 585728, 58653 / 366080, 4617 / 20480],
    [174197 / 959244, -30942 / 79937, 8152137 / 19744439, 666106 / 1039181, -29421 / 29068,
     482048 / 414219],
    [587 / 8064, 0, 4440339 / 15491840, 24353 / 124800, 387 / 44800, 2152 / 5985, 7267 / 94080],
]
RK5BS_B = [587 / 8064, 0, 4440339 / 15491840, 24353 / 124800, 387 / 44800, 2152 / 5985, 7267 / 94080, 0]
RK5BS_C = [0, 1 / 6, 2 / 9, 3 / 7, 2 / 3, 3 / 4, 1, 1]
# embedded fourth-order weights of the pair
RK5BS_BE = [2479 / 34992, 0, 123 / 416, 612941 / 3411720, 43 / 1440, 2272 / 6561, 79937 / 1113912, 3293 / 556956]


# ---- analytic fields -------------------------------------------------------------------------
def tg_alpha(nu, t, dim):
    """taylor_green3d.py:4-9 / taylor_green.py:4-8 (unit box)"""
    return exp(-4 * pi ** 2 * nu * t * (3.0 if dim == 3 else 2.0))


def tg_velocity(X, a):
    x = 2 * pi * X
    if X.shape[1] == 2:                                     # taylor_green.py:10-19
        return np.stack([np.cos(x[:, 0]) * np.sin(x[:, 1]) * a, -np.sin(x[:, 0]) * np.cos(x[:, 1]) * a], 1).ravel()
    return np.stack([np.cos(x[:, 0]) * np.sin(x[:, 1]) * np.sin(x[:, 2]) * a,                # :11-21
                     np.sin(x[:, 0]) * np.cos(x[:, 1]) * np.sin(x[:, 2]) * a,
                     -2 * np.sin(x[:, 0]) * np.sin(x[:, 1]) * np.cos(x[:, 2]) * a], 1).ravel()


def tg_vorticity(X, a):
    x = 2 * pi * X
    if X.shape[1] == 2:                                     # taylor_green.py:21-27
        return -2 * pi * 2.0 * np.cos(x[:, 0]) * np.cos(x[:, 1]) * a
    return np.stack([-2 * pi * 3.0 * np.sin(x[:, 0]) * np.cos(x[:, 1]) * np.cos(x[:, 2]) * a,  # :23-32
                     2 * pi * 3.0 * np.cos(x[:, 0]) * np.sin(x[:, 1]) * np.cos(x[:, 2]) * a,
                     0.0 * x[:, 0]], 1).ravel()


def vtensv(v, dim):
    """base_problem.py:139-154"""
    v = v.reshape(-1, dim)
    if dim == 2:
        return np.stack([v[:, 0] ** 2, v[:, 0] * v[:, 1], v[:, 1] ** 2], 1).ravel()
    return np.stack([v[:, 0] ** 2, v[:, 0] * v[:, 1], v[:, 1] ** 2, v[:, 1] * v[:, 2], v[:, 2] ** 2,
                     v[:, 2] * v[:, 0]], 1).ravel()


class OracleProblem:
    """Free-slip Taylor-Green problem on given mesh tables, CPU only."""

    def __init__(self, dim, ngl, corner_xyz, elem_nodes, coords, bnd_nodes, rho, mu, solver="direct",
                 rtol=1e-10, pc="jacobi", reuse=True, fischer=0):
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
        self.mats = assembly.assemble_kle(tab, corner_xyz, elem_nodes, nnodes, cls, "FS", reuse=reuse)
        self.ops = assembly.assemble_ops(tab, corner_xyz, elem_nodes, nnodes, reuse=reuse)
        self.solver, self.rtol, self.pc = solver, rtol, pc
        self.elem_nodes = np.asarray(elem_nodes, dtype=np.int64)
        self.K = self.mats["K"].tocsr()
        if solver == "direct":
            self._lu = spl.splu(self.K.tocsc())
        else:
            self._setup_pc()
        a0 = tg_alpha(self.nu, 0.0, dim)
        self.vel = tg_velocity(self.coords, a0)
        self.vort = tg_vorticity(self.coords, a0)
        self.iterations = 0
        self.solves = 0
        # PETSc -ksp_fischer_guess 1,<fischer>: A-orthonormal basis of solution increments
        self.fischer, self._fx, self._fax = int(fischer), [], []

    def _setup_pc(self):
        n, d = self.K.shape[0], self.dim
        if self.pc == "jacobi":
            dinv = 1.0 / self.K.diagonal()
            self._M = spl.LinearOperator((n, n), matvec=lambda r: dinv * r)
        elif self.pc in ("asm", "asm+coarse"):
            # PETSc PCASM with one subdomain per spectral element, symmetrically weighted by
            # 1/sqrt(elements sharing the dof):  sum_e R_e^T W inv(R_e K R_e^T) W R_e
            en = self.elem_nodes
            nel, nn = en.shape
            dofs = (en[:, :, None] * d + np.arange(d)[None, None, :]).reshape(nel, nn * d)
            inv = np.stack([np.linalg.inv(self.K[dofs[e]][:, dofs[e]].toarray()) for e in range(nel)])
            wl = (1.0 / np.sqrt(np.bincount(dofs.ravel(), minlength=n)))[dofs]

            def apply(r):
                z = np.einsum("eij,ej->ei", inv, r[dofs] * wl) * wl
                out = np.zeros_like(r)
                np.add.at(out, dofs.ravel(), z.ravel())
                return out
            if self.pc == "asm+coarse":
                apply = self._with_coarse_level(apply)
            self._M = spl.LinearOperator((n, n), matvec=apply)
        else:                       # node-block Jacobi
            import scipy.sparse as sp
            B = sp.bsr_matrix(self.K, blocksize=(d, d))
            B.sort_indices()
            nb = n // d
            blk = np.zeros((nb, d, d))
            rows = np.repeat(np.arange(nb), np.diff(B.indptr))
            sel = B.indices == rows
            blk[rows[sel]] = B.data[sel]
            inv = np.linalg.inv(blk)
            self._M = spl.LinearOperator((n, n), matvec=lambda r: np.einsum("nij,nj->ni", inv, r.reshape(nb, d)).ravel())

    def _with_coarse_level(self, one_level, steps=5):
        """-pc_asm_coarse <steps>: additive Q1 correction on the element corners.  P = the multilinear
        corner functions at the GLL nodes (rows of constrained dofs left out), Ac = P^T K P, solved by
        `steps` Chebyshev steps with point Jacobi on [0.1, 1.1] lambda_max(D^-1 Ac)"""
        import scipy.sparse as sp
        d, ngl = self.dim, self.ngl
        en = self.elem_nodes
        nel, nn = en.shape
        x1 = sem.gauss_lobatto(ngl)[0]
        t = sem.local_order(ngl, d)                                  # tensor index of every local node
        tc = sem.local_order(2, d)                                   # ... of every corner, in {0, 1}
        lin = np.stack([(1.0 - x1) / 2.0, (1.0 + x1) / 2.0], 1)      # [gll point][corner side]
        phi = np.ones((nn, len(tc)))
        for c in range(len(tc)):
            for k in range(d):
                phi[:, c] *= lin[t[:, k], tc[c, k]]
        phi[np.abs(phi) < 1e-14] = 0.0
        corner_local = [int(np.flatnonzero(np.all(t == tc[c] * (ngl - 1), axis=1))[0]) for c in range(len(tc))]
        vnodes = np.unique(en[:, corner_local])
        vid = np.full(en.max() + 1, -1)
        vid[vnodes] = np.arange(len(vnodes))
        first = {}
        for e in range(nel):                                          # a node interpolates through its first element
            for a in range(nn):
                first.setdefault(int(en[e, a]), (e, a))
        rows, cols, vals = [], [], []
        for node, (e, a) in first.items():
            for c in range(len(tc)):
                if phi[a, c] != 0.0:
                    rows.append(node), cols.append(vid[en[e, corner_local[c]]]), vals.append(phi[a, c])
        P = sp.csr_matrix((vals, (rows, cols)), shape=(en.max() + 1, len(vnodes)))
        K = self.K.tocsr()
        off = K - sp.diags(K.diagonal())
        keep = (np.asarray(abs(off).sum(1)).ravel() > 0).astype(float)
        P = (sp.diags(keep) @ sp.kron(P, sp.identity(d))).tocsr()
        Ac = (P.T @ K @ P).tocsr()
        dg = Ac.diagonal()
        dinv = np.where(dg > 1e-12 * dg.max(), 1.0 / np.where(dg == 0, 1.0, dg), 0.0)
        lmax = 1.1 * spl.eigsh(spl.LinearOperator(Ac.shape, matvec=lambda v: dinv * (Ac @ v)), k=1, which="LM",
                               return_eigenvectors=False, tol=1e-4)[0]
        theta, delta = 0.55 * lmax, 0.45 * lmax

        def coarse(b):
            x = np.zeros_like(b)
            r = b.copy()
            sigma = theta / delta
            rho = 1.0 / sigma
            dv = dinv * r / theta
            for _ in range(steps):
                x = x + dv
                r = r - Ac @ dv
                rho_new = 1.0 / (2.0 * sigma - rho)
                dv = rho_new * rho * dv + 2.0 * rho_new / delta * (dinv * r)
                rho = rho_new
            return x
        return lambda r: one_level(r) + P @ coarse(P.T @ r)

    # ---- kle_solver.py:33-37 ---------------------------------------------------------------------
    def solve_kle(self, vort):
        b = self.mats["Rw"] @ vort + self.mats["Krhs"] @ self.vel
        if self.solver == "direct":
            self.vel = self._lu.solve(b)
        else:
            its = [0]

            def cb(_):
                its[0] += 1
            x0 = self.vel
            if self.fischer:                 # the guess REPLACES the incoming one (zero for an empty basis)
                x0 = sum(((fx @ b) * fx for fx in self._fx), np.zeros_like(b))
            r0 = b - self.K @ x0
            # PETSc's default test is on the preconditioned residual; scipy tests the true
            # residual against rtol*||b|| -- same order for this SPD system.
            x, info = spl.cg(self.K, b, x0=x0, rtol=self.rtol, atol=0.0, maxiter=20000, M=self._M,
                             callback=cb)
            if self.fischer:
                if len(self._fx) == self.fischer:          # restart with the current solution
                    self._fx, self._fax = [], []
                    d, ad = x.copy(), self.K @ x
                else:
                    d, ad = x - x0, r0 - (b - self.K @ x)
                a0 = d @ ad
                for fx, fax in zip(self._fx, self._fax):
                    c = fx @ ad
                    d = d - c * fx
                    ad = ad - c * fax
                a = d @ ad
                # keep the direction only if it is well above the solver's noise floor
                tol = max(100.0 * self.rtol, 1e-12)
                if a > 0 and a0 > 0 and a > tol * tol * abs(x @ (self.K @ x)):
                    self._fx.append(d / np.sqrt(a))
                    self._fax.append(ad / np.sqrt(a))
            self.vel = x
            self.iterations += its[0]
        self.solves += 1
        return self.vel

    # ---- base_problem.py:111-136 ------------------------------------------------------------------
    def eval_rhs(self, t, vort_stage=None):
        a = tg_alpha(self.nu, t, self.dim)
        Xb = self.coords[self.bnd]
        self.vort[self.bwdofs] = tg_vorticity(Xb, a)          # :115
        self.vel[self.bdofs] = tg_velocity(Xb, a)             # :116-117
        vel = self.solve_kle(self.vort)                       # :125  (self.vort, not the stage)
        aux = self.ops["SrT"] @ vel                           # :128
        aux *= 2.0 * self.mu                                  # :129
        aux += -1.0 * self.rho * vtensv(vel, self.dim)        # :127,130
        rhs = self.ops["DivSrT"] @ aux                        # :133
        rhs *= 1.0 / self.rho                                 # :134
        return self.ops["Curl"] @ rhs                         # :136

    # ---- TS rk 5bs, fixed step ---------------------------------------------------------------------
    def step(self, t, h, k0=None):
        """one step from self.vort at time t; returns the new time and the FSAL stage"""
        s = len(RK5BS_B)
        k = [None] * s
        u = self.vort
        for i in range(s):
            if i == 0 and k0 is not None:
                k[0] = k0
                continue
            # stage vector (unused by eval_rhs, like the reference; built for fidelity of the BC
            # scatter which writes into self.vort, i.e. u itself)
            k[i] = self.eval_rhs(t + RK5BS_C[i] * h)
        unew = u.copy()
        for j in range(s):
            if RK5BS_B[j] != 0.0:
                unew += h * RK5BS_B[j] * k[j]
        self.vort = unew
        return t + h, k[s - 1]


def step_adaptive(prob, t, h, k0=None, rtol=1e-4, atol=1e-4):
    """One ACCEPTED step of an OracleProblem / OracleProblemNS under the `basic` step controller as PETSc
    documents it (TSAdaptBasic: weighted 2-norm of the embedded error estimate with tolerances atol + rtol
    max(|u|, |u_new|), safety 0.9, halved after a rejection, step factor clipped to [0.1, 10], exponent
    -1/(order-1) with order = 5).  PETSc's sources are not in the reference tree (ts_solver.py:5-23 leaves
    everything to its defaults): UNPINNED against PETSc, restated here so that the device time stepper has
    a CPU statement to agree with.  Returns (t_new, FSAL stage, next step size, rejected attempts)."""
    s = len(RK5BS_B)
    rejected = 0
    while True:
        k = [None] * s
        u0 = prob.vort.copy()
        for i in range(s):
            if i == 0 and k0 is not None:
                k[0] = k0
                continue
            k[i] = prob.eval_rhs(t + RK5BS_C[i] * h)
        u = prob.vort                       # (eval_rhs wrote the boundary values of the last stage time into it)
        unew = u.copy()
        err = np.zeros_like(u)
        for j in range(s):
            if RK5BS_B[j] != 0.0:
                unew += h * RK5BS_B[j] * k[j]
            if RK5BS_B[j] != RK5BS_BE[j]:
                err += h * (RK5BS_B[j] - RK5BS_BE[j]) * k[j]
        tol = atol + rtol * np.maximum(np.abs(u), np.abs(unew))
        enorm = float(np.sqrt(np.mean((err / tol) ** 2)))
        safety = 0.9
        accept = enorm <= 1.0
        if not accept:
            safety *= 0.5
        fac = 10.0 if enorm == 0.0 else safety * enorm ** (-1.0 / 4.0)
        hnew = h * min(max(fac, 0.1), 10.0)
        if accept:
            prob.vort = unew
            return t + h, k[s - 1], hnew, rejected
        rejected += 1
        h = hnew
        k0 = k[0]                           # still f(t, u)
        del u0


# ---- no-slip (NS) path --------------------------------------------------------------------------
SIDE_AXIS = {"left": 0, "right": 0, "down": 1, "up": 1, "back": 2, "front": 2}   # boundary.py:151-162


def side_axis(name):
    """boundary.py:151-162 for the box names; borders beyond those (a body wall) name their normal axis
    by a -x / -y / -z suffix"""
    return SIDE_AXIS[name] if name in SIDE_AXIS else "xyz".index(name[-1])


def ns_dof_classes(sides, side_nodes, nnodes, dim):
    """velocity dof classes of an all-no-slip box (0 free, 2 normal, 3 tangential).

    sides: side names in the order of the case file; side_nodes[name]: node ids.
    Tangential dofs: boundary_conditions.py:214-225; normal dofs with the left/right
    corner rule of :227-243 (single rank, collect=True); a dof that is both is normal
    (mat_ns.py:59-62)."""
    tang, normal, remove = set(), set(), set()
    for s in sides:
        ax = side_axis(s)
        nodes = np.asarray(side_nodes[s], dtype=np.int64)
        for d in range(dim):
            if d != ax:
                tang |= set((nodes * dim + d).tolist())
        loc = set((nodes * dim + ax).tolist())
        if s in ("left", "right"):
            for ind in loc:
                if (ind + 1) in normal:
                    remove.add(ind)
        normal |= loc
    normal -= remove
    cls = np.zeros(nnodes * dim, np.uint8)
    cls[np.fromiter(tang, dtype=np.int64, count=len(tang))] = 3
    cls[np.fromiter(normal, dtype=np.int64, count=len(normal))] = 2
    return cls


class OracleProblemNS:
    """Lid-driven cavity style problem: every side no-slip with a constant wall velocity.
    Restates MatNS (through oracle.assembly), KleSolver.solveFS/solve (kle_solver.py:33-41)
    and the NS branch of evalRHS (base_problem.py:119-123)."""

    def __init__(self, dim, ngl, corner_xyz, elem_nodes, coords, sides, side_nodes, wall_vel, rho, mu):
        self.dim, self.dim_w = dim, (1 if dim == 2 else 3)
        self.rho, self.mu, self.nu = rho, mu, mu / rho
        self.coords = np.asarray(coords, dtype=np.float64)
        n = len(self.coords)
        self.sides = list(sides)
        self.side_nodes = {s: np.asarray(side_nodes[s], dtype=np.int64) for s in self.sides}
        self.wall_vel = {s: np.asarray(wall_vel[s], dtype=np.float64) for s in self.sides}
        self.cls = ns_dof_classes(self.sides, self.side_nodes, n, dim)
        tab = sem.SpectralTables(ngl, dim)
        self.mats = assembly.assemble_kle(tab, corner_xyz, elem_nodes, n, self.cls, "NS")
        self.ops = assembly.assemble_ops(tab, corner_xyz, elem_nodes, n)
        self._lu = spl.splu(self.mats["K"].tocsc())
        self._lufs = spl.splu((self.mats["K"] + self.mats["Kfs"]).tocsc())      # kle_solver.py:25
        self.vel = np.zeros(n * dim)
        self.velfs = np.zeros(n * dim)
        self.vort = np.zeros(n * self.dim_w)

    def _set_sides(self, vec, ncomp, values):
        """setValuesToVec (boundary_conditions.py:252-260): sides in order, later sides overwrite"""
        v = vec.reshape(-1, ncomp)
        for s in self.sides:
            v[self.side_nodes[s]] = values[s]

    def eval_rhs(self, t):
        dim = self.dim
        self._set_sides(self.vort, self.dim_w, {s: np.zeros(self.dim_w) for s in self.sides})   # :115
        self._set_sides(self.vel, dim, self.wall_vel)                                            # :116-117
        m = self.mats
        b = m["Rw"] @ self.vort + m["Rwfs"] @ self.vort + m["Krhsfs"] @ self.vel                 # kle_solver.py:39-41
        self.velfs = self._lufs.solve(b)
        v = self.velfs.reshape(-1, dim)                                                          # :121-122
        for s in self.sides:                    # setTangentialValuesToVec, boundary_conditions.py:262-277
            for d in range(dim):
                if d != side_axis(s):
                    v[self.side_nodes[s], d] = self.wall_vel[s][d]
        self.vort = self.ops["Curl"] @ self.velfs                                                # :123
        self.vel = self._lu.solve(m["Rw"] @ self.vort + m["Krhs"] @ self.vel)                    # :125
        aux = self.ops["SrT"] @ self.vel
        aux *= 2.0 * self.mu
        aux += -1.0 * self.rho * vtensv(self.vel, dim)
        rhs = self.ops["DivSrT"] @ aux
        rhs *= 1.0 / self.rho
        return self.ops["Curl"] @ rhs

    def step(self, t, h, k0=None):
        s = len(RK5BS_B)
        k = [None] * s
        for i in range(s):
            if i == 0 and k0 is not None:
                k[0] = k0
                continue
            k[i] = self.eval_rhs(t + RK5BS_C[i] * h)
        unew = self.vort.copy()
        for j in range(s):
            if RK5BS_B[j] != 0.0:
                unew += h * RK5BS_B[j] * k[j]
        self.vort = unew
        return t + h, k[s - 1]
