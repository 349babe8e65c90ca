"""CPU study (not a test, not product code): how many CG iterations does the KLE system of the
bench workload (3-D Taylor-Green, free-slip box, ngl = 4) need under preconditioners that would
be cheap on the GPU next to one SpMV(K)?  DESIGN section 8 quotes the table this prints.

    python tests/study_precond.py [nelem] [ngl]

Uses the oracle's assembled K (scipy) -- checker code, which is why this lives under tests/.
Preconditioners:
  bjacobi      node-block Jacobi (what the product uses today)
  schwarz      additive Schwarz over elements: sum_e R_e^T (R_e K R_e^T)^-1 R_e; on a uniform box
               the element blocks fall into at most 27 distinct ones (interior, faces, edges,
               corners), so the apply is a gather, one batched dense product per class, a scatter-add
  schwarz-w    the same, symmetrically weighted by 1/sqrt(multiplicity)
  +coarse      additive Q1 coarse-grid correction on the element corners (Galerkin P^T K P)
  +coarse/m    the same on macro cells of m x m x m elements
  +coarse(ngl2)  the coarse operator assembled on the ngl = 2 mesh instead of the Galerkin product
  +coarse~chebK  the Galerkin coarse problem solved by K Chebyshev steps with Jacobi (a fixed polynomial:
               still a linear symmetric preconditioner) instead of exactly
"its warm": the next solve of the time loop (boundary values one stage later) started from the
previous solution.
"""
import os
import sys
import time

import numpy as np
import scipy.sparse as sp
import scipy.sparse.linalg as spl

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

from oracle import timestep  # noqa: E402
from pynama_b200.domain.box_mesh import BoxMesh  # noqa: E402


def node_block_jacobi(K, d):
    B = sp.bsr_matrix(K, blocksize=(d, d))
    B.sort_indices()
    nb = K.shape[0] // d
    blk = np.zeros((nb, d, d))
    rows = np.repeat(np.arange(nb), np.diff(B.indptr))
    sel = B.indices == rows
    blk[rows[sel]] = B.data[sel]
    inv = np.linalg.inv(blk)
    return lambda r: np.einsum("nij,nj->ni", inv, r.reshape(nb, d)).ravel()


def element_schwarz(K, elem_nodes, d, weighted):
    nel, nn = elem_nodes.shape
    dofs = (elem_nodes[:, :, None] * d + np.arange(d)[None, None, :]).reshape(nel, nn * d)
    Kc = K.tocsr()
    inv = np.empty((nel, nn * d, nn * d))
    for e in range(nel):
        inv[e] = np.linalg.inv(Kc[dofs[e]][:, dofs[e]].toarray())
    mult = np.bincount(dofs.ravel(), minlength=K.shape[0]).astype(float)
    w = 1.0 / np.sqrt(mult) if weighted else np.ones_like(mult)
    wl = w[dofs]

    def apply(r):
        loc = r[dofs] * wl
        z = np.einsum("eij,ej->ei", inv, loc) * wl
        out = np.zeros_like(r)
        np.add.at(out, dofs.ravel(), z.ravel())
        return out
    # distinct blocks
    key = np.round(inv / np.abs(inv).max(), 9).reshape(nel, -1)
    nclass = len({k.tobytes() for k in key})
    return apply, nclass


def q1_prolongation(mesh, d, ncell):
    """trilinear interpolation from a regular ncell^3 grid of (macro) cells over the unit box to
    the GLL nodes, per component"""
    X = mesh.coords
    g = ncell + 1
    t = np.clip(X * ncell, 0, ncell)
    i0 = np.minimum(np.floor(t + 1e-12).astype(int), ncell - 1)
    f = t - i0
    rows, cols, vals = [], [], []
    for cx in (0, 1):
        for cy in (0, 1):
            for cz in (0, 1):
                c = np.array([cx, cy, cz])
                w_ = np.prod(np.where(c == 1, f, 1.0 - f), axis=1)
                idx = ((i0[:, 0] + cx) * g + (i0[:, 1] + cy)) * g + (i0[:, 2] + cz)
                sel = w_ > 1e-12
                rows.append(np.flatnonzero(sel)), cols.append(idx[sel]), vals.append(w_[sel])
    P = sp.csr_matrix((np.concatenate(vals), (np.concatenate(rows), np.concatenate(cols))), shape=(len(X), g ** 3))
    assert np.allclose(np.asarray(P.sum(1)).ravel(), 1.0, atol=1e-10)
    return sp.kron(P, sp.identity(d)).tocsr()


def coarse_correction(K, P, bdofs):
    # Dirichlet rows of K are identity rows: keep the coarse space off them
    keep = np.ones(K.shape[0])
    keep[bdofs] = 0.0
    Pk = sp.diags(keep) @ P
    nz = np.flatnonzero(np.asarray(abs(Pk).sum(0)).ravel() > 0)
    Pk = Pk[:, nz].tocsr()
    Ac = (Pk.T @ K @ Pk).tocsc()
    lu = spl.splu(Ac)
    return lambda r: Pk @ lu.solve(Pk.T @ r), Ac.shape[0]


def coarse_rediscretised(mesh, prob, P, d):
    """the same KLE operator assembled on the ngl = 2 mesh of the same corners (what the existing
    assembly kernels could build on the device) instead of the Galerkin product; P maps its
    lexicographic vertex grid to the fine nodes"""
    nel = round(mesh.elem_nodes.shape[0] ** (1 / 3))
    cm = BoxMesh([nel] * 3, [0, 0, 0], [1, 1, 1], 2)
    cp = timestep.OracleProblem(3, 2, cm.corner_xyz, cm.elem_nodes, cm.coords, cm.getBordersNodes(),
                                rho=prob.rho, mu=prob.mu, solver="direct")
    # P's columns are numbered (ix*g + iy)*g + iz on the unit box: match the coarse mesh's nodes by coordinates
    g = nel + 1
    ijk = np.rint(cm.coords * nel).astype(int)
    col_of_node = (ijk[:, 0] * g + ijk[:, 1]) * g + ijk[:, 2]
    # P was built with columns (vertex, component) in grid order; reorder so column = coarse node*d + c
    order = (col_of_node[:, None] * d + np.arange(d)[None, :]).ravel()
    Pn = P.tocsc()[:, order].tocsr()
    lu = spl.splu(cp.K.tocsc())
    free = np.ones(cp.K.shape[0])
    free[cp.bdofs] = 0.0                        # constrained coarse dofs carry no correction
    keep = np.ones(P.shape[0])
    keep[prob.bdofs] = 0.0
    return lambda r: keep * (Pn @ (free * lu.solve(free * (Pn.T @ (keep * r)))))


def chebyshev_inverse(A, dinv_apply, k, lmax, lmin_frac=0.1):
    """k steps of Chebyshev iteration on the Jacobi-preconditioned A, zero start: a fixed
    polynomial in A, hence a linear symmetric operator usable inside CG"""
    lmin = lmin_frac * lmax
    theta, delta = 0.5 * (lmax + lmin), 0.5 * (lmax - lmin)

    def apply(b):
        x = np.zeros_like(b)
        r = b.copy()
        sigma = theta / delta
        rho = 1.0 / sigma
        dvec = dinv_apply(r) / theta
        for _ in range(k):
            x = x + dvec
            r = r - A @ dvec
            rho_new = 1.0 / (2.0 * sigma - rho)
            dvec = rho_new * rho * dvec + 2.0 * rho_new / delta * dinv_apply(r)
            rho = rho_new
        return x
    return apply


def run_cg(K, b, M, rtol, x0=None):
    its = [0]

    def cb(_):
        its[0] += 1
    n = K.shape[0]
    t0 = time.perf_counter()
    x, info = spl.cg(K, b, x0=x0, rtol=rtol, atol=0.0, maxiter=5000, M=spl.LinearOperator((n, n), matvec=M),
                     callback=cb)
    return its[0], info, np.linalg.norm(b - K @ x) / np.linalg.norm(b), time.perf_counter() - t0


def main():
    nel = int(sys.argv[1]) if len(sys.argv) > 1 else 6
    ngl = int(sys.argv[2]) if len(sys.argv) > 2 else 4
    rtol = 1e-10
    mesh = BoxMesh([nel] * 3, [0, 0, 0], [1, 1, 1], ngl)
    t0 = time.perf_counter()
    lean = os.environ.get("STUDY_LEAN") == "1"      # large meshes: no LU anywhere, three rows, cold solves only
    prob = timestep.OracleProblem(3, ngl, mesh.corner_xyz, mesh.elem_nodes, mesh.coords, mesh.getBordersNodes(),
                                  rho=0.5, mu=0.01, solver="cg" if lean else "direct", pc="jacobi")
    K = prob.K
    d = 3
    # the right-hand side of the first solve of a run (kle_solver.py:33-37), zero initial guess
    a0 = timestep.tg_alpha(prob.nu, 0.0, 3)
    Xb = prob.coords[prob.bnd]
    prob.vort[prob.bwdofs] = timestep.tg_vorticity(Xb, a0)
    prob.vel[prob.bdofs] = timestep.tg_velocity(Xb, a0)
    b = prob.mats["Rw"] @ prob.vort + prob.mats["Krhs"] @ prob.vel
    print(f"# {nel}^3 elements, ngl={ngl}: {K.shape[0]} dofs, nnz {K.nnz}, set-up {time.perf_counter() - t0:.1f} s")
    bj = node_block_jacobi(K, d)
    rows = [("bjacobi", bj)]
    if lean:
        swb, ncls = element_schwarz(K, mesh.elem_nodes, d, True)
        print(f"# schwarz-w: {ncls} distinct element blocks", flush=True)
        rows.append(("schwarz-w", swb))
        keepv = np.ones(K.shape[0])
        keepv[prob.bdofs] = 0.0
        Pk = (sp.diags(keepv) @ q1_prolongation(mesh, d, nel)).tocsr()
        nzc = np.flatnonzero(np.asarray(abs(Pk).sum(0)).ravel() > 0)
        Pk = Pk[:, nzc].tocsr()
        Ac = (Pk.T @ K @ Pk).tocsr()
        dinv = 1.0 / Ac.diagonal()
        lmax = 1.1 * spl.eigsh(spl.LinearOperator(Ac.shape, matvec=lambda v: dinv * (Ac @ v)), k=1, which="LM",
                               return_eigenvectors=False, tol=1e-3)[0]
        ch = chebyshev_inverse(Ac, lambda v: dinv * v, 5, lmax)
        rows.append(("schwarz-w+coarse~cheb5", lambda r: swb(r) + Pk @ ch(Pk.T @ r)))
        print(f"{'preconditioner':26s} {'its cold':>9s} {'true rel. residual':>20s} {'s':>7s}", flush=True)
        for name, M in rows:
            its, info, res, secs = run_cg(K, b, M, rtol)
            print(f"{name:26s} {its:9d} {res:20.2e} {secs:7.1f}" + ("" if info == 0 else f"  (info {info})"), flush=True)
        return
    for weighted in (False, True):
        sw, ncls = element_schwarz(K, mesh.elem_nodes, d, weighted)
        name = "schwarz-w" if weighted else "schwarz"
        print(f"# {name}: {ncls} distinct element blocks of {mesh.elem_nodes.shape[1] * d}^2")
        rows.append((name, sw))
    sw, swb = rows[1][1], rows[2][1]
    cc, nc = coarse_correction(K, q1_prolongation(mesh, d, nel), prob.bdofs)
    print(f"# coarse space on the element corners: {nc} dofs")
    rows.append(("bjacobi+coarse", lambda r: bj(r) + cc(r)))
    rows.append(("schwarz+coarse", lambda r: sw(r) + cc(r)))
    rows.append(("schwarz-w+coarse", lambda r: swb(r) + cc(r)))
    P1 = q1_prolongation(mesh, d, nel)
    cr = coarse_rediscretised(mesh, prob, P1, d)
    rows.append(("schwarz-w+coarse(ngl2)", lambda r: swb(r) + cr(r)))
    # the Galerkin coarse problem solved only approximately: k Chebyshev steps with node-block Jacobi
    keepv = np.ones(K.shape[0])
    keepv[prob.bdofs] = 0.0
    Pk = (sp.diags(keepv) @ P1).tocsr()
    nzc = np.flatnonzero(np.asarray(abs(Pk).sum(0)).ravel() > 0)
    Pk = Pk[:, nzc].tocsr()
    Ac = (Pk.T @ K @ Pk).tocsr()
    dinv = 1.0 / Ac.diagonal()
    lmax = 1.1 * spl.eigsh(spl.LinearOperator(Ac.shape, matvec=lambda v: dinv * (Ac @ v)), k=1, which="LM",
                           return_eigenvectors=False, tol=1e-3)[0]
    for kc in (5, 10, 20):
        ch = chebyshev_inverse(Ac, lambda v: dinv * v, kc, lmax)
        rows.append((f"schwarz-w+coarse~cheb{kc}", lambda r, ch=ch: swb(r) + Pk @ ch(Pk.T @ r)))
    for m in (2, 4):
        if nel % m == 0 and nel // m >= 2:
            ccm, ncm = coarse_correction(K, q1_prolongation(mesh, d, nel // m), prob.bdofs)
            print(f"# coarse space on {m}x{m}x{m}-element macro cells: {ncm} dofs")
            rows.append((f"schwarz-w+coarse/{m}", lambda r, ccm=ccm: swb(r) + ccm(r)))
    a1 = timestep.tg_alpha(prob.nu, 1e-3, 3)
    prob.vort[prob.bwdofs] = timestep.tg_vorticity(Xb, a1)
    prob.vel[prob.bdofs] = timestep.tg_velocity(Xb, a1)
    b1 = prob.mats["Rw"] @ prob.vort + prob.mats["Krhs"] @ prob.vel
    xprev = spl.splu(K.tocsc()).solve(b)
    print(f"{'preconditioner':26s} {'its cold':>9s} {'true rel. residual':>20s} {'its warm':>9s} {'s':>7s}")
    for name, M in rows:
        its, info, res, secs = run_cg(K, b, M, rtol)
        itw, infow, _, secw = run_cg(K, b1, M, rtol, x0=xprev)
        print(f"{name:26s} {its:9d} {res:20.2e} {itw:9d} {secs + secw:7.1f}" +
              ("" if info == 0 and infow == 0 else f"  (info {info} {infow})"))


if __name__ == "__main__":
    main()
