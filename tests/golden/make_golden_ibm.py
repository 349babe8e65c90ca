#!/usr/bin/env python
"""Generate tests/golden/ibm_*.npz from the REFERENCE's immersed-boundary code.

Run in the build container only (needs /root/reference):

    python tests/golden/make_golden_ibm.py

The reference modules import petsc4py/mpi4py at the top; neither exists here and none of the
functions recorded below touches them, so both are replaced by empty stand-ins before the import.
Recorded, all computed by the reference's own functions:

* ``fourGrid``/``threeGrid``/``linear`` on a grid of arguments       (domain/immersed_body.py:398-424)
* ``Circle/Line/OpenBox.generateBody``                               (domain/immersed_body.py:293-396)
* ``ImmersedBody.updateBodyParameters`` for a moving body             (domain/immersed_body.py:268-284)
* for three Euler grids, per marker: ``computeClose`` (static and dynamic rule) and ``computeDirac``
  of ``cases/immersed_boundary.py:255-274,346-358`` (which call ``ImmersedBody.getDiracs``)

The only thing supplied from outside is what PETSc's DMPlex would hand back: the Euler node
coordinates (a uniform grid, lexicographic numbering, x fastest) and the marker coordinate
``local + centre`` (immersed_body.py:256-259).
"""
import os
import sys
import tempfile
import types

import numpy as np

REF = os.environ.get("PYNAMA_REFERENCE", "/root/reference")
HERE = os.path.dirname(os.path.abspath(__file__))


def _import_reference():
    stub = tempfile.mkdtemp(prefix="petsc_stub_")
    os.makedirs(os.path.join(stub, "petsc4py"))
    os.makedirs(os.path.join(stub, "mpi4py"))
    with open(os.path.join(stub, "petsc4py", "__init__.py"), "w") as f:
        f.write(
            "import sys, types\n"
            "def init(*a, **k): pass\n"
            "class _Meta(type):\n"
            "    def __getattr__(cls, name):\n"
            "        if name.startswith('__'): raise AttributeError(name)\n"
            "        return _Meta(name, (object,), {})\n"
            "class PETSc(metaclass=_Meta): pass\n"
            "m = types.ModuleType('petsc4py.PETSc')\n"
            "m.__getattr__ = lambda name: getattr(PETSc, name)\n"
            "sys.modules['petsc4py.PETSc'] = m\n")
    with open(os.path.join(stub, "mpi4py", "__init__.py"), "w") as f:
        f.write("MPI = None\n")
    sys.path.insert(0, stub)
    sys.path.insert(0, os.path.join(REF, "src"))
    from domain import immersed_body as ib                                   # noqa: E402
    from cases.immersed_boundary import ImmersedBoundaryStatic, ImmersedBoundaryDynamic  # noqa: E402
    return ib, ImmersedBoundaryStatic, ImmersedBoundaryDynamic


class _Body:
    """what the case asks of ``self.body``; the delta function is the reference object's"""

    def __init__(self, ib, ref_body, xy, dl):
        self._ib, self._ref, self._xy, self._dl = ib, ref_body, xy, dl

    def getNodeCoordinates(self, lag):
        return self._xy[lag]

    def getElementLong(self):
        return self._dl

    def getDiracs(self, dist, h):
        return self._ib.ImmersedBody.getDiracs(self._ref, dist, h)


def grid_nodes(nelem, lower, upper, ngl, jitter=0.0, seed=7):
    n = [ne * (ngl - 1) + 1 for ne in nelem]
    x = np.linspace(lower[0], upper[0], n[0])
    y = np.linspace(lower[1], upper[1], n[1])
    xy = np.stack(np.meshgrid(x, y, indexing="xy"), axis=-1).reshape(-1, 2)   # x fastest
    if jitter:
        h = (upper[0] - lower[0]) / (n[0] - 1)
        xy = xy + jitter * h * np.random.default_rng(seed).uniform(-1, 1, xy.shape)
    return xy


def rows_of(case_cls, ib, ref_body, lag_xy, euler_xy, h, dl):
    fake = types.SimpleNamespace(body=_Body(ib, ref_body, lag_xy, dl), nodeSeparation=h)
    ptr, nodes, w = [0], [], []
    for lag in range(len(lag_xy)):
        mask = case_cls.computeClose(fake, lag, euler_xy.copy())
        keep = np.nonzero(mask > 0)[0]
        d = np.array(case_cls.computeDirac(fake, lag, euler_xy[keep]), dtype=np.float64)
        nodes.append(keep)
        w.append(d)
        ptr.append(ptr[-1] + len(keep))
    return np.array(ptr), np.concatenate(nodes), np.concatenate(w)


def main():
    ib, Static, Dynamic = _import_reference()
    out = {}
    r = np.concatenate([np.linspace(0, 2.5, 251), [0.5, 1.0, 1.5, 2.0]])
    out["r"] = r
    for name in ("fourGrid", "threeGrid", "linear"):
        out[name] = np.array([getattr(ib, name)(float(x)) for x in r], dtype=np.float64)
    np.savez_compressed(os.path.join(HERE, "ibm_dirac.npz"), **out)

    out = {}
    for i, dh in enumerate((0.5, 0.06, 0.0371)):
        c = ib.Circle(vel=[0, 0], center=[0, 0], radius=0.5)
        xy, _, dl = c.generateBody(dh)
        out[f"circle{i}_dh"], out[f"circle{i}_xy"], out[f"circle{i}_dl"] = dh, np.array(xy), dl
        xy, _, dl = ib.Line(vel=[0, 0], center=[0, 0]).generateBody(dh)
        out[f"line{i}_xy"], out[f"line{i}_dl"] = np.array(xy), dl
        xy, _, dl = ib.OpenBox(vel=[0, 0], center=[0, 0]).generateBody(dh)
        out[f"box{i}_xy"], out[f"box{i}_dl"] = np.array(xy), dl
    mover = ib.Circle(vel=[0, 0], center=[0.25, -0.5], radius=0.5)
    mover.setVelRef(2.8)
    mover.setIsMoving()
    mover.updateVelocity = lambda: None           # would write a PETSc Vec
    ts = np.array([0.0, 0.1, 0.77, 1.9, 13.0])
    cen, vel = [], []
    for t in ts:
        mover.updateBodyParameters(float(t))
        cen.append(list(mover.getCenterBody()))
        vel.append(list(mover._ImmersedBody__vel))
    out["mover_t"], out["mover_center"], out["mover_vel"] = ts, np.array(cen), np.array(vel)
    out["mover_uref"], out["mover_start"] = 2.8, np.array([0.25, -0.5])
    np.savez_compressed(os.path.join(HERE, "ibm_bodies.npz"), **out)

    # (name, nelem, lower, upper, ngl, jitter, centre, case class)
    grids = [("coarse", [10, 10], [-5, -5], [5, 5], 3, 0.0, [0.0, 0.0], Static),
             ("static", [50, 50], [-3, -3], [3, 3], 3, 0.0, [0.0, 0.0], Static),
             ("jitter", [24, 24], [-2, -2], [2, 2], 2, 0.2, [0.13, -0.21], Static),
             ("dynamic", [40, 40], [-2, -2], [2, 2], 3, 0.0, [0.0, 0.171], Dynamic)]
    for name, nelem, lo, up, ngl, jit, centre, cls in grids:
        euler_xy = grid_nodes(nelem, lo, up, ngl, jit)
        h = (up[0] - lo[0]) / nelem[0] / (ngl - 1)
        c = ib.Circle(vel=[0, 0], center=centre, radius=0.5)
        loc, _, dl = c.generateBody(h)
        lag_xy = np.array(loc) + np.array(centre)
        ptr, nodes, w = rows_of(cls, ib, c, lag_xy, euler_xy, h, dl)
        np.savez_compressed(os.path.join(HERE, f"ibm_rows_{name}.npz"), nelem=np.array(nelem),
                            lower=np.array(lo, dtype=float), upper=np.array(up, dtype=float), ngl=ngl,
                            jitter=jit, centre=np.array(centre), h=h, dl=dl, lag_xy=lag_xy,
                            euler_xy=euler_xy if jit else np.zeros((0, 2)),
                            dynamic=int(cls is Dynamic), rowptr=ptr, nodes=nodes, weights=w)
        print(name, "markers", len(lag_xy), "entries", len(nodes), "max/row", np.diff(ptr).max())


if __name__ == "__main__":
    main()
