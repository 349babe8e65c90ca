"""CPU restatement of the reference's 2-D immersed-boundary path.  TEST INFRASTRUCTURE ONLY:
imported by tests/ (and nothing in the product); the product path is pynama_b200/csrc/ibm.cu.

What it restates (plain numpy, same arithmetic order as the reference's Python loops):

* the discrete delta functions            src/domain/immersed_body.py:398-424
* the Lagrangian markers of the bodies    src/domain/immersed_body.py:293-396
  (Line, OpenBox, Circle .generateBody; node = local coordinate + body centre, :256-259)
* the marker -> Euler-node search          src/cases/immersed_boundary.py:210-229 (collectNodes),
                                           :263-274 (computeClose, static), :346-358 (dynamic)
* the interpolation matrix H, the spreading matrix S and A = H S
                                           src/cases/immersed_boundary.py:162-208
* the velocity correction                  src/cases/immersed_boundary.py:162-168
* the prescribed motion of a moving body   src/domain/immersed_body.py:268-284

Pinned against the reference itself: tests/golden/make_golden_ibm.py imports the reference's
immersed_body module and ImmersedBoundaryStatic/Dynamic methods (petsc4py/mpi4py stubbed, they are
not touched by these functions) and stores their outputs in tests/golden/ibm_*.npz;
tests/test_oracle_ibm.py compares this file with those fixtures.
"""
from math import ceil, cos, pi, sin, sqrt

import numpy as np


# ---- discrete delta functions (immersed_body.py:398-424) -------------------------------------
def threeGrid(r):
    if r <= 0.5:
        return (1 + sqrt(-3 * r ** 2 + 1)) / 3
    if r <= 1.5:
        return (5 - 3 * r - sqrt(-3 * (1 - r) ** 2 + 1)) / 6
    return 0


def linear(r):
    return (1 - r) if r < 1 else 0


def fourGrid(r):
    if r <= 1:
        return (3 - 2 * r + sqrt(1 + 4 * r - 4 * r ** 2)) / 8
    if r <= 2:
        return (5 - 2 * r - sqrt(-7 + 12 * r - 4 * r ** 2)) / 8
    return 0


DIRACS = {"four": fourGrid, "three": threeGrid, "linear": linear}


def dirac_weight(dist, h, kind="four"):
    """immersed_body.py:261-266: product over the axes of phi(|r|/h)/h, in this order"""
    phi = DIRACS[kind]
    d = 1
    for r in dist:
        d *= phi(abs(r) / h)
        d /= h
    return d


# ---- bodies (immersed_body.py:293-396): local marker coordinates and the arc length ---------
def body_line(dl, longitud=2):
    div = ceil(longitud / dl)
    x = np.linspace(0, longitud, div)
    return np.stack([x, np.zeros(div)], axis=1), dl


def body_box(dl, longitud=1):
    div = ceil(sqrt(2) / dl)
    corners = [(0, longitud), (-longitud, 0), (0, -longitud), (longitud, 0), (0, longitud)]
    xs, ys = [], []
    for (x0, y0), (x1, y1) in zip(corners[:-1], corners[1:]):
        xs.append(np.linspace(x0, x1, div, endpoint=False))
        ys.append(np.linspace(y0, y1, div, endpoint=False))
    return np.stack([np.concatenate(xs), np.concatenate(ys)], axis=1), dl


def body_circle(dh, radius):
    total = 2 * pi * radius
    points = ceil(total / dh)
    ang = np.linspace(0, 2 * pi, points, endpoint=False) + pi / 1000
    xy = np.stack([radius * np.cos(ang), radius * np.sin(ang)], axis=1)
    return xy, 2 * pi * radius / points


def make_body(cfg, h):
    """one entry of the yaml's ``bodies`` list -> (marker coordinates incl. centre, dl)"""
    kind = cfg["type"]
    if kind == "circle":
        xy, dl = body_circle(h, cfg["radius"])
    elif kind == "line":
        xy, dl = body_line(h)
    elif kind == "box":
        xy, dl = body_box(h)
    else:
        raise Exception("not defined")
    return xy + np.asarray(cfg["center"], dtype=np.float64), dl


def moving_body_state(t, uref, start_center):
    """immersed_body.py:268-284: centre and velocity of a 'dynamic' body at time t"""
    f, A = 5, 0.3
    Te = f / uref
    center = [0 + start_center[0], A * sin(2 * pi * t / Te) + start_center[1]]
    vel = [0, 2 * pi * A * cos(2 * pi * t / Te) / Te]
    return center, vel


def box_body_velocity(xy, uref):
    """OpenBox.updateVelocity (immersed_body.py:347-357): the lid of the rotated cavity"""
    v = np.zeros_like(xy)
    lid = (xy[:, 0] >= 0) & (xy[:, 1] >= 0)
    v[lid, 0] = uref / sqrt(2)
    v[lid, 1] = -uref / sqrt(2)
    return v


# ---- marker -> Euler node search and the matrices --------------------------------------------
def affected_cells(centroids, xSide, ySide=None, center=(0, 0)):
    """immersed_boundary.py:231-244: cells whose centroid lies inside the box around ``center``"""
    if not ySide:
        ySide = xSide
    d = np.asarray(centroids)[:, :2] - np.asarray(center, dtype=np.float64)
    return np.nonzero((np.abs(d[:, 0]) < xSide) & (np.abs(d[:, 1]) < ySide))[0]


def close_mask(lag_xy, euler_xy, dl, dynamic=False):
    """computeClose (immersed_boundary.py:263-274, :346-358): 1 where the Euler node is kept"""
    out = np.zeros(len(euler_xy), dtype=bool)
    for i, c in enumerate(euler_xy):
        dist = (c - lag_xy) / dl
        if dynamic:
            out[i] = abs(dist[0]) < 2 and abs(dist[1] * dl) < (1 + 2 * dl)
        else:
            out[i] = abs(dist[0]) < 3 and abs(dist[1]) < 3
    return out


def build_rows(lag_xy, euler_nodes, euler_xy, h, dl, kind="four", dynamic=False):
    """collectNodes + computeDirac (immersed_boundary.py:210-229, :255-262): for every marker the
    kept Euler nodes (ascending) and the unscaled delta weights, explicit zeros included"""
    euler_nodes = np.asarray(euler_nodes)
    order = np.argsort(euler_nodes, kind="stable")
    euler_nodes, euler_xy = euler_nodes[order], np.asarray(euler_xy)[order]
    rows = []
    for lp in lag_xy:
        keep = close_mask(lp, euler_xy, dl, dynamic)
        if not keep.any():
            raise Exception("Lag Node without Euler")
        w = np.array([dirac_weight(c - lp, h, kind) for c in euler_xy[keep]])
        rows.append((euler_nodes[keep], w))
    return rows


def dense_H(rows, n_euler_nodes, dim=2):
    """the unscaled interpolation matrix of buildIBMMatrix (:170-183), dense [nlag*dim, N*dim]"""
    D = np.zeros((len(rows) * dim, n_euler_nodes * dim))
    for l, (nodes, w) in enumerate(rows):
        for d in range(dim):
            D[l * dim + d, nodes * dim + d] = w
    return D


def ibm_matrices(rows, n_euler_nodes, h, dl, dim=2):
    """H, S, A as the reference scales them (:185-190): S = (dl h) D^T, H = h^2 D, A = H S"""
    D = dense_H(rows, n_euler_nodes, dim)
    S = D.T * (dl * h)
    H = D * (h ** 2)
    return H, S, H @ S


def velocity_correction(H, S, A, vel, body_vel):
    """computeVelocityCorrection (:162-168).  Returns (vel + correction, correction, flux)."""
    rhs = H @ vel - body_vel
    q = np.linalg.solve(A, -rhs)
    corr = S @ q
    return vel + corr, corr, q


def affected_nodes(rows):
    """nodes with a positive weight (buildIBMMatrix's return value, :181,:193)"""
    out = set()
    for nodes, w in rows:
        out |= set(nodes[w > 0].tolist())
    return sorted(out)
