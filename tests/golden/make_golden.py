#!/usr/bin/env python
"""Generate tests/golden/*.npz from the REFERENCE implementation.

Run in the build container only (needs /root/reference, which does not exist
on the GPU box):

    python tests/golden/make_golden.py

It imports the reference's own ``Spectral`` element class
(src/domain/elements/spectral.py) with a one-line ``mpi4py`` stub (the only
non-numpy import of that module, unused: element.py:3) and records

* the local node order (as reference coordinates of every node) and the six
  table sets for several ngl,
* K_e / Rw_e / Rd_e and SrT_e / DivSrT_e / Curl_e / W_e for the unit cell and
  for seeded non-affine (perturbed-corner) cells,
* for the large 3-D orders only a seeded sample of entries plus row sums.

Nothing here is used at run time by the product; the oracle is checked against
these files by tests/test_oracle_golden.py, the CUDA path by tests/test_gpu_*.
"""
import os
import sys
import tempfile

import numpy as np

REF = os.environ.get("PYNAMA_REFERENCE", "/root/reference")
HERE = os.path.dirname(os.path.abspath(__file__))


def _import_reference():
    stub = tempfile.mkdtemp(prefix="mpi4py_stub_")
    os.makedirs(os.path.join(stub, "mpi4py"))
    with open(os.path.join(stub, "mpi4py", "__init__.py"), "w") as f:
        f.write("MPI = None\n")
    sys.path.insert(0, stub)
    sys.path.insert(0, os.path.join(REF, "src"))
    from domain.elements.spectral import Spectral          # noqa: E402
    from domain.elements import utilities                    # noqa: E402
    return Spectral, utilities


# corner order of a DMPlex box cell (src/tests/test_dmplex_wrapper.py:23-27,105-115)
CORNERS = {
    2: np.array([[0, 0], [1, 0], [1, 1], [0, 1]], dtype=np.float64),
    3: np.array([[0, 0, 0], [0, 1, 0], [1, 1, 0], [1, 0, 0],
                 [0, 0, 1], [1, 0, 1], [1, 1, 1], [0, 1, 1]], dtype=np.float64),
}


def cells(dim):
    """unit cell, a stretched/shifted affine cell, two seeded non-affine cells"""
    rng = np.random.default_rng(20240229 + dim)
    out = [CORNERS[dim].copy()]
    scale = np.array([0.25, 0.5, 0.125][:dim])
    out.append(CORNERS[dim] * scale + np.array([1.0, -2.0, 0.5][:dim]))
    for _ in range(2):
        out.append(CORNERS[dim] * scale + 0.15 * scale * rng.uniform(-1, 1, CORNERS[dim].shape))
    return out


def tables(el):
    d = {}
    for name in ("H", "Hrs", "HRed", "HrsRed", "HOp", "HrsOp", "HCoo", "HrsCoo",
                 "HCooRed", "HrsCooRed", "HCooOp", "HrsCooOp"):
        d[name] = np.array(getattr(el, name))
    for name in ("gps", "gpsRed", "gpsOp"):
        d[name] = np.array([tuple(g) for g in getattr(el, name)])
    return d


def main():
    Spectral, util = _import_reference()
    out = {}
    # 1-D rules
    for n in range(1, 9):
        x, w = util.gaussPoints(n) if n > 1 else (np.zeros(1), 2 * np.ones(1))
        out[f"gauss{n}_x"], out[f"gauss{n}_w"] = x, w
    for n in range(2, 9):
        x, w = util.lobattoPoints(n)
        out[f"lobatto{n}_x"], out[f"lobatto{n}_w"] = x, w
    np.savez_compressed(os.path.join(HERE, "rules_1d.npz"), **out)

    rng = np.random.default_rng(7)
    for dim, ngls_full, ngls_sample in ((2, (2, 3, 4, 5), (8,)), (3, (2, 3, 4), (5, 6))):
        for ngl in ngls_full + ngls_sample:
            el = Spectral(ngl, dim)
            out = {"ngl": ngl, "dim": dim}
            full = ngl in ngls_full
            tb = tables(el)
            # reference coordinates of the nodes in local order
            out["node_rst"] = tb["gpsOp"][:, :dim]
            out["gps"], out["gpsRed"], out["gpsOp"] = tb["gps"], tb["gpsRed"], tb["gpsOp"]
            if full:
                out.update({k: v for k, v in tb.items()})
            else:
                for k in ("HrsCoo", "HrsCooRed", "HrsCooOp", "HCooOp"):
                    out[k] = tb[k]
            cl = cells(dim)
            out["cells"] = np.array(cl)
            nn = ngl ** dim
            dw, ds = (1, 3) if dim == 2 else (3, 6)
            shapes = {"K": (nn * dim, nn * dim), "Rw": (nn * dim, nn * dim_w(dim)),
                      "Rd": (nn * dim, nn), "SrT": (nn * ds, nn * dim),
                      "Div": (nn * dim, nn * ds), "Curl": (nn * dw, nn * dim)}
            samp = {k: (rng.integers(0, s[0], 4000), rng.integers(0, s[1], 4000))
                    for k, s in shapes.items()}
            for ci, X in enumerate(cl):
                K, Rw, Rd = el.getElemKLEMatrices(X.copy().ravel())
                SrT, Div, Curl, W = el.getElemKLEOperators(X.copy().ravel())
                mats = {"K": K, "Rw": Rw, "Rd": Rd, "SrT": SrT, "Div": Div, "Curl": Curl}
                out[f"c{ci}_W"] = W
                for k, m in mats.items():
                    if full:
                        out[f"c{ci}_{k}"] = m
                    else:
                        i, j = samp[k]
                        out[f"c{ci}_{k}_i"], out[f"c{ci}_{k}_j"] = i, j
                        out[f"c{ci}_{k}_v"] = m[i, j]
                        out[f"c{ci}_{k}_rowsum"] = m.sum(1)
                        out[f"c{ci}_{k}_colsum"] = m.sum(0)
            path = os.path.join(HERE, f"elem_{dim}d_ngl{ngl}.npz")
            np.savez_compressed(path, **out)
            print(path, os.path.getsize(path) // 1024, "kB")

    # local order for more ngl (3-D spectral order straight from the reference)
    out = {}
    for n in range(1, 8):
        out[f"order3d_{n}"] = np.array(Spectral.getSpectralOrder(n))
    np.savez_compressed(os.path.join(HERE, "spectral_order_3d.npz"), **out)


def dim_w(dim):
    return 1 if dim == 2 else 3


if __name__ == "__main__":
    main()
