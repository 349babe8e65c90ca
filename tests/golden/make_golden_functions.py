#!/usr/bin/env python
"""Generate tests/golden/functions.npz from the REFERENCE's analytic field libraries
(src/functions/taylor_green.py, taylor_green3d.py, taylor_green_3d.py: plain numpy, imported as
they are).  Run in the build container only (needs /root/reference):

    python tests/golden/make_golden_functions.py

Recorded: alpha(nu, t) and every array-valued public function of each library at seeded
coordinates, plus the scalar *_test helpers of the 2-D library.  tests/test_boundaries_host.py
compares pynama_b200/functions/* with this file (initial conditions, function boundaries and the
exact fields of the KLE / operator checks all come from these libraries: base_problem.py:181-222,
boundary.py:141-146).
"""
import importlib
import inspect
import os
import sys

import numpy as np

REF = os.environ.get("PYNAMA_REFERENCE", "/root/reference")
HERE = os.path.dirname(os.path.abspath(__file__))


def main():
    sys.path.insert(0, os.path.join(REF, "src"))
    rng = np.random.default_rng(20240301)
    out = {}
    cases = [(0.02, 0.0), (0.02, 0.37), (1.0 / 50.0, 1.25), (100.0, 0.0)]
    out["nu_t"] = np.array(cases)
    for lib_name, dim in (("taylor_green", 2), ("taylor_green3d", 3), ("taylor_green_3d", 3)):
        lib = importlib.import_module(f"functions.{lib_name}")
        coords = rng.uniform(-0.3, 1.4, (23, dim))
        out[f"{lib_name}__coords"] = coords
        out[f"{lib_name}__alpha"] = np.array([lib.alpha(nu, t) for nu, t in cases])
        names = []
        for name, fn in inspect.getmembers(lib, inspect.isfunction):
            if fn.__module__ != lib.__name__ or name == "alpha" or name.endswith("_test"):
                continue
            params = list(inspect.signature(fn).parameters)
            if len(params) != 2:
                continue
            vals = []
            try:
                if fn(coords.copy(), 1.0) is None:          # a stub (taylor_green.convective/diffusive: "pass")
                    continue
                for nu, t in cases:
                    vals.append(np.asarray(fn(coords.copy(), lib.alpha(nu, t)), dtype=np.float64).ravel())
            except Exception as exc:            # a helper with another calling convention
                print(f"skip {lib_name}.{name}: {exc}")
                continue
            out[f"{lib_name}__{name}"] = np.stack(vals)
            names.append(name)
        out[f"{lib_name}__names"] = np.array(names)
        if hasattr(lib, "velocity_test"):
            out[f"{lib_name}__velocity_test"] = np.array(
                [[lib.velocity_test(c, t, nu) for c in coords] for nu, t in cases], dtype=np.float64)
        print(lib_name, names)
    np.savez_compressed(os.path.join(HERE, "functions.npz"), **out)


if __name__ == "__main__":
    main()
