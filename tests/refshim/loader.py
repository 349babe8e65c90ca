"""Imports the reference's own modules (unmodified, from $PYNAMA_REFERENCE, default /root/reference) over the
numpy stand-in for petsc4py (petsc_standin.py) and the box-mesh stand-in for DMPlex (dm_standin.py).

TEST INFRASTRUCTURE; build container only (the GPU box has no copy of the reference).

    with reference_modules() as ref:
        fem = ref.base_problem.BaseProblem(config, case="taylor-green3d")
        fem.setUp(); fem.setUpSolver()

Inside the block `petsc4py`, `mpi4py` and the reference's top-level packages (domain, matrices, solver,
boundaries, cases, functions, common, viewer) are in sys.modules; on exit sys.modules and sys.path are restored,
so the product's tests that alias the same names (test_reference_dropin.py) are not disturbed.
The reference's Paraviewer (HDF5 output through PETSc.Viewer: out of the path's scope) is replaced by a viewer that
writes nothing.
"""
import contextlib
import importlib
import os
import sys
import types

REF = os.environ.get("PYNAMA_REFERENCE", "/root/reference")
SRC = os.path.join(REF, "src")
REF_PACKAGES = ("domain", "matrices", "solver", "boundaries", "cases", "functions", "common", "viewer", "vectors")


def available():
    return os.path.isfile(os.path.join(SRC, "cases", "base_problem.py"))


class _NullParaviewer:
    def __init__(self):
        self.saved = []

    def configure(self, dim, saveDir=None):
        self.dim = dim

    def saveMesh(self, coords, name="mesh"):
        pass

    def saveData(self, step, time, *vecs):
        self.saved.append((step, time))

    def writeXmf(self, name):
        pass


@contextlib.contextmanager
def reference_modules(options=None):
    from . import dm_standin, petsc_standin
    saved_modules = dict(sys.modules)
    saved_path = list(sys.path)
    saved_opts = dict(petsc_standin.Options._db)
    for k in list(sys.modules):
        if k.split(".")[0] in REF_PACKAGES + ("petsc4py", "mpi4py"):
            del sys.modules[k]
    try:
        petsc_standin.Options._db.clear()
        petsc_standin.Options._db.update(options or {})
        p4 = types.ModuleType("petsc4py")
        p4.PETSc = petsc_standin
        p4.init = lambda *a, **k: None
        sys.modules["petsc4py"] = p4
        sys.modules["petsc4py.PETSc"] = petsc_standin
        m4 = types.ModuleType("mpi4py")
        m4.MPI = types.SimpleNamespace(COMM_WORLD=petsc_standin.COMM_WORLD)
        sys.modules["mpi4py"] = m4
        vp = types.ModuleType("viewer.paraviewer")
        vp.Paraviewer = _NullParaviewer
        sys.path.insert(0, SRC)
        importlib.import_module("viewer")
        sys.modules["viewer.paraviewer"] = vp
        dmplex = importlib.import_module("domain.dmplex")
        domain = importlib.import_module("domain.domain")
        domain.BoxDom = dmplex.BoxDom = dm_standin.make_box_dom(dmplex, petsc_standin)
        ns = types.SimpleNamespace(
            PETSc=petsc_standin, dmplex=dmplex, domain=domain,
            spectral=importlib.import_module("domain.elements.spectral"),
            boundary_conditions=importlib.import_module("boundaries.boundary_conditions"),
            mat_fs=importlib.import_module("matrices.mat_fs"),
            mat_ns=importlib.import_module("matrices.mat_ns"),
            kle_solver=importlib.import_module("solver.kle_solver"),
            ts_solver=importlib.import_module("solver.ts_solver"),
            base_problem=importlib.import_module("cases.base_problem"),
            src=SRC)
        yield ns
    finally:
        for k in list(sys.modules):
            if k not in saved_modules:
                del sys.modules[k]
        sys.modules.update(saved_modules)
        sys.path[:] = saved_path
        petsc_standin.Options._db.clear()
        petsc_standin.Options._db.update(saved_opts)


def load_case(name):
    import yaml
    with open(os.path.join(SRC, "cases", f"{name}.yaml")) as f:
        return yaml.load(f, Loader=yaml.Loader)


def run_reference_unittests(ref, filename):
    """Runs one of the reference's own unittest files (src/tests/<filename>) inside an open reference_modules()
    block; the file is loaded by path under a private module name (the repo has a `tests` package of its own)
    with the working directory at the reference root, because the files open 'src/cases/*.yaml' relatively.
    Returns {test id: 'ok' | 'fail: ...' | 'error: ...' | 'skip'}."""
    import importlib.util
    import io
    import unittest
    path = os.path.join(SRC, "tests", filename)
    name = "reference_" + filename[:-3]
    spec = importlib.util.spec_from_file_location(name, path)
    mod = importlib.util.module_from_spec(spec)
    cwd = os.getcwd()
    os.chdir(REF)
    import numpy
    removed = [a for a in ("mat",) if a not in numpy.__dict__]       # the reference's tests predate NumPy 2.0
    try:
        if "mat" in removed:
            numpy.mat = numpy.asmatrix
        spec.loader.exec_module(mod)
        suite = unittest.defaultTestLoader.loadTestsFromModule(mod)
        out = {}

        class R(unittest.TextTestResult):
            def addSuccess(self, test):
                super().addSuccess(test)
                out[test.id().split(".", 1)[1]] = "ok"

            def addFailure(self, test, err):
                super().addFailure(test, err)
                out[test.id().split(".", 1)[1]] = "fail: " + str(err[1])[:200]

            def addError(self, test, err):
                super().addError(test, err)
                out[test.id().split(".", 1)[1]] = "error: " + f"{err[0].__name__}: {err[1]}"[:200]

            def addSkip(self, test, reason):
                super().addSkip(test, reason)
                out[test.id().split(".", 1)[1]] = "skip"
        with contextlib.redirect_stdout(io.StringIO()):
            unittest.TextTestRunner(stream=io.StringIO(), resultclass=R, verbosity=0).run(suite)
        return out
    finally:
        for a in removed:
            delattr(numpy, a)
        os.chdir(cwd)
