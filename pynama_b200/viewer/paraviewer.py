"""Output hook of the post-step callback.  The reference writes HDF5 + XDMF
through PETSc viewers (src/viewer/paraviewer.py:14-78); file output is outside
the hot path, so this stand-in keeps the call surface and writes plain ``.npy``
snapshots only when a save directory is configured and ``-pynama_save 1`` is in
the options database."""
import os

import numpy as np

from ..petsc import COMM_WORLD, Options


class Paraviewer:
    comm = COMM_WORLD

    def __init__(self):
        self.saveDir = None
        self.enabled = bool(Options().getBool("pynama_save", False))

    def configure(self, dim, saveDir=None):
        self.dim = dim
        self.saveDir = saveDir
        if self.enabled and saveDir and not self.comm.rank:
            os.makedirs(saveDir, exist_ok=True)

    def saveMesh(self, coords, name="mesh"):
        if self.enabled and self.saveDir:
            np.save(os.path.join(self.saveDir, f"{name}.npy"), np.asarray(coords))

    def saveData(self, step, time, *vecs):
        if self.enabled and self.saveDir:
            for v in vecs:
                np.save(os.path.join(self.saveDir, f"{v.getName() or 'vec'}-{step:05d}.npy"), v.getArray())

    def writeXmf(self, name):
        pass
