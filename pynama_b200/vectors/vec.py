"""Field container of the hot path.  The reference's ``vectors/vec.py`` is an
unused 18-line stub around ``PETSc.Vec`` (src/vectors/vec.py:3-19); the real
container is ``PETSc.Vec`` itself, whose surface ``pynama_b200.petsc.Vec``
provides on the device."""
from ..petsc import Vec


class Vector(Vec):
    def __init__(self, size=None, name=""):
        super().__init__(size, name=name)
