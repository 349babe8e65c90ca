"""No-slip operator builder on the GPU -- same surface as the reference's
``matrices.mat_ns.MatNS`` (src/matrices/mat_ns.py): adds ``.Kfs .Krhsfs .Rwfs
.Rdfs`` and fills ``.Rd`` (mat_ns.py:47-166)."""
from .mat_fs import MatFS


class MatNS(MatFS):
    bcType = "NS"

    def build(self, buildKLE=True, buildOperators=True):
        if buildKLE:
            self.buildNS()
        if buildOperators:
            self.buildOperators()

    def buildNS(self):
        self._adopt(self.dom.getDeviceMesh().assemble_kle("NS"))
        if not self.comm.rank:
            self.logger.info("KLE Matrices builded")
