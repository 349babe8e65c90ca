"""Free-slip operator builder on the GPU -- same surface as the reference's
``matrices.mat_fs`` (``MatFS``, ``Operators``; src/matrices/mat_fs.py).

``MatFS().setDomain(dom); .build()`` leaves ``.K .Krhs .Rw .Rd`` and
``.getOperators()`` with ``.Curl .SrT .DivSrT`` exactly as the reference does
(mat_fs.py:131-204), but the Python loop over cells (mat_fs.py:150-201) is one
batched element kernel plus one segmented reduction per matrix in
libpynama_b200.so.  Preallocation (mat_fs.py:21-113) has no equivalent: the
symbolic pattern is the device-built node graph.
"""
import logging

from ..petsc import COMM_WORLD, Vec


class MatFS:
    comm = COMM_WORLD
    bcType = "FS"
    # operator storage: True keeps only the structurally non-zero node blocks of
    # Curl/SrT/DivSrT (the exported CSR still shows the reference's explicit zeros)
    compactOperators = True

    def __init__(self):
        self.logger = logging.getLogger(f"[{self.comm.rank}]:MatClass")
        self.kle = list()
        self.operators = list()
        self.operator = None

    def setDomain(self, dom):
        self.dom = dom

    def assembleAll(self):
        for m in self.kle:
            m.assemble()

    def build(self, buildKLE=True, buildOperators=True):
        dm = self.dom.getDeviceMesh()
        if buildKLE:
            self.buildFS()
        if buildOperators:
            self.buildOperators()
        self._dm = dm

    def _adopt(self, mats):
        for name, m in mats.items():
            m.setName(name)
            setattr(self, name, m)
            self.kle.append(m)

    def buildFS(self, globNodesDirichlet=None):
        """K, Krhs, Rw (and the never-filled Rd) -- mat_fs.py:150-192"""
        self._adopt(self.dom.getDeviceMesh().assemble_kle("FS"))
        if not self.comm.rank:
            self.logger.info("KLE Matrices builded")

    def buildOperators(self):
        """Curl, SrT, DivSrT with the lumped-mass projection -- mat_fs.py:194-201, 238-271"""
        self.operator = Operators()
        self.operator.setDimensions(self.dom.getDimensions())
        self.operator.buildFrom(self.dom.getDeviceMesh(), self.compactOperators)
        if not self.comm.rank:
            self.logger.info("Operators Matrices builded")

    def getOperators(self):
        return self.operator

    def printMatsInfo(self):
        self.logger.info(" MATS INFO ")
        self.logger.info("Mat   | Memory Used [B]  | NZ Unneeded")
        for m in self.kle:
            self.logger.info(f" {self.formatMatInfo(m.getName(), m.getInfo())}")

    @staticmethod
    def formatMatInfo(name, info):
        return f"{name:{5}} | {info['memory']:{16}} | {info['nz_unneeded']:{10}}"


class Operators(MatFS):
    def setDimensions(self, dims):
        self.dim, self.dim_w, self.dim_s = dims

    def buildFrom(self, dmesh, compact=True):
        self.Curl, self.SrT, self.DivSrT, w = dmesh.assemble_ops(compact=compact)
        self.kle = [self.Curl, self.SrT, self.DivSrT]
        # the reference destroys its weight vectors after scaling (mat_fs.py:269-271) but
        # BaseProblemTest.OperatorsTests still reads weigCurl (base_problem.py:332): keep
        # the reciprocal lumped mass, repeated per vorticity component, alive
        self.lumpedMass = w
        wc = w.getArray().repeat(self.dim_w)
        self.weigCurl = self.Curl.createVecLeft()
        self.weigCurl.setArray(1.0 / wc)

    def assembleAll(self):
        pass
