"""Stand-in for the reference's `BoxDom` (src/domain/dmplex.py:378-386): the mesh-derived tables that PETSc's
DMPlex would hand to the reference's Domain, from oracle/boxmesh.py.

TEST INFRASTRUCTURE.  `make_box_dom(dmplex_module)` returns a SUBCLASS of the reference's own `DMPlexDom`:
the methods written in plain Python there (getVelocityIndex, getVorticityIndex, getNodesCoordinates,
getBordersNodes, getBordersNames, applyValuesToVec, applyFunctionVecToVec ...) and the reference's
`IndicesManager` are inherited and run unmodified; only the methods that call into PETSc's DMPlex /
PetscSection are replaced by look-ups in the box-mesh tables:

    create, setFemIndexing, getTotalNodes, getTotalElements, computeFullCoordinates (the reference's own
    formula HCooOp @ corners, dmplex.py:73-88, written into a stand-in Vec), getCellCornersCoords,
    createGlobalVec, setLabelToBorders, getBorderNodes, getGlobalNodesFromCell, getGlobalNodesFromEntities, getAllNodes,
    getNodesRange, getConnectivityNodes, getEdgesWidth, getBoundingBox

What is therefore the stand-in's and NOT the reference's: the global node numbering (lexicographic, the
product's; PETSc's own for ngl >= 3 is not in the reference tree), the cell order (x fastest) and the
corner order of a cell (DMPlex closure order as pinned by src/tests/test_dmplex_wrapper.py:23-27,105-115).
"""
import numpy as np

from oracle import boxmesh


def make_box_dom(dmplex, PETSc):
    class StandInBoxDom(dmplex.DMPlexDom):
        def __init__(self):
            super().__init__()
            self._m = None

        # ---- construction --------------------------------------------------------------------------
        def create(self, data):
            self._cfg = {k: list(data[k]) for k in ("nelem", "lower", "upper")}
            # not a key of the reference's case files: interior vertices displaced by perturb * h (seeded), so that
            # the cells are non-affine and no two element matrices are equal
            self._perturb = float(data.get("perturb", 0.0))
            self.dim = len(self._cfg["nelem"])
            self.dim_w = 1 if self.dim == 2 else 3
            self.dim_s = 3 if self.dim == 2 else 6
            self.namingConvention = (["down", "right", "up", "left"] if self.dim == 2 else
                                     ["back", "front", "down", "up", "right", "left"])      # dmplex.py:26-29

        def getDimension(self):
            return self.dim

        def setFemIndexing(self, ngl):
            self.indicesManager = dmplex.IndicesManager(self.dim, ngl, self.comm)
            c = self._cfg
            self._m = boxmesh.box_mesh(c["nelem"], c["lower"], c["upper"], ngl, perturb=self._perturb)
            self.cellStart, self.cellEnd = 0, self._m.nel

        def getTotalNodes(self):
            return self._m.nnodes

        def getTotalElements(self):
            return self._m.nel

        def getBoundingBox(self):
            return tuple(self._cfg["lower"]), tuple(self._cfg["upper"])

        def computeFullCoordinates(self, spElem):
            dim = self.dim
            self.fullCoordVec = PETSc.Vec().createSeq(self._m.nnodes * dim)
            self.fullCoordVec.setName("NodeCoordinates")
            for cell in range(self._m.nel):
                coords = self.getCellCornersCoords(cell)
                coords.shape = (2 ** dim, dim)
                nodes = self.getGlobalNodesFromCell(cell, shared=True)
                inds = self.indicesManager.mapNodesToIndices(nodes, dim)
                tot = np.zeros(dim * spElem.nnode)
                for idx, _gp in enumerate(spElem.gpsOp):                       # dmplex.py:83-86
                    tot[[dim * idx + d for d in range(dim)]] = (spElem.HCooOp[idx] @ coords).T
                self.fullCoordVec.setValues(inds, tot)
            self.fullCoordVec.assemble()
            self.startNode = 0
            self.nodes = list(range(self._m.nnodes))

        def createGlobalVec(self):
            """after computeFullCoordinates the default section is the coordinate section (dmplex.py:66-70):
            dim components per node"""
            return PETSc.Vec().createSeq(self._m.nnodes * self.dim)

        # ---- tables ----------------------------------------------------------------------------------
        def getCellCornersCoords(self, cell):
            if cell + self.cellStart >= self.cellEnd:
                raise Exception('elem parameter must be in local numbering!')
            return np.array(self._m.corner_xyz[cell], dtype=np.float64).ravel()

        def setLabelToBorders(self):
            pass

        def getBorderNodes(self, name):
            if name not in self._m.sides:
                return []
            return [int(n) for n in self._m.sides[name]]

        def getGlobalNodesFromCell(self, cell, shared):
            return [int(n) for n in self._m.elem_nodes[cell]]

        def getGlobalNodesFromEntities(self, entities, shared):
            nodes = set()
            for e in entities:
                nodes |= set(self.getGlobalNodesFromCell(e, shared))
            return nodes

        def getNodesFromLabel(self, label, shared=False):
            return set(int(n) for n in self._m.border_nodes) if label == "External Boundary" else set()

        def getAllNodes(self):
            return list(range(self._m.nnodes))

        def getNodesRange(self):
            return 0, self._m.nnodes

        def getConnectivityNodes(self):
            n = self._m.nnodes
            conn = [set() for _ in range(n)]
            for nodes in self._m.elem_nodes.tolist():
                s = set(nodes)
                for a in nodes:
                    conn[a] |= s
            ind_d = np.zeros(n, dtype=set)
            ind_o = np.zeros(n, dtype=set)
            nnz_diag = np.zeros(n, dtype=np.int32)
            nnz_off = np.zeros(n, dtype=np.int32)
            for a in range(n):
                ind_d[a] = conn[a]
                ind_o[a] = set()
                nnz_diag[a] = len(conn[a])
            return ind_d, ind_o, nnz_diag, nnz_off

        def getEdgesWidth(self):
            c = self._cfg
            return (c["upper"][0] - c["lower"][0]) / c["nelem"][0]

    return StandInBoxDom
