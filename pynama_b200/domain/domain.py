"""Domain facade over the mesh tables, the boundary conditions and the device mesh.

Same public surface as the reference's ``domain.domain.Domain``
(src/domain/domain.py:45-293) for everything the hot path and its callers use
-- ``configure / setOptions / setUp``, dimensions, node and cell ranges,
Dirichlet / no-slip sets, coordinates, ``applyBoundaryConditions*`` -- but the
per-cell ``computeLocalKLEMats`` loop is replaced by one batched device call
(``getDeviceMesh``).  The mesh itself comes from ``BoxMesh`` (structured box
meshes, the 'box-mesh' entry of the case YAML); Gmsh input is not part of this
path yet and is refused loudly.
"""
import copy
import logging

import numpy as np

from ..boundaries.boundary_conditions import BoundaryConditions
from ..petsc import COMM_WORLD, Vec
from .box_mesh import BoxMesh
from .gmsh_mesh import GmshMesh


class Domain:
    comm = COMM_WORLD

    def __init__(self):
        self.logger = logging.getLogger(f"{[self.comm.rank]}:Domain:")
        self._meshType = None
        self._bc = None
        self._ngl = None
        self._mesh = None
        self._dmesh = None
        self._perturb = 0.0
        self.fullCoordVec = None

    # ---- configuration (domain.py:45-75) ----------------------------------------------------
    def configure(self, inp):
        data = copy.deepcopy(inp)
        if "domain" in data:
            self.domData = data["domain"]
        if "boundary-conditions" in data:
            self.bcData = data["boundary-conditions"]

    def setOptions(self, **kwargs):
        for key, val in kwargs.items():
            if key == "ngl":
                self.domData[key] = val
            elif key in ("lower", "upper", "nelem"):
                self.domData["box-mesh"][key] = val
            elif key == "perturb":
                self._perturb = float(val)
            elif key in ("freeSlip", "noSlip", "constant"):
                k = {"freeSlip": "free-slip", "noSlip": "no-slip"}.get(key, key)
                self.bcData[k] = val
            elif key in ("keepCoords", "case", "chart"):
                pass
            else:
                self.logger.warning("Unknown option for domain")

    def create(self):
        assert self.domData, "Domain not defined"
        if "box-mesh" in self.domData:
            self._meshType = "box"
            cfg = self.domData["box-mesh"]
            self._ngl = self.domData["ngl"]
            self._mesh = BoxMesh(cfg["nelem"], cfg["lower"], cfg["upper"], self._ngl,
                                 perturb=self._perturb)
        elif "gmsh-file" in self.domData:
            # unstructured quad/hex mesh (GmshDom, dmplex.py:389-395)
            self._meshType = "gmsh"
            cfg = self.domData["gmsh-file"]
            self._ngl = self.domData["ngl"]
            self._mesh = GmshMesh(cfg if isinstance(cfg, str) else cfg["fileName"], self._ngl)
        else:
            raise Exception("Mesh Type not defined")
        del self.domData
        self.logger.info(f"mesh type: {self._meshType} created")

    def setUp(self):
        self.create()
        self.setUpBoundaryConditions()
        self.setUpBoundaryCoordinates()

    def setUpBoundaryConditions(self):
        assert self.bcData, "Boundary Conditions Not defined"
        bcs = BoundaryConditions(self._mesh.getBordersNames())
        bcs.setBoundaryConditions(self.bcData)
        for name in bcs.getNames():
            bcs.setBoundaryNodes(name, self._mesh.getBorderNodes(name))
        self._bc = bcs
        del self.bcData

    def setUpBoundaryCoordinates(self):
        for name in self._bc.getBordersNeedsCoords():
            inds = self._bc.getIndicesByName(name)
            self._bc.setBoundaryCoords(name, self.getCoordinates(inds))

    # ---- simple queries ----------------------------------------------------------------------
    def getDM(self):
        return self._mesh

    def getMesh(self):
        return self._mesh

    def getMeshType(self):
        return self._meshType

    def getBoundaryType(self):
        if self._bc is None:
            raise Exception("Boundary Type not defined")
        return self._bc.getType()

    def getBoundaryConditions(self):
        return self._bc

    def getDimension(self):
        return self._mesh.dim

    def getDimensions(self):
        dim = self._mesh.dim
        return dim, (1 if dim == 2 else 3), (3 if dim == 2 else 6)

    def getNGL(self):
        return self._ngl

    def getNumOfElements(self):
        return self._mesh.nel

    def getNumOfNodes(self):
        return self._mesh.nnodes

    def getAllNodes(self):
        return list(range(self._mesh.nnodes))

    def getNodesRange(self):
        return 0, self._mesh.nnodes

    def getLocalCellRange(self):
        return 0, self._mesh.nel

    def getExtremeCoords(self):
        return self._mesh.lower, self._mesh.upper

    # ---- coordinates ---------------------------------------------------------------------------
    def getFullCoordArray(self):
        return self._mesh.coords

    def getFullCoordVec(self):
        if self.fullCoordVec is None:
            self.fullCoordVec = Vec().createWithArray(self._mesh.coords.ravel())
            self.fullCoordVec.setName("NodeCoordinates")
        return self.fullCoordVec

    def getNodesCoordinates(self, nodes):
        return self._mesh.coords[np.asarray(nodes, dtype=np.int64)]

    def getCoordinates(self, indices):
        dim = self._mesh.dim
        indices = np.asarray(indices, dtype=np.int64)
        n = len(indices) // dim
        return self._mesh.coords.ravel()[indices[: n * dim]].reshape(n, dim)

    def getCellCentroid(self, cell):
        return self._mesh.corner_xyz[cell].mean(axis=0)

    def getCellCentroids(self):
        """every cell at once: [nel, dim]"""
        return self._mesh.corner_xyz.mean(axis=1)

    def getNodesCoordsFromEntities(self, entities):
        """nodes of a set of cells (ascending ids) and their coordinates (domain.py:185-188)"""
        ents = np.asarray(list(entities), dtype=np.int64)
        nodes = np.unique(self._mesh.elem_nodes[ents].ravel()) if len(ents) else np.zeros(0, np.int64)
        return nodes, self._mesh.coords[nodes]

    def getNodeSeparationIBM(self):
        """width of the first edge of the mesh over the number of GLL intervals (domain.py:212-217)"""
        c = self._mesh.corner_xyz[0]
        return float(np.linalg.norm(c[1] - c[0])) / (self._ngl - 1)

    def getBorderNodes(self, borderName):
        return list(self._mesh.getBorderNodes(borderName))

    def getBoundaryNodes(self):
        return set(self._mesh.getBordersNodes().tolist())

    # ---- Dirichlet / no-slip sets (domain.py:228-250) -----------------------------------------
    def getGlobalIndicesDirichlet(self):
        return self._bc.getIndicesByType("free-slip")

    def getNodesDirichlet(self, collect=False):
        return self._bc.getNodesByType("free-slip", allGather=collect)

    def getGlobalIndicesNoSlip(self):
        return self._bc.getIndicesByType("no-slip")

    def getNodesNoSlip(self, collect=False):
        return self._bc.getNodesByType("no-slip", allGather=collect)

    def getTangDofs(self, collect=False):
        return self._bc.getNoSlipTangDofs(allGather=collect)

    def getNormalDofs(self, collect=False):
        return self._bc.getNoSlipNormalDofs(allGather=collect)

    def getFreeStreamVelocity(self):
        return np.linalg.norm(self._bc.getFreeStreamVelocity())

    def getDofClasses(self):
        """Per velocity dof: 0 free, 1 Dirichlet (free-slip BC nodes, mat_fs.py:158-172),
        2 no-slip normal, 3 no-slip tangential with normal taking precedence
        (mat_ns.py:59-62)."""
        dim = self._mesh.dim
        cls = np.zeros(self._mesh.nnodes * dim, dtype=np.uint8)
        bctype = self._bc.getType()
        if bctype == "FS":
            nodes = np.fromiter(self.getNodesDirichlet(collect=True), dtype=np.int64)
            for d in range(dim):
                cls[nodes * dim + d] = 1
        elif bctype == "NS":
            tang = np.fromiter(self.getTangDofs(collect=True), dtype=np.int64)
            normal = np.fromiter(self.getNormalDofs(collect=True), dtype=np.int64)
            cls[tang] = 3
            cls[normal] = 2
        else:
            raise Exception("FSNS Mat not implemented")
        return cls

    # ---- the batched replacement of computeLocalKLEMats / computeLocalOperators -------------------
    def getDeviceMesh(self):
        if self._dmesh is None:
            from ..device_mesh import DeviceMesh
            m = self._mesh
            cls = self.getDofClasses()
            if self.comm.size > 1:
                # row partition over the GPUs of the box (replaces DMPlex.distribute, dmplex.py:21)
                from ..parallel import Partition
                from ..petsc import Layout
                part = Partition(m, self.comm.rank, self.comm.size)
                self._partition = part
                layout = Layout(part, m.nnodes)
                self._dmesh = DeviceMesh(m.dim, m.ngl, part.corner_xyz, part.elem_nodes, part.n_loc,
                                         layout=layout)
                ldofs = (part.l2g[:, None] * m.dim + np.arange(m.dim)[None, :]).ravel()
                cls = cls[ldofs]
                from ..petsc import Options as _O
                if _O().getString("pc_type", "") == "asm" and _O().getInt("pc_asm_coarse", 0) > 0:
                    # the coarse level is replicated: every rank needs the global element table
                    self._dmesh.set_global_tables(m.elem_nodes, m.nnodes, part.elems)
            else:
                self._dmesh = DeviceMesh(m.dim, m.ngl, m.corner_xyz, m.elem_nodes, m.nnodes)
            from ..petsc import Options
            tol = Options().getReal("pynama_congruence_tol", None)
            if tol is not None:
                # -pynama_congruence_tol 0: evaluate every cell on its own (domain.py:262-276)
                self._dmesh.set_congruence(tol)
            self._dmesh.set_dof_class(cls)
        return self._dmesh

    def computeLocalKLEMats(self, cell):
        """Per-cell view kept for the reference's call shape (domain.py:262-270); values
        come from the device kernels."""
        dm = self.getDeviceMesh()
        K, Rw, Rd = dm.elem_kle(cell, 1)
        nodes = self._mesh.elem_nodes[cell].tolist()
        dim, dim_w, _ = self.getDimensions()
        ivel = [n * dim + d for n in nodes for d in range(dim)]
        iw = [n * dim_w + d for n in nodes for d in range(dim_w)]
        return nodes, (ivel, iw), (K[0], Rw[0], Rd[0])

    def computeLocalOperators(self, cell):
        dm = self.getDeviceMesh()
        S, D, C, W = dm.elem_ops(cell, 1)
        return self._mesh.elem_nodes[cell].tolist(), (S[0], D[0], C[0], W[0])

    # ---- boundary values into Vecs (domain.py:279-285) ------------------------------------------------
    def applyBoundaryConditions(self, vec, varName, t=None, nu=None):
        assert self._bc, "Boundary conditions not defined"
        self._bc.setValuesToVec(vec, varName, t, nu)

    def prefetchBoundaryConditions(self, varNames, t, nu):
        if self._bc is not None and t is not None:
            self._bc.prefetch(varNames, t, nu)

    def applyBoundaryConditionsNS(self, vec, varName, t=None, nu=None):
        assert self._bc, "Boundary conditions not defined"
        self._bc.setTangentialValuesToVec(vec, varName, t, nu)

    def applyValuesToVec(self, nodes, vals, vec):
        dim = self._mesh.dim
        nodes = np.asarray(nodes, dtype=np.int64)
        assert len(vals) <= dim
        if len(vals) == 1:
            vec.set(vals[0])
        else:
            inds = (nodes[:, None] * dim + np.arange(dim)[None, :]).ravel()
            vec.setValues(inds, np.tile(vals, len(nodes)), addv=False)
        return vec

    def view(self):
        print(f"Domain dimensions: {self.getDimension()}  mesh: {self._meshType}  ngl: {self._ngl}")
        print(f"elements: {self.getNumOfElements()}  nodes: {self.getNumOfNodes()}")
        print(self._bc)
