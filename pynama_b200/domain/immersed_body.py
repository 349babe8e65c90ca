"""Lagrangian bodies of the 2-D immersed-boundary path: host mirror of
src/domain/immersed_body.py (BodiesContainer :8-128, ImmersedBody :130-291, Line :293-313,
OpenBox :315-363, Circle :365-396, the delta functions :398-424).

The reference stores a body as a 1-D DMPlex and its velocity as a PETSc Vec; here a body is a
numpy array of marker coordinates (set-up data, a few hundred points) and its velocity a device
``Vec``.  The delta weights themselves are evaluated on the device (csrc/ibm.cu); the scalar
functions below are the reference's ``body.dirac`` hooks kept for callers that probe them.
"""
import logging
from math import ceil, cos, pi, sin, sqrt

import numpy as np

from ..petsc import Vec


# ---- delta functions, scalar (immersed_body.py:398-424) ---------------------------------------
def threeGrid(r):
    """three-cell support"""
    if r > 1.5:
        return 0
    if r > 0.5:
        return (5 - 3 * r - sqrt(-3 * (1 - r) ** 2 + 1)) / 6
    return (1 + sqrt(-3 * r ** 2 + 1)) / 3


def linear(r):
    """hat function"""
    return (1 - r) if r < 1 else 0


def fourGrid(r):
    """four-cell support (Peskin)"""
    if r > 2:
        return 0
    if r > 1:
        return (5 - 2 * r - sqrt(-7 + 12 * r - 4 * r ** 2)) / 8
    return (3 - 2 * r + sqrt(1 + 4 * r - 4 * r ** 2)) / 8


DIRAC_KIND = {"fourGrid": 0, "threeGrid": 1, "linear": 2}


class ImmersedBody:
    """One body: markers = local coordinates + centre (immersed_body.py:256-259)."""

    def __init__(self, vel=(0, 0), center=(0, 0)):
        self.dirac = fourGrid
        self._center = list(center)
        self._startCenter = list(center)
        self._dl = None
        self._vel = list(vel)
        self._Uref = None
        self._isStatic = True
        self._local = None                 # [npoints, 2] marker coordinates relative to the centre
        self._velVec = None
        self._history = {"times": [], "displ": [], "vel": []}
        self._lagNodes = []
        self.logger = logging.getLogger("Body Immersed")

    # -- geometry --------------------------------------------------------------------------------
    def generateBody(self, dh):
        """-> (local coordinates [n, 2], arc length dl); overridden per shape"""
        raise NotImplementedError

    def generateDMPlex(self, dh, dim=1):
        """Name kept from the reference (:169-180): builds the markers and the velocity Vec."""
        self._local, self._dl = self.generateBody(dh)
        self._L = self.getLong()
        self.firstNode, self.lastNode = 0, len(self._local)
        self._lagNodes = [0] * len(self._local)
        self._velVec = Vec(2 * len(self._local))
        self.updateVelocity()

    def getTotalNodes(self):
        return self.lastNode - self.firstNode

    def getNodeCoordinates(self, node):
        return self._local[node] + np.asarray(self._center, dtype=np.float64)

    def getAllCoordinates(self):
        """every marker at once: [n, 2] (same sum as getNodeCoordinates, row by row)"""
        return self._local + np.asarray(self._center, dtype=np.float64)

    def getElementLong(self):
        return self._dl

    def setElementLong(self, dl, normals):
        self._normals, self._dl = normals, dl

    def getLong(self):
        return None

    def getRegion(self):
        return None

    def setCaracteristicLong(self, val):
        self._L = val

    def getCaracteristicLong(self):
        return self._L

    def getCenterBody(self):
        return self._center

    def setCenter(self, val):
        self._center = val

    def setEulerNodes(self, lagPoi, NNZNodes):
        self._lagNodes[lagPoi] = NNZNodes

    # -- kinematics ------------------------------------------------------------------------------
    def setVelRef(self, vel):
        self._Uref = vel

    def getVelRef(self):
        return self._Uref

    def setIsMoving(self):
        self._isStatic = False

    def isStatic(self):
        return self._isStatic

    def getVelocity(self):
        return self._velVec

    def setVelocity(self, inds, values):
        self._velVec.setValues(inds, values)
        self._velVec.assemble()

    def updateVelocity(self):
        n = self.getTotalNodes()
        self.setVelocity(np.arange(2 * n), np.tile(np.asarray(self._vel, dtype=np.float64), n))

    def updateBodyParameters(self, t):
        """prescribed heave of a 'dynamic' body (:268-284): y = A sin(2 pi t / Te), Te = f / Uref"""
        if self._isStatic:
            return
        f, A = 5, 0.3
        Te = f / self._Uref
        self._center = [0 + self._startCenter[0], A * sin(2 * pi * t / Te) + self._startCenter[1]]
        self._vel = [0, 2 * pi * A * cos(2 * pi * t / Te) / Te]
        self.updateVelocity()
        for key, val in (("times", t), ("displ", self._center), ("vel", self._vel)):
            self._history[key].append(val)

    # -- forces / weights ------------------------------------------------------------------------
    def computeForce(self, q):
        q = np.asarray(q)
        fx = fy = 0
        for poi in range(self.getTotalNodes()):          # sequential sum, as the reference
            fx += q[poi * 2]
            fy += q[poi * 2 + 1]
        return fx, fy

    def getDiracs(self, dist, h):
        d = 1
        for r in dist:
            d *= self.dirac(abs(r) / h)
            d /= h
        return d

    def view(self):
        self.logger.info(f"Arc len: {self._dl} | Dirac Type: {self.dirac.__name__} | "
                         f"Vel Fluid Reference: {self._Uref} ")

    def viewState(self):
        self.logger.info(f"{self.__class__.__name__} vel: {self._vel} | Body center position {self._center}")


class Line(ImmersedBody):
    def generateBody(self, dl, longitud=2):
        self._region = longitud
        n = ceil(longitud / dl)
        xy = np.zeros((n, 2))
        xy[:, 0] = np.linspace(0, longitud, n)
        return xy, dl

    def getLong(self):
        return 1

    def getRegion(self):
        return self._region


class OpenBox(ImmersedBody):
    """the rotated square of the immersed lid-driven cavity"""

    def generateBody(self, dl, longitud=1):
        n = ceil(sqrt(2) / dl)
        L = longitud
        path = [(0, L), (-L, 0), (0, -L), (L, 0), (0, L)]
        sides = [np.stack([np.linspace(p[0], q[0], n, endpoint=False),
                           np.linspace(p[1], q[1], n, endpoint=False)], axis=1)
                 for p, q in zip(path[:-1], path[1:])]
        return np.concatenate(sides), dl

    def updateVelocity(self):
        """lid = the markers of the first quadrant, moving along the lid (:347-357)"""
        uref = self.getVelRef()
        xy = self.getAllCoordinates()
        vals = np.zeros((len(xy), 2))
        if uref is not None:
            lid = (xy[:, 0] >= 0) & (xy[:, 1] >= 0)
            vals[lid, 0] = uref / sqrt(2)
            vals[lid, 1] = -uref / sqrt(2)
        self.setVelocity(np.arange(2 * len(xy)), vals.ravel())

    def getLong(self):
        return 1

    def getRegion(self):
        return 1


class Circle(ImmersedBody):
    def __init__(self, vel, center, radius):
        super().__init__(vel, center)
        self._radius = radius

    def generateBody(self, dh):
        r = self._radius
        n = ceil(2 * pi * r / dh)
        ang = np.linspace(0, 2 * pi, n, endpoint=False) + pi / 1000
        return np.stack([r * np.cos(ang), r * np.sin(ang)], axis=1), 2 * pi * r / n

    def getLong(self):
        return self._radius * 2

    def getRegion(self):
        return self._radius + 2 * self.getElementLong()


BODY_TYPES = {"circle": lambda c: Circle(vel=[0, 0], center=c["center"], radius=c["radius"]),
              "line": lambda c: Line(vel=[0, 0], center=c["center"]),
              "box": lambda c: OpenBox(vel=[0, 0], center=c["center"])}


class BodiesContainer:
    types = ["circle", "line", "box"]

    def __init__(self, bodies):
        self.logger = logging.getLogger("Bodies Container")
        self.bodies = []
        for cfg in bodies:
            if cfg["type"] not in BODY_TYPES:
                raise Exception("not defined")
            body = BODY_TYPES[cfg["type"]](cfg)
            if cfg["vel"] == "dynamic":
                body.setIsMoving()
            self.bodies.append(body)

    def createBodies(self, h):
        for i, body in enumerate(self.bodies):
            body.generateDMPlex(h)
            self.logger.info(f"Body number: {i}")
            body.viewState()
        self.locTotalNodes = self.bodies[-1].getTotalNodes()
        self._offsets = np.concatenate([[0], np.cumsum([b.getTotalNodes() for b in self.bodies])])

    def bodyNumbers(self):
        return len(self.bodies)

    def getTotalNodes(self):
        return sum(b.getTotalNodes() for b in self.bodies)

    def mapGlobToLocal(self, globNode):
        """two-body rule of the reference (:46-53): bodies of equal size"""
        if globNode >= self.locTotalNodes:
            return globNode - self.locTotalNodes, 1
        return globNode, 0

    def getNodeCoordinates(self, globNode):
        loc, num = self.mapGlobToLocal(globNode)
        return self.bodies[num].getNodeCoordinates(loc)

    def getAllCoordinates(self):
        return np.concatenate([b.getAllCoordinates() for b in self.bodies])

    def getDiracs(self, dist, h):
        return self.bodies[0].getDiracs(dist, h)

    def getDiracKind(self):
        return DIRAC_KIND[self.bodies[0].dirac.__name__]

    def getElementLong(self):
        return self.bodies[0].getElementLong()

    def getRegion(self):
        return 0.5 + 2 + self.getElementLong() * 4

    def getCenters(self):
        return [b.getCenterBody() for b in self.bodies]

    def setEulerNodes(self, glNode, eulerNodesNum):
        loc, num = self.mapGlobToLocal(glNode)
        self.bodies[num].setEulerNodes(loc, eulerNodesNum)

    def computeForce(self, vec, dt):
        q = vec.getArray()
        fxs, fys = [], []
        for b, o in zip(self.bodies, self._offsets[:-1]):
            fx, fy = b.computeForce(q[2 * o: 2 * (o + b.getTotalNodes())])
            fxs.append(float(fx / dt))
            fys.append(float(fy / dt))
        return fxs, fys

    def setVelRef(self, vel):
        for b in self.bodies:
            b.setVelRef(vel)

    def getVelRef(self, bodyNum=0):
        return self.bodies[bodyNum].getVelRef()

    def updateBodyParameters(self, t):
        for b in self.bodies:
            b.updateBodyParameters(t)

    def updateVelocity(self):
        for b in self.bodies:
            b.updateVelocity()

    def getVelocity(self):
        """the bodies' velocities one after the other (the reference nests the Vecs, :116-124)"""
        if len(self.bodies) == 1:
            return self.bodies[0].getVelocity()
        return Vec().createWithArray(np.concatenate([b.getVelocity().getArray() for b in self.bodies]))

    def viewBodies(self):
        for i, b in enumerate(self.bodies):
            self.logger.info(f"Body num: {i}")
            b.viewState()
