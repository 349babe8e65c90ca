"""Output hook of the post-step callback.

The reference writes one HDF5 file per saved step through PETSc viewers plus an XDMF index that
ParaView opens (src/viewer/paraviewer.py:14-78, src/viewer/xml_generator.py).  PETSc viewers and
h5py do not exist here, so the heavy data goes to raw little-endian float64 files and the XDMF
index references them with ``Format="Binary"`` -- the same temporal collection of point clouds
(``Polyvertex`` topology, node coordinates as geometry, one ``Vector`` or ``Scalar`` attribute per
saved Vec), readable by the same ParaView reader.

File output is outside the hot path: nothing is written unless ``-pynama_save 1`` is in the options
database and the case gives a save directory.  Same call surface as the reference class
(``configure``, ``saveMesh``, ``saveData``, ``writeXmf``, ``writeYaml``).  On a row-partitioned
run every Vec is gathered and rank 0 writes.
"""
import os

import numpy as np

from ..petsc import COMM_WORLD, Options


class Paraviewer:
    comm = COMM_WORLD

    def __init__(self):
        self.saveDir = None
        self.enabled = bool(Options().getBool("pynama_save", False))
        self._steps = []               # [(step, time, [(name, kind, ncomp, file)])]
        self._nnodes = 0
        self._mesh_file = None          # set by saveMesh; xdmf() leaves the geometry out while it is None
        self.dim = None

    def configure(self, dim, saveDir=None):
        self.dim = dim
        self.saveDir = saveDir
        if self.enabled and saveDir and not self.comm.rank:
            os.makedirs(saveDir, exist_ok=True)

    @property
    def _active(self):
        return self.enabled and bool(self.saveDir)

    def _write(self, fname, arr):
        if not self.comm.rank:
            np.ascontiguousarray(arr, dtype="<f8").tofile(os.path.join(self.saveDir, fname))

    def saveMesh(self, coords, name="mesh"):
        if not self._active:
            return
        c = np.asarray(coords.getGlobalArray() if hasattr(coords, "getGlobalArray") else coords, dtype=np.float64)
        self._nnodes = c.size // self.dim
        self._write(f"{name}.bin", c)
        self._mesh_file = f"{name}.bin"

    def saveData(self, step, time, *vecs):
        if not self._active:
            return
        items = []
        for v in vecs:
            a = v.getGlobalArray() if hasattr(v, "getGlobalArray") else np.asarray(v)
            vname = (v.getName() if hasattr(v, "getName") else None) or "vec"
            ncomp = max(a.size // max(self._nnodes, 1), 1)
            fname = f"{vname}-{int(step):05d}.bin"
            self._write(fname, a)
            items.append((vname, "Scalar" if ncomp == 1 else "Vector", ncomp, fname))
        self._steps.append((int(step), float(time), items))

    def xdmf(self):
        """the XDMF index of everything saved so far, as a string"""
        n, dim = self._nnodes, self.dim
        geo = "XY" if dim == 2 else "XYZ"
        out = ['<?xml version="1.0" ?>', '<Xdmf Version="2.0">', ' <Domain>',
               '  <Grid Name="TimeSeries" GridType="Collection" CollectionType="Temporal">']
        for step, time, items in self._steps:
            out += [f'   <Grid Name="mesh1" GridType="Uniform">',
                    f'    <Time Value="{time!r}"/>',
                    f'    <Topology TopologyType="Polyvertex" Dimensions="{n}"/>']
            if self._mesh_file is not None:            # createMesh(saveMesh=False): data without a geometry block
                out += [f'    <Geometry GeometryType="{geo}">',
                        f'     <DataItem Dimensions="{n} {dim}" NumberType="Float" Precision="8" Format="Binary" '
                        f'Endian="Little">{self._mesh_file}</DataItem>',
                        '    </Geometry>']
            for vname, kind, ncomp, fname in items:
                dims = f"{n}" if ncomp == 1 else f"{n} {ncomp}"
                out += [f'    <Attribute Name="{vname}" AttributeType="{kind}" Center="Node">',
                        f'     <DataItem Dimensions="{dims}" NumberType="Float" Precision="8" Format="Binary" '
                        f'Endian="Little">{fname}</DataItem>',
                        '    </Attribute>']
            out.append('   </Grid>')
        out += ['  </Grid>', ' </Domain>', '</Xdmf>', '']
        return "\n".join(out)

    def writeXmf(self, name):
        if self._active and not self.comm.rank:
            with open(os.path.join(self.saveDir, f"{name}.xmf"), "w") as f:
                f.write(self.xdmf())

    def writeYaml(self, name, data):
        if self._active and not self.comm.rank:
            import yaml
            data = dict(data, dir=self.saveDir)
            with open(os.path.join(self.saveDir, f"{name}.yaml"), "w") as f:
                yaml.dump(data, f, default_flow_style=False)
