"""CPU: the XDMF + raw-binary writer behind the post-step hook (the reference's
viewer/paraviewer.py:14-78 writes HDF5 through PETSc viewers, which do not exist here)."""
import os
import xml.etree.ElementTree as ET

import numpy as np


class _V:
    def __init__(self, a, name):
        self.a, self.name = np.asarray(a, dtype=float), name

    def getGlobalArray(self):
        return self.a

    def getName(self):
        return self.name


def test_xdmf_index_and_binary_payload(tmp_path):
    from pynama_b200.petsc import Options
    from pynama_b200.viewer.paraviewer import Paraviewer
    Options().setValue("pynama_save", 1)
    try:
        v = Paraviewer()
    finally:
        Options().delValue("pynama_save")
    d = str(tmp_path / "out")
    v.configure(3, d)
    n = 17
    rng = np.random.default_rng(0)
    xyz = rng.uniform(0, 1, (n, 3))
    v.saveMesh(xyz.ravel())
    data = {}
    for step, t in ((0, 0.0), (3, 0.125)):
        vel, vort, p = rng.uniform(-1, 1, (n, 3)), rng.uniform(-1, 1, (n, 3)), rng.uniform(-1, 1, n)
        v.saveData(step, t, _V(vel.ravel(), "velocity"), _V(vort.ravel(), "vorticity"), _V(p, "pressure"))
        data[step] = (vel, vort, p)
    v.writeXmf("case")
    root = ET.parse(os.path.join(d, "case.xmf")).getroot()
    assert root.tag == "Xdmf"
    coll = root.find("Domain/Grid")
    assert coll.get("CollectionType") == "Temporal"
    grids = coll.findall("Grid")
    assert [float(g.find("Time").get("Value")) for g in grids] == [0.0, 0.125]
    for g, step in zip(grids, (0, 3)):
        assert g.find("Topology").get("TopologyType") == "Polyvertex" and g.find("Topology").get("Dimensions") == str(n)
        geo = g.find("Geometry/DataItem")
        assert geo.get("Format") == "Binary" and geo.get("Precision") == "8"
        assert np.array_equal(np.fromfile(os.path.join(d, geo.text), dtype="<f8").reshape(n, 3), xyz)
        attrs = {a.get("Name"): a for a in g.findall("Attribute")}
        assert {k: a.get("AttributeType") for k, a in attrs.items()} == \
            {"velocity": "Vector", "vorticity": "Vector", "pressure": "Scalar"}
        for name, want in zip(("velocity", "vorticity", "pressure"), data[step]):
            item = attrs[name].find("DataItem")
            got = np.fromfile(os.path.join(d, item.text), dtype="<f8")
            assert np.array_equal(got, np.asarray(want).ravel())
            assert [int(x) for x in item.get("Dimensions").split()] == list(np.asarray(want).shape)


def test_nothing_is_written_by_default(tmp_path):
    from pynama_b200.viewer.paraviewer import Paraviewer
    v = Paraviewer()
    d = str(tmp_path / "out")
    v.configure(2, d)
    v.saveMesh(np.zeros(8))
    v.saveData(0, 0.0, _V(np.zeros(8), "velocity"))
    v.writeXmf("case")
    assert not os.path.exists(d)
