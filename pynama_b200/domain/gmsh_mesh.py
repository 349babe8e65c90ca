"""Unstructured quad/hex spectral-element mesh tables from a Gmsh 4.1 ASCII file
(the reference's GmshDom, src/domain/dmplex.py:389-395, which hands the file to
DMPlexCreateFromFile).

Same table surface as ``BoxMesh`` (inputs A0 of the hot path): corner coordinates per cell in
DMPlex closure-vertex order, element -> global node table in the reference's local order, node
coordinates (corner basis at the nodal points times the corner coordinates, dmplex.py:73-88: the
elements are straight-sided), and the node sets of the borders.

* Cells: 4-node quadrangles (Gmsh type 3) or 8-node hexahedra (type 5).  Gmsh's hexahedron vertex
  order (bottom face counter-clockwise, then top face) is mapped to the closure order pinned by
  src/tests/test_dmplex_wrapper.py:105-115.
* Global GLL nodes: generated per element and identified across elements by their coordinates, so
  shared edges and faces match whatever the relative orientation of the neighbouring cells is.
  Numbering = lexicographic order of the coordinates (z, y, x); like the box numbering for
  ngl >= 3 it differs from PETSc's by a permutation ("numbering").
* Borders: the boundary elements of the file (lines in 2-D, quadrangles in 3-D) carry physical
  tags, which DMPlex turns into the "Face Sets" label; the reference maps border *names* to face-set
  ids by position in its naming convention (dmplex.py:26-29,167-170: "down" = 1, "right" = 2, ...).
  A physical group whose name equals a border name takes precedence over the positional rule.
"""
import numpy as np

from .box_mesh import CORNERS, SIDES, corner_basis_at_nodes

GMSH_HEX_TO_CLOSURE = [0, 3, 2, 1, 4, 5, 6, 7]      # closure corner c is Gmsh vertex GMSH_HEX_TO_CLOSURE[c]
CELL_TYPE = {2: 3, 3: 5}                             # dim -> Gmsh element type of the cells
FACE_TYPE = {2: 1, 3: 3}                             # dim -> Gmsh element type of the boundary elements
NODES_OF_TYPE = {15: 1, 1: 2, 3: 4, 5: 8}


def read_msh(path):
    """Minimal Gmsh ASCII reader (MSH 4.1 and MSH 2.2) -> dict(physical_names, entity_physical, nodes, elements).
    nodes: {tag: xyz}; elements: {type: (entity_dim, entity_tag, node tags)[]}"""
    with open(path) as f:
        tok = f.read().split("\n")
    sections, name, buf = {}, None, []
    for line in tok:
        s = line.strip()
        if s.startswith("$End"):
            sections[name] = buf
            name, buf = None, []
        elif s.startswith("$"):
            name, buf = s[1:], []
        elif name is not None and s:
            buf.append(s)
    ver = float(sections["MeshFormat"][0].split()[0])
    if int(sections["MeshFormat"][0].split()[1]) != 0 or not (2.0 <= ver < 3.0 or 4.0 <= ver < 5.0):
        raise ValueError(f"{path}: only Gmsh 2.x and 4.x ASCII files are supported (found {sections['MeshFormat'][0]})")
    phys = {}
    for s in sections.get("PhysicalNames", [])[1:]:
        d, tag, nm = s.split(None, 2)
        phys[(int(d), int(tag))] = nm.strip().strip('"')
    if ver < 3.0:
        return _read_msh2(sections, phys)
    ent_phys = {}
    ent = sections.get("Entities", [])
    if ent:
        counts = [int(v) for v in ent[0].split()]
        i = 1
        for d, n in enumerate(counts):
            for _ in range(n):
                v = ent[i].split()
                i += 1
                nb = 4 if d == 0 else 7                      # tag + point (3) or bounding box (6)
                nphys = int(v[nb])
                ent_phys[(d, int(v[0]))] = [int(t) for t in v[nb + 1:nb + 1 + nphys]]
    nodes = {}
    sec = sections["Nodes"]
    nblocks = int(sec[0].split()[0])
    i = 1
    for _ in range(nblocks):
        _, _, parametric, n = (int(v) for v in sec[i].split())
        tags = [int(sec[i + 1 + k]) for k in range(n)]
        for k, t in enumerate(tags):
            nodes[t] = [float(v) for v in sec[i + 1 + n + k].split()[:3]]
        i += 1 + 2 * n
    elements = {}
    sec = sections["Elements"]
    nblocks = int(sec[0].split()[0])
    i = 1
    for _ in range(nblocks):
        ed, et, typ, n = (int(v) for v in sec[i].split())
        nn = NODES_OF_TYPE.get(typ)
        for k in range(n):
            v = [int(t) for t in sec[i + 1 + k].split()]
            if nn is not None:
                elements.setdefault(typ, []).append((ed, et, v[1:1 + nn]))
        i += 1 + n
    return {"physical_names": phys, "entity_physical": ent_phys, "nodes": nodes, "elements": elements}


DIM_OF_TYPE = {15: 0, 1: 1, 3: 2, 5: 3}


def _read_msh2(sections, phys):
    """MSH 2.x ASCII (what DMPlexCreateFromFile of PETSc 3.12 also reads): `$Nodes` = count, then `id x y z`;
    `$Elements` = count, then `id type ntags physical elementary ... node ids`.  Mapped to the structure of the
    4.1 reader: the elementary tag is the entity, its physical tag the entity's physical group."""
    nodes = {}
    for s in sections["Nodes"][1:]:
        v = s.split()
        nodes[int(v[0])] = [float(x) for x in v[1:4]]
    elements, ent_phys = {}, {}
    for s in sections["Elements"][1:]:
        v = [int(t) for t in s.split()]
        typ, ntags = v[1], v[2]
        nn = NODES_OF_TYPE.get(typ)
        if nn is None:
            continue
        physical = v[3] if ntags >= 1 else 0
        elementary = v[4] if ntags >= 2 else physical
        d = DIM_OF_TYPE[typ]
        elements.setdefault(typ, []).append((d, elementary, v[3 + ntags:3 + ntags + nn]))
        if physical:
            tags = ent_phys.setdefault((d, elementary), [])
            if physical not in tags:
                tags.append(physical)
    return {"physical_names": phys, "entity_physical": ent_phys, "nodes": nodes, "elements": elements}


def _identify_points(xyz, tol):
    """class id per point: points closer than ~tol are one node.  Rounding to a grid of pitch tol
    can split two copies of a point that straddle a cell boundary, so two grids offset by half a
    cell are used and the classes are the connected components of 'shares a cell in either grid'
    (distinct GLL nodes are >> tol apart, so nothing else gets merged)."""
    from scipy.sparse import coo_matrix
    from scipy.sparse.csgraph import connected_components
    n = len(xyz)
    _, a = np.unique(np.floor(xyz / tol).astype(np.int64), axis=0, return_inverse=True)
    _, b = np.unique(np.floor(xyz / tol + 0.5).astype(np.int64), axis=0, return_inverse=True)
    a, b = a.ravel(), b.ravel()
    na, nb = a.max() + 1, b.max() + 1
    rows = np.concatenate([np.arange(n), np.arange(n)])
    cols = np.concatenate([n + a, n + na + b])
    g = coo_matrix((np.ones(2 * n, dtype=np.int8), (rows, cols)), shape=(n + na + nb,) * 2)
    _, lab = connected_components(g, directed=False)
    _, inv = np.unique(lab[:n], return_inverse=True)
    return inv.ravel()


class GmshMesh:
    def __init__(self, path, ngl, data=None):
        msh = data if data is not None else read_msh(path)
        el = msh["elements"]
        self.dim = dim = 3 if 5 in el else 2
        if CELL_TYPE[dim] not in el:
            raise ValueError(f"{path}: no quadrangle / hexahedron cells found")
        self.ngl = int(ngl)
        p = self.ngl - 1
        self.nn = self.ngl ** dim
        self.sides = SIDES[dim]
        tags = sorted(msh["nodes"])
        vid = {t: k for k, t in enumerate(tags)}
        vxyz = np.array([msh["nodes"][t][:dim] for t in tags], dtype=np.float64)
        cells = np.array([[vid[t] for t in c[2]] for c in el[CELL_TYPE[dim]]], dtype=np.int64)
        if dim == 3:
            cells = cells[:, GMSH_HEX_TO_CLOSURE]
        self.nel = len(cells)
        self.cell_vertices = cells
        self.corner_xyz = np.ascontiguousarray(vxyz[cells], dtype=np.float64)          # [nel, 2^dim, dim]
        self._check_orientation()
        self.lower, self.upper = vxyz.min(axis=0), vxyz.max(axis=0)

        # GLL nodes of every element, identified by coordinates
        Hc, nodes1d = corner_basis_at_nodes(dim, self.ngl)
        xyz = np.einsum("ac,ecd->ead", Hc, self.corner_xyz).reshape(-1, dim)
        span = np.ptp(vxyz, axis=0).max()
        hmin = self._min_edge()
        tol = 1e-6 * hmin
        inv = _identify_points(xyz, tol)
        # number the nodes in lexicographic order of their coordinates (z, y, x)
        first = np.full(inv.max() + 1, -1, dtype=np.int64)
        first[inv[::-1]] = np.arange(len(inv))[::-1]
        order = np.lexsort(np.round(xyz[first] / tol).astype(np.int64).T)
        rank = np.empty_like(order)
        rank[order] = np.arange(len(order))
        inv = rank[inv]
        first = first[order]
        self.nnodes = len(first)
        self.elem_nodes = np.ascontiguousarray(inv.reshape(self.nel, self.nn), dtype=np.int32)
        self.coords = xyz[first]
        assert span > 0 and len(np.unique(self.elem_nodes[0])) == self.nn, "degenerate element"

        # local nodes on each local face of the reference cell: tensor index 0 or p along an axis
        unit = Hc @ CORNERS[dim].astype(np.float64)
        ref = (1.0 + nodes1d) / 2.0
        self.local_offset = np.abs(unit[:, :, None] - ref[None, None, :]).argmin(axis=2)
        corner_bits = CORNERS[dim]
        self._faces = []                                   # (local corner ids, local node ids)
        for axis in range(dim):
            for side in (0, 1):
                cs = np.nonzero(corner_bits[:, axis] == side)[0]
                ns = np.nonzero(self.local_offset[:, axis] == side * p)[0]
                self._faces.append((cs, ns))

        # boundary elements -> physical tag; cell faces matched by their vertex sets
        face_of = {}
        for e in range(self.nel):
            for f, (cs, _) in enumerate(self._faces):
                face_of.setdefault(frozenset(cells[e, cs].tolist()), []).append((e, f))
        self._exterior = [v[0] for v in face_of.values() if len(v) == 1]
        self._border_faces = {}                            # physical tag -> [(cell, local face)]
        for ed, et, vt in el.get(FACE_TYPE[dim], []):
            hit = face_of.get(frozenset(vid[t] for t in vt))
            if hit is None:
                continue
            for tag in msh["entity_physical"].get((ed, et), []) or [et]:
                self._border_faces.setdefault(tag, []).extend(hit)
        self._phys_by_name = {nm.lower(): tag for (d, tag), nm in msh["physical_names"].items() if d == dim - 1}
        # borders beyond the reference's six (four) box names: named physical groups of boundary faces,
        # e.g. the wall of a body inside the channel (cylinder_mesh.py); in tag order after the box sides
        # (a tag within 1..len(sides) is already a box side by the reference's positional rule, whatever it is
        # called in the file -- src/tests/test.msh names its whole boundary "External" with tag 1 = "down")
        nbox = len(self.sides)
        extra = sorted((tag, nm) for (d, tag), nm in msh["physical_names"].items()
                       if d == dim - 1 and nm.lower() not in self.sides and tag > nbox and tag in self._border_faces)
        self.sides = list(self.sides) + [nm.lower() for _, nm in extra]

    def _min_edge(self):
        c = self.corner_xyz
        d = np.linalg.norm(c[:, :, None, :] - c[:, None, :, :], axis=3)
        d[d == 0] = np.inf
        return float(d.min())

    def _check_orientation(self):
        """closure order must give a positive Jacobian at the cell centre; left-handed cells
        (a file written with the other winding) are mirrored"""
        c = self.corner_xyz
        if self.dim == 2:
            a, b = c[:, 1] - c[:, 0], c[:, 3] - c[:, 0]
            det = a[:, 0] * b[:, 1] - a[:, 1] * b[:, 0]
            flip = [0, 3, 2, 1]
        else:
            # closure order: c0=(0,0,0), c3=(1,0,0), c1=(0,1,0), c4=(0,0,1)
            det = np.einsum("ij,ij->i", np.cross(c[:, 3] - c[:, 0], c[:, 1] - c[:, 0]), c[:, 4] - c[:, 0])
            flip = [4, 7, 6, 5, 0, 3, 2, 1]
        bad = det < 0
        if bad.any():
            self.cell_vertices[bad] = self.cell_vertices[bad][:, flip]
            self.corner_xyz[bad] = self.corner_xyz[bad][:, flip]

    # ---- the L2 queries the builders use (same as BoxMesh) -----------------------------------
    def getDimension(self):
        return self.dim

    def getBordersNames(self):
        return self.sides

    def _nodes_of_faces(self, faces):
        if not faces:
            return np.zeros(0, dtype=np.int64)
        out = [self.elem_nodes[e, self._faces[f][1]] for e, f in faces]
        return np.unique(np.concatenate(out)).astype(np.int64)

    def getBorderNodes(self, name):
        tag = self._phys_by_name.get(name.lower())
        if tag is None:
            tag = self.sides.index(name) + 1                # dmplex.py:167-170
        return self._nodes_of_faces(self._border_faces.get(tag, []))

    def getBordersNodes(self):
        """nodes of the 'External Boundary' label: faces that belong to one cell only (dmplex.py:19-20)"""
        return self._nodes_of_faces(self._exterior)

    def getCellCornersCoords(self, cell):
        return self.corner_xyz[cell].ravel().copy()

    def getGlobalNodesFromCell(self, cell, shared=True):
        return self.elem_nodes[cell].tolist()

    def getNodesCoordinates(self, nodes):
        return self.coords[np.asarray(nodes, dtype=np.int64)]


def write_msh(path, vxyz, cells, boundary, physical_names=None, version=4.1):
    """Gmsh 4.1 ASCII writer for tests and synthetic unstructured inputs.  vxyz [nv, 2|3];
    cells [nel, 4|8] vertex ids in GMSH order; boundary: list of (physical tag, vertex ids)."""
    vxyz = np.asarray(vxyz, dtype=np.float64)
    dim = vxyz.shape[1]
    cells = np.asarray(cells)
    tags = sorted({t for t, _ in boundary})
    if version < 3.0:
        return _write_msh2(path, vxyz, cells, boundary, physical_names or {}, dim)
    with open(path, "w") as f:
        f.write("$MeshFormat\n4.1 0 8\n$EndMeshFormat\n")
        names = physical_names or {}
        f.write(f"$PhysicalNames\n{len(names)}\n")
        for (d, t), nm in names.items():
            f.write(f'{d} {t} "{nm}"\n')
        f.write("$EndPhysicalNames\n")
        # one entity per boundary physical tag (dimension dim-1) + one for the cells
        counts = [0, 0, 0, 0]
        counts[dim - 1] = len(tags)
        counts[dim] = 1
        f.write("$Entities\n" + " ".join(str(c) for c in counts) + "\n")
        for t in tags:
            f.write(f"{t} 0 0 0 1 1 1 1 {t} 0\n")
        f.write("1 0 0 0 1 1 1 1 1000 0\n$EndEntities\n")
        nv = len(vxyz)
        f.write(f"$Nodes\n1 {nv} 1 {nv}\n{dim} 1 0 {nv}\n")
        f.write("\n".join(str(i + 1) for i in range(nv)) + "\n")
        for x in vxyz:
            xx = list(x) + [0.0] * (3 - dim)
            f.write(f"{xx[0]:.17g} {xx[1]:.17g} {xx[2]:.17g}\n")
        f.write("$EndNodes\n")
        nel = len(cells) + len(boundary)
        f.write(f"$Elements\n{len(tags) + 1} {nel} 1 {nel}\n")
        k = 1
        for t in tags:
            faces = [v for tt, v in boundary if tt == t]
            f.write(f"{dim - 1} {t} {FACE_TYPE[dim]} {len(faces)}\n")
            for v in faces:
                f.write(f"{k} " + " ".join(str(int(i) + 1) for i in v) + "\n")
                k += 1
        f.write(f"{dim} 1 {CELL_TYPE[dim]} {len(cells)}\n")
        for c in cells:
            f.write(f"{k} " + " ".join(str(int(i) + 1) for i in c) + "\n")
            k += 1
        f.write("$EndElements\n")


def _write_msh2(path, vxyz, cells, boundary, names, dim):
    """the same mesh as an MSH 2.2 ASCII file (boundary elements first, physical tag = elementary tag)"""
    with open(path, "w") as f:
        f.write("$MeshFormat\n2.2 0 8\n$EndMeshFormat\n")
        f.write(f"$PhysicalNames\n{len(names)}\n")
        for (d, t), nm in names.items():
            f.write(f'{d} {t} "{nm}"\n')
        f.write("$EndPhysicalNames\n")
        f.write(f"$Nodes\n{len(vxyz)}\n")
        for i, x in enumerate(vxyz):
            xx = list(x) + [0.0] * (3 - dim)
            f.write(f"{i + 1} {xx[0]:.17g} {xx[1]:.17g} {xx[2]:.17g}\n")
        f.write("$EndNodes\n")
        f.write(f"$Elements\n{len(cells) + len(boundary)}\n")
        k = 1
        for t, v in boundary:
            f.write(f"{k} {FACE_TYPE[dim]} 2 {t} {t} " + " ".join(str(int(i) + 1) for i in v) + "\n")
            k += 1
        for c in cells:
            f.write(f"{k} {CELL_TYPE[dim]} 2 1000 1 " + " ".join(str(int(i) + 1) for i in c) + "\n")
            k += 1
        f.write("$EndElements\n")
