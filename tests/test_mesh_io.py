"""CPU: PLY mesh writer round trip on a CPU marching-cubes mesh (oracle)."""
import numpy as np

from mast3r_slam_quality_dualtsdf_b200 import mesh_io
from tests.test_mc_oracle import _sphere_grid
from oracle import oracle as orc


def test_ply_round_trip_and_weld(tmp_path):
    keys, t, _c, _r, v = _sphere_grid()
    tri = orc.marching_cubes(keys, t, np.ones(len(t), np.float32), v)
    assert len(tri) > 100
    p = tmp_path / "m.ply"
    nv, nf = mesh_io.save_mesh_ply(str(p), tri)
    assert nf == len(tri) and nv < 3 * nf  # shared vertices are stored once
    # closed surface: Euler characteristic 2 (V - E + F with E = 3F/2)
    assert nv - 3 * nf // 2 + nf == 2
    verts, faces = mesh_io.load_mesh_ply(str(p))
    assert verts.shape == (nv, 3) and faces.shape == (nf, 3)
    assert np.array_equal(verts[faces].reshape(-1, 3, 3), np.asarray(tri, np.float32).reshape(-1, 3, 3))
    nv2, nf2 = mesh_io.save_mesh_ply(str(p), tri, weld_vertices=False)
    assert (nv2, nf2) == (3 * nf, nf)
    e, f = mesh_io.weld(np.zeros((0, 3, 3), np.float32))
    assert e.shape == (0, 3) and f.shape == (0, 3)


def _parse_ply_generic(path):
    """An independent PLY reader for this test: walks the header grammar of the PLY
    format (elements, scalar and list properties, binary_little_endian) the way
    `plyfile` -- the library the reference writes its PLYs with (evaluate.py:88-106) --
    does, without assuming the layout `mesh_io.save_mesh_ply` emits."""
    import struct

    fmt = {"char": "b", "uchar": "B", "short": "h", "ushort": "H", "int": "i", "uint": "I",
           "float": "f", "double": "d", "int8": "b", "uint8": "B", "int16": "h", "uint16": "H",
           "int32": "i", "uint32": "I", "float32": "f", "float64": "d"}
    data = open(path, "rb").read()
    assert data.startswith(b"ply\n")
    end = data.index(b"end_header\n") + len(b"end_header\n")
    lines = data[:end].decode("ascii").strip().split("\n")
    assert lines[1].split() == ["format", "binary_little_endian", "1.0"]
    elements = []
    for ln in lines[2:-1]:
        tok = ln.split()
        if tok[0] == "comment":
            continue
        if tok[0] == "element":
            elements.append((tok[1], int(tok[2]), []))
        elif tok[0] == "property":
            if tok[1] == "list":
                elements[-1][2].append(("list", fmt[tok[2]], fmt[tok[3]], tok[4]))
            else:
                elements[-1][2].append(("scalar", fmt[tok[1]], None, tok[2]))
        else:
            raise AssertionError("unknown header line: " + ln)
    out, off = {}, end
    for name, count, props in elements:
        rows = []
        for _ in range(count):
            row = {}
            for kind, a, b, pname in props:
                if kind == "scalar":
                    (row[pname],) = struct.unpack_from("<" + a, data, off)
                    off += struct.calcsize("<" + a)
                else:
                    (n,) = struct.unpack_from("<" + a, data, off)
                    off += struct.calcsize("<" + a)
                    row[pname] = struct.unpack_from("<%d%s" % (n, b), data, off)
                    off += struct.calcsize("<%d%s" % (n, b))
            rows.append(row)
        out[name] = rows
    assert off == len(data), "trailing bytes after the last element"
    return out


def test_ply_file_is_valid_for_a_generic_reader(tmp_path):
    """the mesh PLY is checked by a reader that did not write it --
    vertex element with float x / y / z first (the reference's point-cloud layout,
    evaluate.py:92-99), then faces as `list uchar int vertex_indices`; every index in
    range, the triangles equal to the marching-cubes soup they came from."""
    keys, t, _c, _r, v = _sphere_grid(n=14, r=0.15)
    tri = np.asarray(orc.marching_cubes(keys, t, np.ones(len(t), np.float32), v), np.float32)
    p = tmp_path / "g.ply"
    nv, nf = mesh_io.save_mesh_ply(str(p), tri)
    ply = _parse_ply_generic(str(p))
    assert list(ply) == ["vertex", "face"] and len(ply["vertex"]) == nv and len(ply["face"]) == nf
    assert list(ply["vertex"][0]) == ["x", "y", "z"]
    verts = np.array([[r["x"], r["y"], r["z"]] for r in ply["vertex"]], np.float32)
    faces = np.array([r["vertex_indices"] for r in ply["face"]], np.int64)
    assert faces.shape == (nf, 3) and faces.min() >= 0 and faces.max() < nv
    assert np.array_equal(verts[faces], tri.reshape(-1, 3, 3))
    assert len(np.unique(verts, axis=0)) == nv  # welded: no vertex stored twice
    # an indexed mesh handed over as (verts, faces) is written as is
    nv2, nf2 = mesh_io.save_mesh_ply(str(p), verts, faces=faces.astype(np.int32))
    ply2 = _parse_ply_generic(str(p))
    assert (nv2, nf2) == (nv, nf) and ply2 == ply
