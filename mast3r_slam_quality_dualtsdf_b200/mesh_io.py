"""Binary PLY writer for `TSDFVolume.extract_mesh` output.

The reference writes point-cloud PLYs through `plyfile`
(mast3r_slam/evaluate.py:88-106, :173-187: little-endian records, `vertex`
element with float x/y/z); it has no mesh output.  This writer lays a mesh out
the same way by hand -- `vertex` (float x, y, z) plus `face` (uchar count +
3 int indices) -- and welds the triangle soup so shared vertices are stored
once.  Pure numpy: it runs on the host after one D2H copy of the soup.
"""
import numpy as np


def weld(tri_soup, decimals=7):
    """(T,3,3) or (3T,3) float32 soup -> (verts (V,3) f32, faces (T,3) i32).
    Marching-cubes vertices on a shared edge are bit-identical (same float32
    operation order), so welding is an exact `unique`."""
    v = np.ascontiguousarray(np.asarray(tri_soup, np.float32).reshape(-1, 3))
    if v.shape[0] == 0:
        return v, np.zeros((0, 3), np.int32)
    key = v.view(np.dtype((np.void, 12))).reshape(-1)
    _, first, inv = np.unique(key, return_index=True, return_inverse=True)
    return v[first], inv.reshape(-1, 3).astype(np.int32)


def save_mesh_ply(path, tri_soup, weld_vertices=True, faces=None):
    """Write the mesh; returns (n_vertices, n_faces).  `tri_soup` is a triangle
    soup (welded here on the host) or, with `faces` (T,3) int32, the vertex
    array of an already indexed mesh (`TSDFVolume.extract_mesh_indexed`: the
    weld ran on the device)."""
    if hasattr(tri_soup, "detach"):
        tri_soup = tri_soup.detach().cpu().numpy()
    if faces is not None:
        if hasattr(faces, "detach"):
            faces = faces.detach().cpu().numpy()
        verts = np.ascontiguousarray(np.asarray(tri_soup, np.float32).reshape(-1, 3))
        faces = np.ascontiguousarray(np.asarray(faces, np.int32).reshape(-1, 3))
    elif weld_vertices:
        verts, faces = weld(tri_soup)
    else:
        verts = np.ascontiguousarray(np.asarray(tri_soup, np.float32).reshape(-1, 3))
        faces = np.arange(verts.shape[0], dtype=np.int32).reshape(-1, 3)
    header = ("ply\nformat binary_little_endian 1.0\n"
              f"element vertex {verts.shape[0]}\nproperty float x\nproperty float y\n"
              f"property float z\nelement face {faces.shape[0]}\n"
              "property list uchar int vertex_indices\nend_header\n")
    rec = np.empty(faces.shape[0], dtype=[("n", "u1"), ("idx", "<i4", (3,))])
    rec["n"] = 3
    rec["idx"] = faces
    with open(path, "wb") as f:
        f.write(header.encode("ascii"))
        f.write(verts.astype("<f4").tobytes())
        f.write(rec.tobytes())
    return int(verts.shape[0]), int(faces.shape[0])


def load_mesh_ply(path):
    """Reader for the files `save_mesh_ply` writes (tests, round trips)."""
    with open(path, "rb") as f:
        data = f.read()
    end = data.index(b"end_header\n") + len(b"end_header\n")
    head = data[:end].decode("ascii").split("\n")
    nv = int([h for h in head if h.startswith("element vertex")][0].split()[-1])
    nf = int([h for h in head if h.startswith("element face")][0].split()[-1])
    verts = np.frombuffer(data, "<f4", nv * 3, end).reshape(nv, 3)
    rec = np.frombuffer(data, np.dtype([("n", "u1"), ("idx", "<i4", (3,))]), nf, end + nv * 12)
    return verts.copy(), rec["idx"].copy()
