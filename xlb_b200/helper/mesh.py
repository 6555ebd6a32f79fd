"""Triangle-soup helpers for mesh-based boundary conditions.

The reference's examples read STL files through ``trimesh`` (examples/cfd/windtunnel_3d.py:78-81:
``trimesh.load_mesh(file, process=False).vertices`` -- three consecutive vertex rows per facet) and then scale / shift
the vertices into lattice units (:83-100).  ``read_stl`` returns exactly that array without the trimesh dependency, and
``fit_to_grid`` restates the example's transformation.
"""

import struct

import numpy as np


def read_stl(path):
    """(3 * n_facets, 3) float32 vertices of a binary or ASCII STL file, facet order preserved."""
    with open(path, "rb") as fh:
        data = fh.read()
    if len(data) >= 84:
        (n,) = struct.unpack_from("<I", data, 80)
        if len(data) == 84 + 50 * n:  # binary: 80-byte header, count, 50-byte records (normal, 3 vertices, attribute)
            rec = np.frombuffer(data, dtype=np.dtype([("n", "<f4", (3,)), ("v", "<f4", (3, 3)), ("a", "<u2")]), count=n, offset=84)
            return np.ascontiguousarray(rec["v"].reshape(-1, 3), dtype=np.float32)
    verts = []
    for line in data.decode("ascii", errors="replace").splitlines():
        tok = line.split()
        if len(tok) == 4 and tok[0] == "vertex":
            verts.append([float(tok[1]), float(tok[2]), float(tok[3])])
    if not verts or len(verts) % 3:
        raise ValueError(f"{path}: not a binary or ASCII STL file")
    return np.asarray(verts, dtype=np.float32)


def write_stl(path, vertices):
    """Binary STL of a (3 * n, 3) triangle soup (normals left zero; readers recompute them)."""
    tri = np.asarray(vertices, dtype=np.float32).reshape(-1, 3, 3)
    rec = np.zeros(len(tri), dtype=np.dtype([("n", "<f4", (3,)), ("v", "<f4", (3, 3)), ("a", "<u2")]))
    rec["v"] = tri
    with open(path, "wb") as fh:
        fh.write(b"xlb_b200".ljust(80, b" "))
        fh.write(struct.pack("<I", len(tri)))
        fh.write(rec.tobytes())


def fit_to_grid(vertices, grid_shape, length_fraction=0.25, shift_z=2):
    """windtunnel_3d.py:83-100: move the mesh to the origin, scale its longest extent to ``length_fraction`` of the grid's
    x extent, place it a quarter of the way down the tunnel, centred in y, ``shift_z`` cells above the floor.
    Returns (vertices in lattice units, dx, frontal cross-section in lattice units)."""
    v = np.asarray(vertices, dtype=np.float64).copy()
    v -= v.min(axis=0)
    extents = v.max(axis=0)
    dx = extents.max() / (grid_shape[0] * length_fraction)
    v /= dx
    shift = np.array([grid_shape[0] / 4, (grid_shape[1] - extents[1] / dx) / 2, shift_z])
    return (v + shift).astype(np.float32), dx, float(np.prod(extents[1:]) / dx**2)


def ellipsoid_soup(radii=(1.0, 1.0, 1.0), centre=(0.0, 0.0, 0.0), subdivisions=3):
    """Closed triangle soup of an ellipsoid (subdivided icosahedron pushed onto the surface): a stand-in bluff body for
    the wind-tunnel example when no STL file is at hand."""
    p = (1.0 + 5.0**0.5) / 2.0
    verts = np.array(
        [[-1, p, 0], [1, p, 0], [-1, -p, 0], [1, -p, 0], [0, -1, p], [0, 1, p], [0, -1, -p], [0, 1, -p], [p, 0, -1], [p, 0, 1], [-p, 0, -1], [-p, 0, 1]],
        dtype=np.float64,
    )
    faces = np.array(
        [[0, 11, 5], [0, 5, 1], [0, 1, 7], [0, 7, 10], [0, 10, 11], [1, 5, 9], [5, 11, 4], [11, 10, 2], [10, 7, 6], [7, 1, 8],
         [3, 9, 4], [3, 4, 2], [3, 2, 6], [3, 6, 8], [3, 8, 9], [4, 9, 5], [2, 4, 11], [6, 2, 10], [8, 6, 7], [9, 8, 1]]
    )
    tri = verts[faces]
    for _ in range(subdivisions):
        a, b, c = tri[:, 0], tri[:, 1], tri[:, 2]
        ab, bc, ca = (a + b) / 2, (b + c) / 2, (c + a) / 2
        tri = np.concatenate([np.stack([a, ab, ca], 1), np.stack([b, bc, ab], 1), np.stack([c, ca, bc], 1), np.stack([ab, bc, ca], 1)])
    tri = tri / np.linalg.norm(tri, axis=-1, keepdims=True)
    return (tri * np.asarray(radii, dtype=np.float64) + np.asarray(centre, dtype=np.float64)).reshape(-1, 3).astype(np.float32)
