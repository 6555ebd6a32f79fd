"""Second, independent implementation of the lattice-link / surface test behind the ray-cast mesh masker (TEST
INFRASTRUCTURE -- see oracle/__init__.py).  It exists to PIN ``oracle/mesh_masker.py`` (and through it the CUDA kernel):
the reference delegates the query to ``warp.mesh_query_ray`` (ray.py:59-63), which is absent here, so the restated
watertight test (Woop, Benthin, Wald 2013; FP32) is cross-checked against a different algorithm in a different precision.

Algorithm: Moeller & Trumbore, "Fast, Minimum Storage Ray/Triangle Intersection" (JGT 2(1), 1997), FP64, two-sided,
brute force.  For a ray ``o + t d`` (``d`` = unit lattice direction, ``0 <= t < |c_l|``, ray.py:57-63) and a triangle
``(a, b, c)``:

    e1 = b - a, e2 = c - a, p = d x e2, det = e1 . p          det == 0  -> parallel: no hit
    s = o - a, u = (s . p) / det, q = s x e1, v = (d . q) / det, t = (e2 . q) / det
    hit  <=>  u >= 0 and v >= 0 and u + v <= 1 and 0 <= t < max_t
(evaluated with d = c_l, so t runs over [0, 1), and on the numerators u det, v det, t det -- see ``link_hits``).

Edge and vertex handling, stated because it is the contract the comparison relies on: the barycentric bounds are
INCLUSIVE, so a ray through a shared edge or vertex hits every triangle that shares it (the link counts as crossing the
surface; a watertight caster reports at least one of them, the masker only needs "any hit").  A ray that lies IN a
triangle's plane is parallel to it and misses that triangle; it still hits the neighbouring faces it pierces.  ``t == 0``
(cell centre on the surface) hits here, ``t == max_t`` (surface exactly at the neighbouring cell centre) does not -- the
half-open interval of ray.py:63 (``t < max_length``).  These two cases -- a cell centre lying exactly ON a face -- are NOT
part of the cross-check: an FP32 caster (the oracle's, and Warp's) normalises the diagonal directions, and whether its t
comes out as +0 or -5e-8, as sqrt(2) or one ulp below, is rounding; such links are reported with margin -1 and skipped.
"""

import numpy as np


def link_hits(origins, c, tri):
    """(n,) bool: does the lattice link from each origin (n,3) to origin + ``c`` (``c`` = the integer lattice vector, i.e.
    the ray of ray.py:57-63 parametrised over [0, 1) instead of [0, |c|)) cross ANY triangle (m,3,3); and (n,) float64
    ``margin``: how far the closest call among the triangles is from flipping the decision (barycentric / parametric
    units) -- comparisons across precisions skip links with a tiny margin.  The decision itself is taken on the
    NUMERATORS (u det, v det, t det against det), without a division or a normalised direction, so that for meshes and
    cell centres with dyadic-rational coordinates every product and sum is exact in FP64 and the decision is exact."""
    o = np.asarray(origins, dtype=np.float64)[:, None, :]
    d = np.asarray(c, dtype=np.float64)
    a = tri[None, :, 0, :].astype(np.float64)
    e1 = tri[None, :, 1, :].astype(np.float64) - a
    e2 = tri[None, :, 2, :].astype(np.float64) - a
    p = np.cross(d, e2)
    det = np.einsum("nmk,nmk->nm", e1, p)
    s = o - a
    q = np.cross(s, e1)
    sgn = np.sign(det)
    nu = np.einsum("nmk,nmk->nm", s, p) * sgn  # u |det|
    nv = np.einsum("k,nmk->nm", d, q) * sgn    # v |det|
    nt = np.einsum("nmk,nmk->nm", e2, q) * sgn  # t |det|
    ad = np.abs(det)
    par = det == 0.0
    inside = (nu >= 0.0) & (nv >= 0.0) & (nu + nv <= ad)
    in_t = (nt >= 0.0) & (nt < ad)
    hit = ~par & inside & in_t
    # distance of every triangle's decision from its boundary: the smallest |quantity| among those that decide it
    with np.errstate(divide="ignore", invalid="ignore"):
        u, v, t = nu / ad, nv / ad, nt / ad
        bary = np.minimum(np.minimum(u, v), 1.0 - u - v)  # >= 0 inside
        para = np.minimum(t, 1.0 - t)                     # >= 0 inside the interval
        near_b = np.where(para > -1e-3, np.abs(bary), np.inf)   # a barycentric sign matters only if t is (almost) in range
        near_t = np.where(bary > -1e-3, np.abs(para), np.inf)   # the interval matters only if the point is (almost) inside
        m = np.minimum(near_b, near_t)
        m = np.where(par | ~np.isfinite(m), np.inf, m)
        # a cell centre that lies ON a triangle (the link's start, t == 0, or its end, t == max_t, exactly) is outside the
        # contract: in FP32 which side of the bound t lands on is decided by the rounding of the normalised direction (the
        # reference's Warp query has the same ambiguity); mark such links
        m = np.where(~par & inside & ((nt == 0.0) | (nt == ad)), -1.0, m)
    return hit.any(axis=1), m.min(axis=1)


def mesh_links(lat, vertices, shape, chunk=2048):
    """For every lattice direction l != rest: (cells (n,3) int, hit (n,) bool, margin (n,) float64) over the cells of the
    mesh's bounding box grown by one cell -- the same candidate set the masker walks."""
    tri = np.asarray(vertices, dtype=np.float64).reshape(-1, 3, 3)
    nx, ny, nz = shape
    lo = np.maximum(np.floor(tri.reshape(-1, 3).min(axis=0)).astype(int) - 1, 0)
    hi = np.minimum(np.ceil(tri.reshape(-1, 3).max(axis=0)).astype(int) + 1, [nx, ny, nz])
    g = np.meshgrid(np.arange(lo[0], hi[0]), np.arange(lo[1], hi[1]), np.arange(lo[2], hi[2]), indexing="ij")
    cells = np.stack([x.ravel() for x in g], axis=1)
    centres = cells + 0.5
    out = {}
    for l in range(lat.q):
        c = lat.c[:, l].astype(np.float64)
        if not c.any():
            continue
        hits, margins = [], []
        for s in range(0, len(cells), chunk):
            h, m = link_hits(centres[s : s + chunk], c, tri)
            hits.append(h)
            margins.append(m)
        out[l] = (cells, np.concatenate(hits), np.concatenate(margins))
    return out


# ----------------------------------------------------------------------------- non-convex / degenerate test meshes
def torus(R, r, centre, n_major=24, n_minor=12):
    """(3*m, 3) float32 triangle soup of a torus around the z axis (non-convex, genus 1)."""
    th = np.linspace(0.0, 2.0 * np.pi, n_major, endpoint=False)
    ph = np.linspace(0.0, 2.0 * np.pi, n_minor, endpoint=False)
    T, P = np.meshgrid(th, ph, indexing="ij")
    pts = np.stack([(R + r * np.cos(P)) * np.cos(T), (R + r * np.cos(P)) * np.sin(T), r * np.sin(P)], axis=-1) + np.asarray(centre, dtype=np.float64)
    tris = []
    for i in range(n_major):
        for j in range(n_minor):
            a, b = pts[i, j], pts[(i + 1) % n_major, j]
            c, d = pts[(i + 1) % n_major, (j + 1) % n_minor], pts[i, (j + 1) % n_minor]
            tris += [a, b, c, a, c, d]
    return np.array(tris).astype(np.float32)


def extruded_polygon(poly_xy, z0, z1):
    """Closed prism over a simple polygon given as a list of (x, y) corners (ear-free fan is NOT used: the caps are
    triangulated per axis-aligned rectangle strip, so the polygon must be rectilinear with corners listed counter-clockwise)."""
    poly = [tuple(map(float, p)) for p in poly_xy]
    n = len(poly)
    tris = []
    for i in range(n):  # side walls
        (x0, y0), (x1, y1) = poly[i], poly[(i + 1) % n]
        tris += [(x0, y0, z0), (x1, y1, z0), (x1, y1, z1), (x0, y0, z0), (x1, y1, z1), (x0, y0, z1)]
    xs = sorted({p[0] for p in poly})
    ys = sorted({p[1] for p in poly})

    def inside(px, py):  # even-odd rule
        c = False
        for i in range(n):
            (x0, y0), (x1, y1) = poly[i], poly[(i + 1) % n]
            if (y0 > py) != (y1 > py) and px < (x1 - x0) * (py - y0) / (y1 - y0) + x0:
                c = not c
        return c

    for xa, xb in zip(xs, xs[1:]):  # caps: one rectangle per occupied cell of the corner grid
        for ya, yb in zip(ys, ys[1:]):
            if inside(0.5 * (xa + xb), 0.5 * (ya + yb)):
                for z in (z0, z1):
                    tris += [(xa, ya, z), (xb, ya, z), (xb, yb, z), (xa, ya, z), (xb, yb, z), (xa, yb, z)]
    return np.array(tris, dtype=np.float32)


def l_bracket(origin, arm=6.0, thick=2.5, height=4.0):
    """Non-convex L-shaped prism (re-entrant edge along z)."""
    ox, oy, oz = origin
    poly = [(ox, oy), (ox + arm, oy), (ox + arm, oy + thick), (ox + thick, oy + thick), (ox + thick, oy + arm), (ox, oy + arm)]
    return extruded_polygon(poly, oz, oz + height)


def touching_boxes(lo, size):
    """Two boxes sharing a face (the shared face is present twice, once per box: coincident triangles inside the solid)."""
    from .mesh_masker import box_mesh

    lo = np.asarray(lo, dtype=np.float64)
    size = np.asarray(size, dtype=np.float64)
    a = box_mesh(lo, lo + size)
    b = box_mesh(lo + [size[0], 0, 0], lo + [2 * size[0], size[1], size[2]])
    return np.concatenate([a, b]).astype(np.float32)


def thin_plate(lo, extent, thickness=0.3):
    """A plate thinner than one cell (both faces fall between two neighbouring cell centres)."""
    from .mesh_masker import box_mesh

    lo = np.asarray(lo, dtype=np.float64)
    return box_mesh(lo, lo + [thickness, extent, extent])
