"""Oracle of the ray-cast mesh masker (TEST INFRASTRUCTURE -- see oracle/__init__.py).

Restates ``MeshMaskerRay`` (xlb/operator/boundary_masker/ray.py:43-77) and the shared post-pass
``resolve_out_of_bound_kernel`` (mesh_boundary_masker.py:157-179): for every cell centre ``x + 0.5`` and every
lattice direction ``l != centre`` a ray of length ``|c_l|`` is cast along ``c_l``; on a hit the cell is tagged with the
BC id, ``missing_mask[opp l]`` is set and (optionally) the fractional distance ``|hit - centre| / |c_l|`` is stored.

PARITY UNPINNED.  The reference has no JAX path for this operator and delegates the intersection to
``warp.mesh_query_ray`` (third-party ``warp-lang >= 1.10.0``, setup.py:96; not vendored, not installable offline).
What is restated here is the published algorithm behind that call -- Woop, Benthin, Wald, "Watertight Ray/Triangle
Intersection", JCGT 2(1), 2013 -- in FP32 with the FP64 fallback for zero edge functions, two-sided, accepting the
closest hit with ``0 <= t < max_t`` (brute force over triangles instead of a BVH, which changes no result).  The CUDA
kernel is compared bit-exactly with this restatement; neither is checked against Warp itself.
"""

import numpy as np

F = np.float32


def _max_dim(a):
    """index of the first largest component (warp/native/intersect.h ``max_dim``)."""
    x = max(a[0], a[1], a[2])
    return 0 if x == a[0] else (1 if x == a[1] else 2)


def ray_hits(origins, direction, max_t, tri):
    """Watertight ray/triangle test of every origin (n,3) against every triangle (m,3,3), all FP32.
    Returns (hit bool (n,m), t (n,m), u (n,m), v (n,m))."""
    d = direction.astype(F)
    kz = _max_dim(np.abs(d))
    kx = (kz + 1) % 3
    ky = (kx + 1) % 3
    if d[kz] < 0:
        kx, ky = ky, kx
    sx = F(d[kx] / d[kz])
    sy = F(d[ky] / d[kz])
    sz = F(F(1.0) / d[kz])
    o = origins.astype(F)[:, None, :]
    a = tri[None, :, 0, :] - o
    b = tri[None, :, 1, :] - o
    c = tri[None, :, 2, :] - o
    ax = a[..., kx] - sx * a[..., kz]
    ay = a[..., ky] - sy * a[..., kz]
    bx = b[..., kx] - sx * b[..., kz]
    by = b[..., ky] - sy * b[..., kz]
    cx = c[..., kx] - sx * c[..., kz]
    cy = c[..., ky] - sy * c[..., kz]
    u = cx * by - cy * bx
    v = ax * cy - ay * cx
    w = bx * ay - by * ax
    zero = (u == 0) | (v == 0) | (w == 0)
    if np.any(zero):  # FP64 fallback of the edge functions (Woop et al., section 3.3)
        D = np.float64
        u64 = (cx.astype(D) * by.astype(D) - cy.astype(D) * bx.astype(D)).astype(F)
        v64 = (ax.astype(D) * cy.astype(D) - ay.astype(D) * cx.astype(D)).astype(F)
        w64 = (bx.astype(D) * ay.astype(D) - by.astype(D) * ax.astype(D)).astype(F)
        u = np.where(zero, u64, u)
        v = np.where(zero, v64, v)
        w = np.where(zero, w64, w)
    miss = ((u < 0) | (v < 0) | (w < 0)) & ((u > 0) | (v > 0) | (w > 0))
    det = (u + v) + w
    miss |= det == 0
    az = sz * a[..., kz]
    bz = sz * b[..., kz]
    cz = sz * c[..., kz]
    t_scaled = (u * az + v * bz) + w * cz
    with np.errstate(divide="ignore", invalid="ignore"):
        rcp = F(1.0) / det
        t = t_scaled * rcp
        uu = u * rcp
        vv = v * rcp
    hit = ~miss & (t >= 0) & (t < F(max_t))
    return hit, t, uu, vv


def _closest_hit_weight(hit, t, u, v, tri, centres, max_len):
    """ray.py:72-76: position of the hit from the barycentrics (wp.mesh_eval_position), its distance to the cell centre
    divided by |c_l|.  The closest hit wins; equal-t ties between triangles sharing an edge go to the smaller weight."""
    with np.errstate(invalid="ignore", over="ignore"):
        pos = u[..., None] * tri[None, :, 0, :] + v[..., None] * tri[None, :, 1, :] + ((F(1.0) - u) - v)[..., None] * tri[None, :, 2, :]
        dv = pos - centres[:, None, :]
        dist = np.sqrt((dv[..., 0] * dv[..., 0] + dv[..., 1] * dv[..., 1]) + dv[..., 2] * dv[..., 2], dtype=F)
        wgt = np.where(hit, dist / max_len, F(np.inf)).astype(F)
    tt = np.where(hit, t, F(np.inf))
    tmin = tt.min(axis=1, keepdims=True)
    return np.where(tt == tmin, wgt, F(np.inf)).min(axis=1)


def mesh_masker_ray(lat, vertices, bc_id, bc_mask, missing_mask, distances=None, chunk=4096):
    """ray.py:43-77 + mesh_boundary_masker.py:157-179.  ``vertices`` (3*m, 3) triangle soup in lattice units;
    ``bc_mask`` uint8 [1,nx,ny,nz], ``missing_mask`` bool [q,nx,ny,nz] and ``distances`` float32 [q,nx,ny,nz] (or None)
    are updated in place and returned."""
    assert lat.d == 3
    tri = np.asarray(vertices, dtype=F).reshape(-1, 3, 3)
    nx, ny, nz = bc_mask.shape[1:]
    lo = np.maximum(np.floor(tri.reshape(-1, 3).min(axis=0)).astype(int) - 1, 0)
    hi = np.minimum(np.ceil(tri.reshape(-1, 3).max(axis=0)).astype(int) + 1, [nx, ny, nz])
    gx, gy, gz = np.meshgrid(np.arange(lo[0], hi[0]), np.arange(lo[1], hi[1]), np.arange(lo[2], hi[2]), indexing="ij")
    cells = np.stack([gx.ravel(), gy.ravel(), gz.ravel()], axis=1)
    centres = cells.astype(F) + F(0.5)  # helper_functions_masker.py:47-52
    for l in range(lat.q):
        cl = lat.c[:, l].astype(F)
        if not np.any(cl):
            continue
        max_len = np.sqrt(F(cl @ cl), dtype=F)
        direction = (cl / max_len).astype(F)
        for s in range(0, len(cells), chunk):
            cs, ce = cells[s : s + chunk], centres[s : s + chunk]
            hit, t, u, v = ray_hits(ce, direction, max_len, tri)
            anyhit = hit.any(axis=1)
            if not anyhit.any():
                continue
            ix, iy, iz = cs[anyhit].T
            bc_mask[0, ix, iy, iz] = bc_id
            missing_mask[lat.opp_indices[l], ix, iy, iz] = True
            if distances is not None:
                best = _closest_hit_weight(hit, t, u, v, tri, ce, max_len)
                distances[l, ix, iy, iz] = best[anyhit]
    # resolve_out_of_bound_kernel: tagged cells also miss every direction that pulls from outside the domain
    tagged = bc_mask[0] == bc_id
    ii, jj, kk = np.nonzero(tagged)
    for l in range(1, lat.q):
        px, py, pz = ii - lat.c[0, l], jj - lat.c[1, l], kk - lat.c[2, l]
        oob = (px < 0) | (px >= nx) | (py < 0) | (py >= ny) | (pz < 0) | (pz >= nz)
        missing_mask[l, ii[oob], jj[oob], kk[oob]] = True
    return bc_mask, missing_mask, distances


# ----------------------------------------------------------------------------- test meshes (triangle soups)
def icosphere(radius, centre, subdivisions=2):
    """(3*m, 3) float32 triangle soup of an icosphere."""
    p = (1.0 + 5.0**0.5) / 2.0
    v = np.array(
        [[-1, p, 0], [1, p, 0], [-1, -p, 0], [1, -p, 0], [0, -1, p], [0, 1, p], [0, -1, -p], [0, 1, -p], [p, 0, -1], [p, 0, 1], [-p, 0, -1], [-p, 0, 1]],
        dtype=np.float64,
    )
    f = [(0, 11, 5), (0, 5, 1), (0, 1, 7), (0, 7, 10), (0, 10, 11), (1, 5, 9), (5, 11, 4), (11, 10, 2), (10, 7, 6), (7, 1, 8),
         (3, 9, 4), (3, 4, 2), (3, 2, 6), (3, 6, 8), (3, 8, 9), (4, 9, 5), (2, 4, 11), (6, 2, 10), (8, 6, 7), (9, 8, 1)]
    tris = v[np.array(f)]
    for _ in range(subdivisions):
        a, b, c = tris[:, 0], tris[:, 1], tris[:, 2]
        ab, bc, ca = (a + b) / 2, (b + c) / 2, (c + a) / 2
        tris = np.concatenate([np.stack([a, ab, ca], 1), np.stack([b, bc, ab], 1), np.stack([c, ca, bc], 1), np.stack([ab, bc, ca], 1)])
    tris = tris / np.linalg.norm(tris, axis=-1, keepdims=True)
    return (tris * radius + np.asarray(centre, dtype=np.float64)).reshape(-1, 3).astype(F)


def box_mesh(lo, hi):
    """(36, 3) float32 triangle soup of an axis-aligned box (12 large triangles; exercises the exact-zero fallbacks)."""
    lo, hi = np.asarray(lo, dtype=np.float64), np.asarray(hi, dtype=np.float64)
    c = np.array([[lo[0] if not (i & 1) else hi[0], lo[1] if not (i & 2) else hi[1], lo[2] if not (i & 4) else hi[2]] for i in range(8)])
    quads = [(0, 1, 3, 2), (4, 6, 7, 5), (0, 4, 5, 1), (2, 3, 7, 6), (0, 2, 6, 4), (1, 5, 7, 3)]
    tris = []
    for q in quads:
        tris += [c[q[0]], c[q[1]], c[q[2]], c[q[0]], c[q[2]], c[q[3]]]
    return np.array(tris).astype(F)
