"""Ray-cast mesh masker: MeshMaskerRay (ray.py:43-77, mesh_boundary_masker.py:157-245).

The reference has no JAX path and no tests for this operator, and its intersection routine lives in warp-lang (absent), so
the oracle (oracle/mesh_masker.py) restates the published watertight ray/triangle algorithm.  It is pinned here against
geometry it must reproduce (analytic sphere shell, exact box faces, exact segment clipping of a convex mesh) and against a
second, independent implementation (FP64 Moeller-Trumbore, oracle/mesh_moller.py) on non-convex meshes, thin plates,
coincident faces and grazing rays; what remains unpinned is agreement with Warp's own BVH query, which cannot run here.
GPU tests compare the CUDA kernel with the oracle bit for bit (masks) / to the last bit (distance weights) on all of
these meshes, and run a stepper on a mesh BC.
"""

import numpy as np
import pytest

from oracle import lattice as olat
from oracle import mesh_masker as omm


def _fields(q, shape):
    return np.zeros((1,) + shape, np.uint8), np.zeros((q,) + shape, bool)


def test_oracle_sphere_shell():
    """Tagged cells lie within one diagonal ray (sqrt 3) of the sphere surface; every cell whose axis ray crosses it is tagged."""
    lat = olat.D3Q27()
    shape, c, r = (22, 20, 21), np.array([10.3, 9.6, 10.1]), 5.2
    bc, ms = _fields(27, shape)
    omm.mesh_masker_ray(lat, omm.icosphere(r, c, 3), 5, bc, ms)
    tag = bc[0] == 5
    ii, jj, kk = np.nonzero(tag)
    d = np.sqrt((ii + 0.5 - c[0]) ** 2 + (jj + 0.5 - c[1]) ** 2 + (kk + 0.5 - c[2]) ** 2)
    assert d.min() > r - np.sqrt(3) - 0.1 and d.max() < r + np.sqrt(3)
    # cells just outside along +x from the centre row: the -x ray (length 1) must hit
    j, k = int(c[1]), int(c[2])
    i_out = int(np.ceil(c[0] + np.sqrt(r**2 - (j + 0.5 - c[1]) ** 2 - (k + 0.5 - c[2]) ** 2) - 0.5))
    assert tag[i_out, j, k]
    l_minus_x = [l for l in range(27) if tuple(lat.c[:, l]) == (-1, 0, 0)][0]
    assert ms[lat.opp_indices[l_minus_x], i_out, j, k]
    assert not ms[:, ~tag].any()  # only tagged cells get missing directions (the mesh is strictly inside the domain)


def test_oracle_box_faces_exact():
    """Axis-aligned box on integer planes: the cells adjacent to a face (inside and outside) are the ones tagged along
    the face normal, and the fractional distance of the normal ray is exactly 0.5."""
    lat = olat.D3Q19()
    shape = (14, 12, 13)
    bc, ms = _fields(19, shape)
    dist = np.zeros((19,) + shape, np.float32)
    omm.mesh_masker_ray(lat, omm.box_mesh((4, 3, 5), (9, 8, 10)), 2, bc, ms, dist)
    tag = bc[0] == 2
    assert tag[3, 5, 7] and tag[4, 5, 7] and not tag[2, 5, 7] and not tag[6, 5, 7]
    lpx = [l for l in range(19) if tuple(lat.c[:, l]) == (1, 0, 0)][0]
    assert ms[lat.opp_indices[lpx], 3, 5, 7] and dist[lpx, 3, 5, 7] == np.float32(0.5)
    lmx = lat.opp_indices[lpx]
    assert ms[lpx, 4, 5, 7] and dist[lmx, 4, 5, 7] == np.float32(0.5)


def test_oracle_out_of_bound_resolution():
    """mesh_boundary_masker.py:157-179: a tagged cell on the domain edge also misses the directions that pull from outside."""
    lat = olat.D3Q19()
    shape = (10, 9, 8)
    bc, ms = _fields(19, shape)
    omm.mesh_masker_ray(lat, omm.box_mesh((0.25, 2, 2), (3, 6, 6)), 4, bc, ms)
    assert bc[0, 0, 4, 4] == 4
    for l in range(1, 19):
        if lat.c[0, l] == 1:  # pulls from x = -1
            assert ms[l, 0, 4, 4]


# ----------------------------------------------------------------------------- oracle vs a second implementation
# the watertight FP32 restatement is pinned against an independent FP64 Moeller-Trumbore link test
# (oracle/mesh_moller.py, which states its edge / vertex / interval conventions) on non-convex meshes, a plate thinner
# than a cell, coincident faces, and rays that graze shared edges and vertices.
def _grazing_box():
    # corners at (k + 0.25) on all three axes: the face-diagonal rays from the cell centres pass exactly through the box's
    # edges, the cube-diagonal rays exactly through its corners (at t = 0.75), and no cell centre lies on a face
    return omm.box_mesh((4.25, 4.25, 4.25), (9.25, 9.25, 9.25))


def _box_faces_on_cell_centres():
    # faces on half-integer planes: axis rays from the cell centres (x + 0.5) run IN the faces' planes and along the box's edges
    return omm.box_mesh((4.5, 3.5, 5.5), (9.5, 8.5, 10.5))


from oracle import mesh_moller as omt  # noqa: E402

PIN_MESHES = [
    # name, lattice, shape, mesh factory, exact (True: coordinates are dyadic rationals, both implementations are exact -> no link may differ)
    ("torus", "D3Q27", (26, 26, 14), lambda: omt.torus(7.3, 2.6, (12.7, 13.1, 6.6)), False),
    ("l_bracket", "D3Q27", (16, 16, 12), lambda: omt.l_bracket((3.25, 4.25, 2.75)), True),
    ("l_bracket_on_cell_centres", "D3Q19", (16, 16, 12), lambda: omt.l_bracket((3.5, 4.5, 2.5), arm=6.0, thick=3.0, height=5.0), True),
    ("touching_boxes", "D3Q27", (20, 12, 12), lambda: omt.touching_boxes((3.25, 2.75, 3.25), (4.5, 5.0, 4.5)), True),
    ("thin_plate", "D3Q27", (12, 14, 14), lambda: omt.thin_plate((5.6, 2.25, 2.75), 8.0, thickness=0.3), True),
    ("thin_plate_between_centres", "D3Q19", (12, 14, 14), lambda: omt.thin_plate((5.625, 2.25, 2.75), 8.0, thickness=0.25), True),
    ("grazing_box", "D3Q27", (15, 15, 15), _grazing_box, True),
    ("box_faces_on_cell_centres", "D3Q27", (16, 14, 15), _box_faces_on_cell_centres, True),
]


@pytest.mark.parametrize("name,lattice,shape,mesh,exact", PIN_MESHES, ids=[m[0] for m in PIN_MESHES])
def test_oracle_agrees_with_moller_trumbore(name, lattice, shape, mesh, exact):
    lat = _lat(lattice)
    verts = mesh()
    bc, ms = _fields(lat.q, shape)
    omm.mesh_masker_ray(lat, verts, 3, bc, ms)
    links = omt.mesh_links(lat, verts, shape)
    any_hit = np.zeros(shape, bool)
    n_hits = n_skipped = 0
    for l, (cells, hit, margin) in links.items():
        got = ms[lat.opp_indices[l]][cells[:, 0], cells[:, 1], cells[:, 2]]
        # exact meshes: every link is compared, grazing ones included (margin 0), except cell centres lying ON a face (margin -1)
        safe = margin >= 0.0 if exact else margin > 1e-4
        bad = np.flatnonzero((got != hit) & safe)
        assert bad.size == 0, f"{name}, direction {l} {tuple(lat.c[:, l])}: {bad.size} links differ, first at cell {cells[bad[0]]} (watertight {got[bad[0]]}, Moeller-Trumbore {hit[bad[0]]}, margin {margin[bad[0]]:.2e})"
        n_hits += int(hit.sum())
        n_skipped += int((~safe).sum())
        any_hit[cells[hit, 0], cells[hit, 1], cells[hit, 2]] = True
    # "on_cell_centres": faces pass through cell centres on purpose; the links that start or end ON a face are outside the
    # contract (see oracle/mesh_moller.py) and only the remaining ones are compared
    degenerate = "on_cell_centres" in name
    assert n_hits > 200 and (degenerate or n_skipped < 0.02 * sum(len(v[0]) for v in links.values()))
    if exact and not degenerate:
        assert np.array_equal(bc[0] == 3, any_hit)  # the tagged set is exactly the set of cells with a crossing link
    # the mesh is strictly inside the domain: nothing outside the candidate box is touched
    assert np.array_equal(bc[0] == 3, ms.any(axis=0))


def test_thin_plate_is_seen_from_both_sides():
    """A plate thinner than a cell between two cell centres: the link across it crosses the surface twice; both cells are
    tagged and miss the direction towards each other (no leak through thin features)."""
    lat = olat.D3Q19()
    shape = (12, 14, 14)
    bc, ms = _fields(19, shape)
    omm.mesh_masker_ray(lat, omt.thin_plate((5.625, 2.25, 2.75), 8.0, thickness=0.25), 6, bc, ms)
    lpx = [l for l in range(19) if tuple(lat.c[:, l]) == (1, 0, 0)][0]
    lmx = lat.opp_indices[lpx]
    assert bc[0, 5, 6, 6] == 6 and bc[0, 6, 6, 6] == 6 and bc[0, 4, 6, 6] == 0 and bc[0, 7, 6, 6] == 0
    assert ms[lmx, 5, 6, 6] and ms[lpx, 6, 6, 6]  # (5,6,6) casts along +x and hits -> missing[opp(+x)]


# ----------------------------------------------------------------------------- CUDA kernel vs oracle
MESHES = [
    ("icosphere_q27", "D3Q27", (24, 20, 22), lambda: omm.icosphere(5.3, (11.2, 9.7, 10.4), 2)),
    ("icosphere_q19", "D3Q19", (20, 21, 19), lambda: omm.icosphere(4.1, (9.4, 10.2, 9.3), 3)),
    ("box_integer_planes", "D3Q27", (16, 14, 15), lambda: omm.box_mesh((4, 3, 5), (9, 8, 10))),
    ("box_half_planes", "D3Q19", (16, 14, 15), lambda: omm.box_mesh((4.5, 3.5, 5.5), (9.5, 8.5, 10.5))),
    ("box_touching_edge", "D3Q19", (10, 9, 8), lambda: omm.box_mesh((0.25, 2, 2), (3, 6, 6))),
] + [(m[0], m[1], m[2], m[3]) for m in PIN_MESHES]  # the meshes the oracle is pinned on: non-convex, thin, coincident faces, grazing rays


def _lat(name):
    return {"D3Q19": olat.D3Q19, "D3Q27": olat.D3Q27}[name]()


@pytest.mark.gpu
@pytest.mark.parametrize("name,lattice,shape,mesh", MESHES, ids=[m[0] for m in MESHES])
def test_cuda_mesh_masker_matches_oracle(name, lattice, shape, mesh):
    import torch

    from tests.util import init_xlb
    from xlb_b200.operator.boundary_condition import HalfwayBounceBackBC
    from xlb_b200.operator.boundary_masker import MeshMaskerRay, MeshVoxelizationMethod

    lat = _lat(lattice)
    verts = mesh()
    bc_o, ms_o = _fields(lat.q, shape)
    bc_o[0, 1, 1, 1] = 9  # pre-existing tags of other BCs stay
    dist_o = np.zeros((lat.q,) + shape, np.float32)
    init_xlb(lattice, "FP32FP32")
    bc = HalfwayBounceBackBC(mesh_vertices=verts.copy(), voxelization_method=MeshVoxelizationMethod("RAY"))
    omm.mesh_masker_ray(lat, verts, bc.id, bc_o, ms_o, dist_o)
    bc_mask = torch.zeros((1,) + shape, dtype=torch.uint8, device="cuda")
    bc_mask[0, 1, 1, 1] = 9
    missing = torch.zeros((lat.q,) + shape, dtype=torch.bool, device="cuda")
    distances = torch.zeros((lat.q,) + shape, dtype=torch.float32, device="cuda")
    bc.needs_mesh_distance = True
    masker = MeshMaskerRay()
    d, b, m = masker(bc, distances, bc_mask, missing)
    assert bc.__dict__.get("mesh_vertices") is None  # consumed, like mesh_boundary_masker.py:203-204
    assert np.array_equal(b.cpu().numpy(), bc_o)
    assert np.array_equal(m.cpu().numpy(), ms_o)
    assert np.array_equal(d.cpu().numpy().view(np.uint32), dist_o.view(np.uint32))


@pytest.mark.gpu
def test_cuda_mesh_masker_slab_equals_whole():
    """Voxelising per x-slab (global triangle coordinates, local block) gives the rows of the whole-domain result."""
    import torch

    from tests.util import init_xlb
    from xlb_b200.operator.boundary_condition import HalfwayBounceBackBC
    from xlb_b200.operator.boundary_masker import MeshMaskerRay

    shape = (24, 20, 22)
    verts = omm.icosphere(5.3, (11.2, 9.7, 10.4), 2)
    init_xlb("D3Q27", "FP32FP32")
    whole_b = torch.zeros((1,) + shape, dtype=torch.uint8, device="cuda")
    whole_m = torch.zeros((27,) + shape, dtype=torch.bool, device="cuda")
    bc = HalfwayBounceBackBC(mesh_vertices=verts.copy())
    MeshMaskerRay()(bc, None, whole_b, whole_m)
    for x0, x1 in [(0, 9), (9, 13), (13, 24)]:
        lb = torch.zeros((1, x1 - x0) + shape[1:], dtype=torch.uint8, device="cuda")
        lm = torch.zeros((27, x1 - x0) + shape[1:], dtype=torch.bool, device="cuda")
        bc2 = HalfwayBounceBackBC(mesh_vertices=verts.copy())
        bc2.id = bc.id
        MeshMaskerRay()(bc2, None, lb, lm, (x0, 0, 0), shape)
        assert torch.equal(lb, whole_b[:, x0:x1]) and torch.equal(lm, whole_m[:, x0:x1])


@pytest.mark.gpu
def test_cuda_mesh_masker_error_paths():
    import torch

    from tests.util import init_xlb
    from xlb_b200.operator.boundary_condition import HalfwayBounceBackBC
    from xlb_b200.operator.boundary_masker import MeshMaskerRay

    init_xlb("D3Q19", "FP32FP32")
    shape = (8, 8, 8)
    bc_mask = torch.zeros((1,) + shape, dtype=torch.uint8, device="cuda")
    missing = torch.zeros((19,) + shape, dtype=torch.bool, device="cuda")
    bc = HalfwayBounceBackBC(mesh_vertices=omm.box_mesh((2, 2, 2), (9, 5, 5)))  # sticks out of the domain
    with pytest.raises(Exception):  # ValueError wrapped by Operator.__call__ like the reference (operator.py:128-133)
        MeshMaskerRay()(bc, None, bc_mask, missing)
    init_xlb("D2Q9", "FP32FP32")
    with pytest.raises(NotImplementedError):
        MeshMaskerRay()


@pytest.mark.gpu
def test_stepper_with_mesh_obstacle_runs_and_matches_generic():
    """windtunnel_3d.py:77-113 in miniature: regularized inlet, extrapolation outlet, full-way walls and a half-way
    bounce-back obstacle given as a MESH.  The masks equal index-masker + oracle ray masker, and the fused kernel
    equals the per-cell generic kernel bitwise over 20 steps (mass stays finite and positive)."""
    import torch

    import xlb_b200 as xlb
    from tests.util import init_xlb
    from xlb_b200 import _lib
    from xlb_b200.compute_backend import ComputeBackend
    from xlb_b200.grid import grid_factory
    from xlb_b200.operator.boundary_condition import ExtrapolationOutflowBC, FullwayBounceBackBC, HalfwayBounceBackBC, RegularizedBC
    from xlb_b200.operator.boundary_condition.boundary_condition_registry import boundary_condition_registry
    from xlb_b200.operator.boundary_masker import MeshVoxelizationMethod
    from xlb_b200.operator.stepper import IncompressibleNavierStokesStepper

    shape = (40, 24, 24)
    verts = omm.icosphere(4.2, (12.3, 11.8, 12.1), 2)
    outs = []
    for variant in (_lib.VARIANT_AUTO, _lib.VARIANT_GENERIC):
        init_xlb("D3Q27", "FP32FP32")
        boundary_condition_registry.next_id = 1
        grid = grid_factory(shape, compute_backend=ComputeBackend.CUDA)
        box = grid.bounding_box_indices()
        box_ne = grid.bounding_box_indices(remove_edges=True)
        walls = [box["bottom"][i] + box["top"][i] + box["front"][i] + box["back"][i] for i in range(3)]
        walls = np.unique(np.array(walls), axis=-1).tolist()
        bcs = [
            FullwayBounceBackBC(indices=walls),
            RegularizedBC("velocity", prescribed_value=(0.03, 0.0, 0.0), indices=box_ne["left"]),
            ExtrapolationOutflowBC(indices=box_ne["right"]),
            HalfwayBounceBackBC(mesh_vertices=verts.copy(), voxelization_method=MeshVoxelizationMethod("RAY")),
        ]
        st = IncompressibleNavierStokesStepper(grid=grid, boundary_conditions=bcs, collision_type="KBC", backend_config={"kernel_variant": variant})
        f_0, f_1, bc_mask, missing = st.prepare_fields()
        if variant == _lib.VARIANT_AUTO:
            lat = olat.D3Q27()
            car = (bc_mask[0] == bcs[3].id).cpu().numpy()
            bc_o, ms_o = _fields(27, shape)
            omm.mesh_masker_ray(lat, verts, bcs[3].id, bc_o, ms_o)
            assert np.array_equal(car, bc_o[0] == bcs[3].id) and car.sum() > 100
            assert np.array_equal(missing.cpu().numpy()[:, car], ms_o[:, car])
        for i in range(20):
            f_0, f_1 = st(f_0, f_1, bc_mask, missing, 1.7, i)
            f_0, f_1 = f_1, f_0
        outs.append(f_0.clone())
        rho = f_0.sum(dim=0)
        assert torch.isfinite(f_0).all() and rho.min() > 0.8 and rho.max() < 1.2
    assert torch.equal(outs[0], outs[1])


# ----------------------------------------------------------------------------- host-side pieces (no GPU)
def test_stl_roundtrip_and_grid_fit(tmp_path):
    """read_stl returns what trimesh.load_mesh(..., process=False).vertices gives the reference's example
    (windtunnel_3d.py:80-81): three consecutive vertex rows per facet, binary or ASCII."""
    from xlb_b200.helper import ellipsoid_soup, fit_to_grid, read_stl, write_stl

    v = ellipsoid_soup(radii=(2.0, 1.0, 0.5), centre=(1.0, -2.0, 3.0), subdivisions=2)
    assert v.shape == (3 * 20 * 4**2, 3) and v.dtype == np.float32
    # closed surface: every undirected edge is shared by exactly two triangles
    tri = np.round(v.reshape(-1, 3, 3).astype(np.float64), 5)
    edges = {}
    for t in tri:
        for a, b in ((0, 1), (1, 2), (2, 0)):
            key = tuple(sorted((tuple(t[a]), tuple(t[b]))))
            edges[key] = edges.get(key, 0) + 1
    assert set(edges.values()) == {2}
    p = str(tmp_path / "a.stl")
    write_stl(p, v)
    assert np.array_equal(read_stl(p), v)
    q = str(tmp_path / "b.stl")
    with open(q, "w") as fh:
        fh.write("solid s\n")
        for t in v.reshape(-1, 3, 3)[:5]:
            fh.write("facet normal 0 0 0\n outer loop\n")
            for a in t:
                fh.write(f"  vertex {float(a[0])!r} {float(a[1])!r} {float(a[2])!r}\n")
            fh.write(" endloop\nendfacet\n")
        fh.write("endsolid s\n")
    assert np.array_equal(read_stl(q), v[:15])
    with open(str(tmp_path / "c.stl"), "w") as fh:
        fh.write("not an stl\n")
    with pytest.raises(ValueError):
        read_stl(str(tmp_path / "c.stl"))
    w, dx, area = fit_to_grid(v, (64, 32, 32))  # windtunnel_3d.py:83-100
    ext = w.max(axis=0) - w.min(axis=0)
    assert abs(ext.max() - 16.0) < 1e-4 and abs(w[:, 0].min() - 16.0) < 1e-4 and abs(w[:, 2].min() - 2.0) < 1e-4
    assert abs((w[:, 1].min() + w[:, 1].max()) / 2 - 16.0) < 1e-4 and abs(area - ext[1] * ext[2]) < 1e-3 and dx > 0


def test_mesh_bc_construction_and_method_registry():
    """Mesh BCs keep their vertices until a masker consumes them; method ids are the reference's
    (mesh_voxelization_method.py:12-17)."""
    import xlb_b200 as xlb
    from xlb_b200.compute_backend import ComputeBackend
    from xlb_b200.operator.boundary_condition import HalfwayBounceBackBC
    from xlb_b200.operator.boundary_masker import MeshVoxelizationMethod
    from xlb_b200.precision_policy import PrecisionPolicy

    pol = PrecisionPolicy.FP32FP32
    vs = xlb.velocity_set.D3Q19(precision_policy=pol, compute_backend=ComputeBackend.CUDA)
    xlb.init(velocity_set=vs, default_backend=ComputeBackend.CUDA, default_precision_policy=pol)
    assert [MeshVoxelizationMethod(n).id for n in ("AABB", "RAY", "AABB_CLOSE", "WINDING")] == [1, 2, 3, 4]
    assert MeshVoxelizationMethod("AABB_CLOSE", close_voxels=3).options == {"close_voxels": 3}
    with pytest.raises(AssertionError):
        MeshVoxelizationMethod("SDF")
    verts = omm.box_mesh((2, 2, 2), (5, 5, 5))
    bc = HalfwayBounceBackBC(mesh_vertices=verts, voxelization_method=MeshVoxelizationMethod("RAY"))
    assert bc.indices is None and bc.mesh_vertices is verts and bc.voxelization_method.name == "RAY" and not bc.needs_mesh_distance


def test_oracle_ray_hits_equal_exact_segment_clipping():
    """Independent geometric pin of the ray masker: for a closed CONVEX mesh, whether the lattice link from a cell centre
    p to p + c_l crosses the surface can be decided without any ray/triangle test, by clipping the segment against the
    face planes (Cyrus-Beck) in FP64: it crosses iff the clipped interval is non-empty and not both end points are inside.
    Every link of every cell must agree with the oracle's watertight FP32 ray caster (no leaks through shared edges or
    vertices, no phantom hits); links whose decision hinges on less than 1e-4 are left out."""
    lat = olat.D3Q27()
    shape = (20, 19, 21)
    centre = np.array([9.63, 9.41, 10.27])
    verts = omm.icosphere(5.17, centre, 2)
    tri = verts.reshape(-1, 3, 3).astype(np.float64)
    normals = np.cross(tri[:, 1] - tri[:, 0], tri[:, 2] - tri[:, 0])
    normals *= (np.sign(np.einsum("fk,fk->f", normals, tri[:, 0] - centre)) / np.linalg.norm(normals, axis=1))[:, None]  # outward, unit
    bc, ms = _fields(27, shape)
    omm.mesh_masker_ray(lat, verts, 7, bc, ms)
    # candidate cells: everything within 2 cells of the surface (all others have no hit and no crossing)
    grid = np.stack(np.meshgrid(*[np.arange(n) for n in shape], indexing="ij"), axis=-1).reshape(-1, 3)
    r = np.linalg.norm(grid + 0.5 - centre, axis=1)
    near = grid[np.abs(r - 5.17) < 2.5]
    assert not ms.reshape(27, -1)[:, np.abs(r - 5.17) >= 2.5].any()
    pts = near + 0.5
    a = np.einsum("fk,nfk->nf", normals, pts[:, None, :] - tri[None, :, 0, :])  # signed height of p above each face
    checked = 0
    for l in range(27):
        c = lat.c[:, l].astype(np.float64)
        if not c.any():
            continue
        b = normals @ c  # change of the height along the link
        with np.errstate(divide="ignore", invalid="ignore"):
            tc = -a / b[None, :]
        t_hi = np.where(b[None, :] > 0, tc, np.inf).min(axis=1)
        t_lo = np.where(b[None, :] < 0, tc, -np.inf).max(axis=1)
        par_out = ((b[None, :] == 0) & (a > 0)).any(axis=1)
        lo_, hi_ = np.maximum(t_lo, 0.0), np.minimum(t_hi, 1.0)
        through = (lo_ <= hi_) & ~par_out
        p_in = np.all(a < 0, axis=1)
        q_in = np.all(a + b[None, :] < 0, axis=1)
        crosses = through & ~(p_in & q_in)
        margin = np.minimum.reduce([np.abs(hi_ - lo_), np.abs(t_lo), np.abs(t_hi), np.abs(t_lo - 1.0), np.abs(t_hi - 1.0), np.min(np.abs(a), axis=1)])
        safe = margin > 1e-4
        got = ms[lat.opp_indices[l]][near[:, 0], near[:, 1], near[:, 2]]
        bad = np.count_nonzero(got[safe] != crosses[safe])
        assert bad == 0, f"direction {l}: {bad} links differ"
        checked += int(np.count_nonzero(crosses & safe))
    assert checked > 3000
    assert np.array_equal(bc[0] == 7, ms.any(axis=0))
