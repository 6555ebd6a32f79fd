"""The reference's own unit tests (tests/kernels/*, tests/boundary_conditions/* of hsalehipour/XLB, JAX variants),
restated against the NumPy oracle.  They are analytic identities -- the reference stores no golden vectors
 -- and pin the oracle's equilibrium, macroscopic, BGK, stream, EquilibriumBC, FullwayBounceBackBC
and the ``bc_mask`` of IndicesBoundaryMasker.  Shapes / parameters are the reference's where affordable.
"""

import numpy as np
import pytest

from oracle import bc as obc
from oracle import ops
from oracle.lattice import make_lattice
from oracle.masker import indices_boundary_masker

LATS = [("D2Q9", (50, 50)), ("D2Q9", (100, 100)), ("D3Q19", (50, 50, 50)), ("D3Q27", (50, 50, 50))]


@pytest.mark.parametrize("name,shape", LATS)
def test_equilibrium_rest_state(name, shape):
    """tests/kernels/equilibrium/test_equilibrium_jax.py:19-48: rho=1,u=0 -> feq == w, sum == 1."""
    lat = make_lattice(name)
    rho = np.ones((1,) + shape, np.float32)
    u = np.zeros((lat.d,) + shape, np.float32)
    feq = ops.quadratic_equilibrium(lat, rho, u, dtype=np.float32)
    assert feq.dtype == np.float32
    assert np.allclose(np.sum(feq, axis=0), 1.0)
    w = lat.w.reshape((-1,) + (1,) * lat.d)
    assert np.allclose(feq, w)


@pytest.mark.parametrize("name,shape", LATS)
@pytest.mark.parametrize("rho,vel", [(1.0, 0.0), (1.1, 1.0)])
def test_macroscopic_of_equilibrium(name, shape, rho, vel):
    """tests/kernels/macroscopic/test_macroscopic_jax.py:23-50."""
    lat = make_lattice(name)
    r = np.full((1,) + shape, rho, np.float32)
    u = np.full((lat.d,) + shape, vel, np.float32)
    feq = ops.quadratic_equilibrium(lat, r, u, dtype=np.float32)
    rc, uc = ops.macroscopic(lat, feq)
    assert rc.dtype == np.float32 and uc.dtype == np.float32  # JAX promotion int32 x f32 -> f32
    assert np.allclose(rc, rho)
    assert np.allclose(uc, vel, atol=1e-6)


@pytest.mark.parametrize("name,shape", LATS)
@pytest.mark.parametrize("omega", [0.6, 1.0])
def test_bgk_identity(name, shape, omega):
    """tests/kernels/collision/test_bgk_collision_jax.py:20-50."""
    lat = make_lattice(name)
    rho = np.ones((1,) + shape, np.float32)
    u = np.zeros((lat.d,) + shape, np.float32)
    feq = ops.quadratic_equilibrium(lat, rho, u, dtype=np.float32)
    f = np.ones((lat.q,) + shape, np.float32)
    out = ops.bgk(f, feq, omega)
    assert np.allclose(out, f - omega * (f - feq))


@pytest.mark.parametrize("name,shape", LATS)
def test_stream_one_hot_plane(name, shape):
    """tests/kernels/stream/test_stream_jax.py:19-65."""
    lat = make_lattice(name)
    f = np.zeros((lat.q,) + shape, np.float32)
    f[..., shape[-1] // 2] = 1
    out = ops.stream(lat, f)
    for l in range(lat.q):
        exp = np.roll(f[l], tuple(lat.c[:, l]), axis=tuple(range(lat.d)))
        assert np.array_equal(out[l], exp)


def _sphere(shape, d):
    nr = shape[0]
    r = nr // 4
    ax = [np.arange(nr)] * d
    M = np.meshgrid(*ax)
    idx = np.where(sum((m - nr // 2) ** 2 for m in M) < r**2)
    return [list(idx[i]) for i in range(d)]


@pytest.mark.parametrize("name,shape", [("D2Q9", (4, 4)), ("D2Q9", (50, 50)), ("D3Q19", (50, 50, 50)), ("D3Q27", (50, 50, 50))])
def test_indices_masker_bc_mask(name, shape):
    """tests/boundary_conditions/mask/test_bc_indices_masker_jax.py:19-79."""
    lat = make_lattice(name)
    idx = _sphere(shape, lat.d)
    bc = obc.FullwayBounceBackBC(lat, "FP32FP32", indices=idx, registry=obc.Registry())
    bc.id = 5
    bc_mask = np.zeros((1,) + shape, np.uint8)
    missing = np.zeros((lat.q,) + shape, bool)
    bc_mask, missing = indices_boundary_masker(lat, [bc], bc_mask, missing)
    assert bc_mask.dtype == np.uint8 and missing.dtype == bool
    assert bc_mask.shape == (1,) + shape and missing.shape == (lat.q,) + shape
    sel = (0,) + tuple(np.array(i) for i in idx)
    assert np.all(bc_mask[sel] == 5)
    bc_mask[sel] = 0
    assert np.all(bc_mask == 0)


@pytest.mark.parametrize("name,shape", [("D2Q9", (50, 50)), ("D3Q19", (50, 50, 50)), ("D3Q27", (50, 50, 50))])
def test_equilibrium_bc(name, shape):
    """tests/boundary_conditions/bc_equilibrium/test_bc_equilibrium_jax.py:20-90."""
    lat = make_lattice(name)
    idx = _sphere(shape, lat.d)
    reg = obc.Registry()
    bc = obc.EquilibriumBC(lat, "FP32FP32", 1.0, (0.0,) * lat.d, indices=[list(i) for i in idx], registry=reg)
    bc_mask = np.zeros((1,) + shape, np.uint8)
    missing = np.zeros((lat.q,) + shape, bool)
    bc_mask, missing = indices_boundary_masker(lat, [bc], bc_mask, missing)
    f_pre = np.full((lat.q,) + shape, 0.0, np.float32)
    f_post = np.full((lat.q,) + shape, 2.0, np.float32)
    f = bc(f_pre, f_post, bc_mask, missing)
    sel = tuple(np.array(i) for i in idx)
    inside = np.zeros(shape, bool)
    inside[sel] = True
    w = lat.w.astype(np.float32).reshape((-1,) + (1,) * lat.d)
    assert np.allclose(f[:, inside], np.broadcast_to(w, f.shape)[:, inside])
    assert np.all(f[:, ~inside] == 2.0)


@pytest.mark.parametrize("name,shape", [("D2Q9", (50, 50)), ("D3Q19", (50, 50, 50)), ("D3Q27", (50, 50, 50))])
def test_fullway_bc(name, shape):
    """tests/boundary_conditions/bc_fullway_bounce_back/test_bc_fullway_bounce_back_jax.py:20-80."""
    lat = make_lattice(name)
    idx = _sphere(shape, lat.d)
    bc = obc.FullwayBounceBackBC(lat, "FP32FP32", indices=[list(i) for i in idx], registry=obc.Registry())
    bc_mask = np.zeros((1,) + shape, np.uint8)
    missing = np.zeros((lat.q,) + shape, bool)
    bc_mask, missing = indices_boundary_masker(lat, [bc], bc_mask, missing)
    rng = np.random.default_rng(0)
    f_pre = rng.random((lat.q,) + shape, dtype=np.float32)
    f_post = np.full((lat.q,) + shape, 2.0, np.float32)
    f = bc(f_pre, f_post, bc_mask, missing)
    inside = np.zeros(shape, bool)
    inside[tuple(np.array(i) for i in idx)] = True
    assert np.all(f[:, ~inside] == 2.0)
    assert np.array_equal(f[:, inside], f_pre[lat.opp_indices][:, inside])


@pytest.mark.parametrize("policy,tol", [("FP32FP32", 2e-6), ("FP64FP64", 1e-13)])
def test_torch_cpu_transcription_matches_oracle(policy, tol):
    """oracle/torch_cpu.py (the multi-threaded CPU baseline of bench.py) == the NumPy oracle on the benchmark cavity."""
    import torch

    from oracle import cases
    from oracle.torch_cpu import TorchCavityStepper, to_torch

    case = cases.cavity3d(12, policy=policy, omega=1.7)
    ost, _ = cases.build_oracle(case)
    f0, f1, bc, mm = ost.prepare_fields()
    f0 = cases.random_state(case, seed=4)
    ts = TorchCavityStepper(ost)
    a, b = f0, None
    ta, bcm = to_torch(f0), to_torch(bc)
    for i in range(5):
        _, b = ost(a, None, bc, mm, case["omega"], i)
        _, tb = ts(ta, None, bcm, None, case["omega"], i)
        a, ta = b, tb
    err = np.max(np.abs(ta.numpy().astype(np.float64) - a.astype(np.float64))) / np.max(np.abs(a))
    assert err <= tol
