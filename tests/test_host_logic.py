"""Host-side mirror of the reference API: velocity sets, grid helpers, BC descriptors, registry, dispatch.
CPU only -- nothing here launches a kernel."""

import numpy as np
import pytest

import xlb_b200 as xlb
from oracle.lattice import make_lattice
from oracle.masker import bounding_box_indices
from xlb_b200.compute_backend import ComputeBackend
from xlb_b200.precision_policy import Precision, PrecisionPolicy


def _init(name="D3Q19", policy="FP32FP32"):
    pol = PrecisionPolicy[policy]
    vs = getattr(xlb.velocity_set, name)(precision_policy=pol, compute_backend=ComputeBackend.CUDA)
    xlb.init(velocity_set=vs, default_backend=ComputeBackend.CUDA, default_precision_policy=pol)
    return vs


@pytest.mark.parametrize("name", ["D2Q9", "D3Q19", "D3Q27"])
def test_velocity_set_matches_oracle(name):
    vs = _init(name)
    lat = make_lattice(name)
    assert np.array_equal(vs.c, lat.c)
    assert np.allclose(vs.w, lat.w, rtol=0, atol=0)
    assert np.array_equal(vs.opp_indices, lat.opp_indices)
    assert np.array_equal(vs.cc, lat.cc)
    assert np.allclose(vs.qi, lat.qi)
    assert np.array_equal(vs.main_indices, lat.main_indices)
    assert np.array_equal(vs.right_indices, lat.right_indices)
    assert np.array_equal(vs.left_indices, lat.left_indices)
    assert vs.center_index == lat.center_index == 0


def test_d3q19_table_of_survey():
    """explicit ordering and opposite table."""
    vs = _init("D3Q19")
    assert vs.c[:, 1].tolist() == [0, 0, -1] and vs.c[:, 9].tolist() == [-1, 0, 0] and vs.c[:, 18].tolist() == [1, 1, 0]
    assert vs.opp_indices.tolist() == [0, 2, 1, 6, 8, 7, 3, 5, 4, 14, 16, 15, 18, 17, 9, 11, 10, 13, 12]
    assert vs.right_indices.tolist() == [14, 15, 16, 17, 18] and vs.left_indices.tolist() == [9, 10, 11, 12, 13]
    vs27 = _init("D3Q27")
    assert vs27.opp_indices.tolist() == [0, 2, 1, 6, 8, 7, 3, 5, 4, 18, 20, 19, 24, 26, 25, 21, 23, 22, 9, 11, 10, 15, 17, 16, 12, 14, 13]


@pytest.mark.parametrize("shape", [(6, 5), (5, 6, 7)])
@pytest.mark.parametrize("remove_edges", [False, True])
def test_bounding_box_indices(shape, remove_edges):
    _init("D2Q9" if len(shape) == 2 else "D3Q19")
    g = xlb.grid.grid_factory(shape, compute_backend=ComputeBackend.CUDA)
    mine = g.bounding_box_indices(remove_edges=remove_edges)
    ref = bounding_box_indices(shape, remove_edges=remove_edges)
    assert set(mine) == set(ref)
    for k in ref:
        assert np.array_equal(np.array(mine[k]), np.array(ref[k])), k


def test_precision_policy():
    assert PrecisionPolicy.FP32FP16.compute_precision == Precision.FP32
    assert PrecisionPolicy.FP32FP16.store_precision == Precision.FP16
    assert PrecisionPolicy.FP64FP32.compute_precision == Precision.FP64
    assert [p.name for p in PrecisionPolicy] == ["FP64FP64", "FP64FP32", "FP64FP16", "FP32FP32", "FP32FP16"]


def test_other_backends_rejected():
    with pytest.raises(ValueError):
        xlb.init(velocity_set=None, default_backend=ComputeBackend.JAX, default_precision_policy=PrecisionPolicy.FP32FP32)


def test_bc_registry_and_descriptors():
    from xlb_b200.operator.boundary_condition import EquilibriumBC, ExtrapolationOutflowBC, FullwayBounceBackBC, HalfwayBounceBackBC, ZouHeBC
    from xlb_b200.operator.boundary_condition import ImplementationStep
    from xlb_b200.operator.boundary_condition.boundary_condition_registry import boundary_condition_registry as reg
    from oracle import bc as obc

    vs = _init("D3Q27")
    reg.next_id = 1
    box = bounding_box_indices((8, 6, 6), remove_edges=True)
    a = FullwayBounceBackBC(indices=box["bottom"])
    b = EquilibriumBC(rho=1.0, u=(0.02, 0.0, 0.0), indices=box["top"])
    c = HalfwayBounceBackBC(indices=box["front"], prescribed_value=(0.01, 0.0, 0.02))
    d = ZouHeBC("velocity", prescribed_value=(0.04, 0.0, 0.0), indices=box["left"])
    e = ExtrapolationOutflowBC(indices=box["right"])
    assert [x.id for x in (a, b, c, d, e)] == [1, 2, 3, 4, 5]
    assert a.implementation_step == ImplementationStep.COLLISION
    assert all(x.implementation_step == ImplementationStep.STREAMING for x in (b, c, d, e))
    assert c.needs_padding and d.needs_padding and not a.needs_padding
    assert e.normal.tolist() == [1, 0, 0]
    lat = make_lattice("D3Q27")
    ob = obc.EquilibriumBC(lat, "FP32FP32", 1.0, (0.02, 0.0, 0.0), registry=obc.Registry())
    assert np.array_equal(b.feq_constants().astype(np.float32), ob.feq_constants().astype(np.float32))
    oc = obc.HalfwayBounceBackBC(lat, "FP32FP32", prescribed_value=(0.01, 0.0, 0.02), registry=obc.Registry())
    cu = lat.w.astype(np.float32)[:, None] * np.tensordot(lat.c.astype(np.float32), oc.u_wall, axes=(0, 0))
    assert np.allclose(c.moving_wall_table(), (6.0 * cu).reshape(-1), rtol=0, atol=0)
    with pytest.raises(AssertionError):
        ZouHeBC("temperature", prescribed_value=1.0)
    with pytest.raises(ValueError):
        ZouHeBC("velocity", prescribed_value=(0.1, 0, 0), profile=lambda: None)
    # padded index set == oracle's
    c2 = HalfwayBounceBackBC(indices=[[3, 4], [3, 3], [2, 2]])
    oc2 = obc.HalfwayBounceBackBC(lat, "FP32FP32", indices=[[3, 4], [3, 3], [2, 2]], registry=obc.Registry())
    assert np.array_equal(c2.pad_indices(), oc2.pad_indices())


def test_kbc_rejects_d3q19_and_push_rejected():
    from xlb_b200.operator.collision import KBC
    from xlb_b200.operator.stepper import IncompressibleNavierStokesStepper

    _init("D3Q19")
    with pytest.raises(NotImplementedError):
        KBC()
    g = xlb.grid.grid_factory((8, 8, 8), compute_backend=ComputeBackend.CUDA)
    with pytest.raises(NotImplementedError):
        IncompressibleNavierStokesStepper(grid=g, boundary_conditions=[], streaming_scheme="push")
    with pytest.raises(NotImplementedError):
        IncompressibleNavierStokesStepper(grid=g, boundary_conditions=[], collision_type="KBC")


from xlb_b200.operator import Operator  # noqa: E402


class Toy(Operator):
    # module level: the registry key is the first component of __qualname__ (operator.py:74-87)
    @Operator.register_backend(ComputeBackend.CUDA)
    def two(self, a, b):
        return ("two", a, b)

    @Operator.register_backend(ComputeBackend.CUDA)
    def one(self, a):
        return ("one", a)


def test_operator_dispatch_by_signature():
    """operator.py:89-133: the first registered implementation whose signature binds is called."""
    _init("D3Q19")
    t = Toy()
    assert t(1) == ("one", 1)
    assert t(1, 2) == ("two", 1, 2)
    seen = []
    t(5, callback=seen.append)
    assert seen == [("one", 5)]
    with pytest.raises(Exception):
        t(1, 2, 3)


def test_fields_need_a_gpu():
    import torch

    if torch.cuda.is_available():
        pytest.skip("GPU present")
    _init("D3Q19")
    g = xlb.grid.grid_factory((4, 4, 4), compute_backend=ComputeBackend.CUDA)
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        g.create_field(19)
