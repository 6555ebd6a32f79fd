"""Neutral descriptions of the parity / benchmark cases (TEST INFRASTRUCTURE).

A case is plain data (lattice name, policy, shape, collision, omega, ordered BC specs) so that
the same description can be turned into oracle objects (``build_oracle`` below) and into the
product's objects (``tests/util.py``).  The shapes and BC lists follow the reference scripts:

* ``cavity3d``  examples/performance/mlups_3d.py:194-204, 222   (BASELINE configs 1, 2, 4)
* ``cavity2d``  examples/cfd/lid_driven_cavity_2d.py:44-55
* ``sphere3d``  examples/cfd/flow_past_sphere_3d.py:38-113 with Zou-He instead of Regularized
                (BASELINE configs 3, 5)
"""

import numpy as np

from . import bc as obc
from .lattice import make_lattice
from .masker import bounding_box_indices
from .stepper import Stepper


def _union(faces, names, d):
    idx = [sum((faces[n][i] for n in names), []) for i in range(d)]
    return np.unique(np.array(idx), axis=-1).tolist()


def _force_tag(force):
    return "" if force is None else "_F" + "_".join(f"{v:g}" for v in force)


def cavity3d(n, policy="FP32FP32", lattice="D3Q19", collision="BGK", omega=1.0, shape=None, force=None):
    shape = tuple(shape) if shape else (n, n, n)
    box = bounding_box_indices(shape)
    box_ne = bounding_box_indices(shape, remove_edges=True)
    lid = box_ne["top"]
    walls = _union(box, ["bottom", "left", "right", "front", "back"], 3)
    return dict(
        name=f"cavity3d_{'x'.join(map(str, shape))}_{lattice}_{collision}_{policy}_w{omega}{_force_tag(force)}",
        lattice=lattice,
        policy=policy,
        shape=shape,
        collision=collision,
        omega=omega,
        force=force,
        bcs=[
            dict(kind="EquilibriumBC", rho=1.0, u=(0.02, 0.0, 0.0), indices=lid),
            dict(kind="FullwayBounceBackBC", indices=walls),
        ],
    )


def cavity2d(n, policy="FP32FP32", collision="BGK", omega=1.6, u_lid=0.05, shape=None, force=None):
    shape = tuple(shape) if shape else (n, n)
    box = bounding_box_indices(shape)
    box_ne = bounding_box_indices(shape, remove_edges=True)
    lid = box_ne["top"]
    walls = _union(box, ["bottom", "left", "right"], 2)
    return dict(
        name=f"cavity2d_{'x'.join(map(str, shape))}_{collision}_{policy}_w{omega}{_force_tag(force)}",
        lattice="D2Q9",
        policy=policy,
        shape=shape,
        collision=collision,
        omega=omega,
        force=force,
        bcs=[
            dict(kind="HalfwayBounceBackBC", indices=walls),
            dict(kind="EquilibriumBC", rho=1.0, u=(u_lid, 0.0), indices=lid),
        ],
    )


def sphere_indices(shape, radius=None, centre=None):
    """flow_past_sphere_3d.py:45-54."""
    radius = shape[1] // 12 if radius is None else radius
    centre = (shape[0] // 6, shape[1] // 2, shape[2] // 2) if centre is None else centre
    X, Y, Z = np.meshgrid(*(np.arange(s) for s in shape), indexing="ij")
    idx = np.where((X - centre[0]) ** 2 + (Y - centre[1]) ** 2 + (Z - centre[2]) ** 2 < radius**2)
    return [idx[i].tolist() for i in range(3)]


def poiseuille_profile(shape, u_max, dtype):
    """flow_past_sphere_3d.py:58-81 (bc_profile_jax) -> array (3, ny, nz)."""
    H_y = float(shape[1] - 1)
    H_z = float(shape[2] - 1)
    Y, Z = np.meshgrid(np.arange(shape[1]), np.arange(shape[2]), indexing="ij")
    y_c = Y - (H_y / 2.0)
    z_c = Z - (H_z / 2.0)
    r2 = (2.0 * y_c / H_y) ** 2.0 + (2.0 * z_c / H_z) ** 2.0
    u_x = u_max * np.maximum(0.0, 1.0 - r2)
    return np.stack([u_x, np.zeros_like(u_x), np.zeros_like(u_x)]).astype(dtype)


def sphere3d(
    shape,
    policy="FP32FP32",
    lattice="D3Q27",
    collision="KBC",
    omega=1.6,
    u_max=0.04,
    inlet="velocity",
    outlet="extrapolation",
    profile=False,
    radius=None,
    moving_sphere=None,
    force=None,
):
    shape = tuple(shape)
    box = bounding_box_indices(shape)
    box_ne = bounding_box_indices(shape, remove_edges=True)
    walls = _union(box, ["bottom", "top", "front", "back"], 3)
    if radius is None:
        radius = max(shape[1] // 12, 2)
    sph = sphere_indices(shape, radius=radius)
    if inlet == "velocity":
        if profile:
            from .ops import policy_dtypes
            _, sdt = policy_dtypes(policy)
            inlet_bc = dict(kind="ZouHeBC", bc_type="velocity", profile_array=poiseuille_profile(shape, u_max, sdt), indices=box_ne["left"])
        else:
            inlet_bc = dict(kind="ZouHeBC", bc_type="velocity", prescribed_value=(u_max, 0.0, 0.0), indices=box_ne["left"])
    elif inlet == "pressure":
        inlet_bc = dict(kind="ZouHeBC", bc_type="pressure", prescribed_value=1.01, indices=box_ne["left"])
    elif inlet == "regularized_velocity":  # what examples/cfd/flow_past_sphere_3d.py:107 ships
        if profile:
            from .ops import policy_dtypes
            _, sdt = policy_dtypes(policy)
            inlet_bc = dict(kind="RegularizedBC", bc_type="velocity", profile_array=poiseuille_profile(shape, u_max, sdt), indices=box_ne["left"])
        else:
            inlet_bc = dict(kind="RegularizedBC", bc_type="velocity", prescribed_value=(u_max, 0.0, 0.0), indices=box_ne["left"])
    elif inlet == "regularized_pressure":
        inlet_bc = dict(kind="RegularizedBC", bc_type="pressure", prescribed_value=1.01, indices=box_ne["left"])
    else:
        raise ValueError(inlet)
    if outlet == "extrapolation":
        outlet_bc = dict(kind="ExtrapolationOutflowBC", indices=box_ne["right"])
    elif outlet == "pressure":
        outlet_bc = dict(kind="ZouHeBC", bc_type="pressure", prescribed_value=1.0, indices=box_ne["right"])
    elif outlet == "donothing":
        outlet_bc = dict(kind="DoNothingBC", indices=box_ne["right"])
    else:
        raise ValueError(outlet)
    sphere_bc = dict(kind="HalfwayBounceBackBC", indices=sph)
    if moving_sphere is not None:
        sphere_bc["prescribed_value"] = tuple(moving_sphere)
    return dict(
        name=f"sphere3d_{'x'.join(map(str, shape))}_{lattice}_{collision}_{policy}_{inlet}_{outlet}{'_prof' if profile else ''}{'_moving' if moving_sphere is not None else ''}_w{omega}{_force_tag(force)}",
        lattice=lattice,
        policy=policy,
        shape=shape,
        collision=collision,
        omega=omega,
        force=force,
        bcs=[dict(kind="FullwayBounceBackBC", indices=walls), inlet_bc, outlet_bc, sphere_bc],
    )


def periodic(shape, policy="FP32FP32", lattice="D3Q19", collision="BGK", omega=1.7, force=None):
    return dict(
        name=f"periodic_{'x'.join(map(str, shape))}_{lattice}_{collision}_{policy}_w{omega}{_force_tag(force)}",
        lattice=lattice,
        policy=policy,
        shape=tuple(shape),
        collision=collision,
        omega=omega,
        force=force,
        bcs=[],
    )


def build_oracle(case):
    """Turn a case description into (Stepper, [oracle BC objects]) with ids 1.. in list order."""
    lat = make_lattice(case["lattice"])
    reg = obc.Registry()
    bcs = []
    for spec in case["bcs"]:
        kind = spec["kind"]
        idx = [list(c) for c in spec["indices"]]
        pol = case["policy"]
        if kind == "EquilibriumBC":
            b = obc.EquilibriumBC(lat, pol, spec["rho"], spec["u"], indices=idx, registry=reg)
        elif kind == "FullwayBounceBackBC":
            b = obc.FullwayBounceBackBC(lat, pol, indices=idx, registry=reg)
        elif kind == "HalfwayBounceBackBC":
            b = obc.HalfwayBounceBackBC(lat, pol, indices=idx, prescribed_value=spec.get("prescribed_value"), registry=reg)
        elif kind == "DoNothingBC":
            b = obc.DoNothingBC(lat, pol, indices=idx, registry=reg)
        elif kind == "ExtrapolationOutflowBC":
            b = obc.ExtrapolationOutflowBC(lat, pol, indices=idx, registry=reg)
        elif kind in ("ZouHeBC", "RegularizedBC"):
            cls = obc.ZouHeBC if kind == "ZouHeBC" else obc.RegularizedBC
            if "profile_array" in spec:
                arr = spec["profile_array"]
                b = cls(lat, pol, spec["bc_type"], profile=lambda arr=arr: arr, indices=idx, registry=reg)
            else:
                b = cls(lat, pol, spec["bc_type"], prescribed_value=spec["prescribed_value"], indices=idx, registry=reg)
        else:
            raise ValueError(kind)
        bcs.append(b)
    st = Stepper(lat, case["policy"], case["shape"], bcs, case["collision"], force_vector=case.get("force"))
    return st, bcs


def random_state(case, seed=0, amp=0.01):
    """f = w_l * (1 + amp * U(-1, 1)), numpy default_rng(seed), in the store dtype."""
    from .ops import policy_dtypes

    lat = make_lattice(case["lattice"])
    _, sdt = policy_dtypes(case["policy"])
    rng = np.random.default_rng(seed)
    r = rng.uniform(-1.0, 1.0, size=(lat.q,) + tuple(case["shape"]))
    w = lat.w.reshape((-1,) + (1,) * lat.d)
    return (w * (1.0 + amp * r)).astype(sdt)
