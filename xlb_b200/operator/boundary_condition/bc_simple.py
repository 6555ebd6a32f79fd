"""FullwayBounceBackBC, HalfwayBounceBackBC, EquilibriumBC, DoNothingBC (bc_fullway_bounce_back.py,
bc_halfway_bounce_back.py, bc_equilibrium.py, bc_do_nothing.py of the reference)."""

import numpy as np

from ... import _lib
from .boundary_condition import BoundaryCondition, ImplementationStep


class FullwayBounceBackBC(BoundaryCondition):
    """COLLISION step: f_post[l] = f_post_stream[opp l] on tagged cells (bc_fullway_bounce_back.py:50-56)."""

    abi_kind = _lib.BC_FULLWAY

    def __init__(self, velocity_set=None, precision_policy=None, compute_backend=None, indices=None, mesh_vertices=None, voxelization_method=None):
        super().__init__(ImplementationStep.COLLISION, velocity_set, precision_policy, compute_backend, indices, mesh_vertices, voxelization_method)


class DoNothingBC(BoundaryCondition):
    """STREAMING step: f_post = f_pre on tagged cells (bc_do_nothing.py:50-54)."""

    abi_kind = _lib.BC_DONOTHING

    def __init__(self, velocity_set=None, precision_policy=None, compute_backend=None, indices=None, mesh_vertices=None, voxelization_method=None):
        super().__init__(ImplementationStep.STREAMING, velocity_set, precision_policy, compute_backend, indices, mesh_vertices, voxelization_method)


class HalfwayBounceBackBC(BoundaryCondition):
    """STREAMING step: f_post[l] = f_pre[opp l] (+ 6 w_l c_l.u_wall) where missing[l] (bc_halfway_bounce_back.py:116-134).

    ``prescribed_value`` is a constant wall velocity (moving wall).  A callable ``profile`` must return that
    constant ``(d,)`` / ``(d, 1)`` array (the reference's JAX path only supports a space-independent value too)."""

    abi_kind = _lib.BC_HALFWAY

    def __init__(
        self, velocity_set=None, precision_policy=None, compute_backend=None, indices=None, mesh_vertices=None, voxelization_method=None,
        profile=None, prescribed_value=None,
    ):
        super().__init__(ImplementationStep.STREAMING, velocity_set, precision_policy, compute_backend, indices, mesh_vertices, voxelization_method)
        self.needs_padding = True
        self.profile = profile
        self.needs_moving_wall_treatment = (profile is not None) or (prescribed_value is not None)
        if prescribed_value is None and profile is None:
            print(f"WARNING! Assuming no-slip condition for BC type = {self.__class__.__name__}!")
            prescribed_value = [0] * self.velocity_set.d
        if prescribed_value is not None:
            if profile is not None:
                raise ValueError("Cannot specify both profile and prescribed_value")
            if not isinstance(prescribed_value, (tuple, list, np.ndarray)):
                raise ValueError("Velocity prescribed_value must be a tuple, list, or array")
            value = np.asarray(prescribed_value, dtype=np.float64)
            self.profile = lambda: value.astype(self.precision_policy.store_precision.np_dtype).reshape(-1, 1)

    def moving_wall_table(self):
        """6 * w_l * (c_l . u_wall) with the dtype promotions of the JAX code (bc_halfway_bounce_back.py:123-128)."""
        vs = self.velocity_set
        if not self.needs_moving_wall_treatment:
            return np.zeros(vs.q)
        u_wall = np.asarray(self.profile())
        if u_wall.size != vs.d:
            raise NotImplementedError("HalfwayBounceBackBC: only a space-independent wall velocity is supported")
        u_wall = u_wall.reshape(-1, 1)
        cdt = self.precision_policy.compute_precision.np_dtype
        cu = vs._w.astype(cdt)[:, None] * np.tensordot(vs._c.astype(u_wall.dtype), u_wall, axes=(0, 0))
        return (6.0 * cu).reshape(-1).astype(np.float64)

    def bc_table_row(self, row, field_shape=None):
        super().bc_table_row(row, field_shape)
        row["table"][: self.velocity_set.q] = self.moving_wall_table()


class EquilibriumBC(BoundaryCondition):
    """STREAMING step: f_post = feq(rho, u) constants on tagged cells (bc_equilibrium.py:41-80)."""

    abi_kind = _lib.BC_EQUILIBRIUM

    def __init__(
        self, rho, u, equilibrium_operator=None, velocity_set=None, precision_policy=None, compute_backend=None, indices=None,
        mesh_vertices=None, voxelization_method=None,
    ):
        from ..equilibrium import Equilibrium, QuadraticEquilibrium

        self.rho = rho
        self.u = u
        self.equilibrium_operator = QuadraticEquilibrium() if equilibrium_operator is None else equilibrium_operator
        if not issubclass(type(self.equilibrium_operator), Equilibrium):
            raise ValueError("Equilibrium operator must be a subclass of Equilibrium")
        super().__init__(ImplementationStep.STREAMING, velocity_set, precision_policy, compute_backend, indices, mesh_vertices, voxelization_method)

    def feq_constants(self):
        """The q constants, evaluated once on the host in the compute dtype with the operation order of
        quadratic_equilibrium.py:25-30 (what bc_equilibrium.py:74-76 computes under jit)."""
        vs = self.velocity_set
        cdt = self.precision_policy.compute_precision.np_dtype
        rho = np.array([self.rho], dtype=cdt)
        u = np.array(self.u, dtype=cdt)
        cu = 3.0 * np.tensordot(vs._c.astype(cdt), u, axes=(0, 0))
        usqr = 1.5 * np.sum(np.square(u), axis=0, keepdims=True)
        return (rho * vs._w.astype(cdt) * (1.0 + cu * (1.0 + 0.5 * cu) - usqr)).astype(np.float64)

    def bc_table_row(self, row, field_shape=None):
        super().bc_table_row(row, field_shape)
        row["table"][: self.velocity_set.q] = self.feq_constants()
