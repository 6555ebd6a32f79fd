"""Oracle stepper (TEST INFRASTRUCTURE -- see oracle/__init__.py).

Restates ``IncompressibleNavierStokesStepper.prepare_fields`` / ``jax_implementation_pull``
(xlb/operator/stepper/nse_stepper.py:99-148, 238-282) with NumPy whole-array operations in
the same order.  Every cell is updated (the JAX path has no BC_SOLID early-out).
"""

import numpy as np

from . import ops
from .bc import STREAMING, COLLISION
from .masker import indices_boundary_masker


class Stepper:
    def __init__(self, lat, policy, shape, boundary_conditions=(), collision_type="BGK", force_vector=None):
        self.lat = lat
        self.policy = policy
        self.cdt, self.sdt = ops.policy_dtypes(policy)
        self.shape = tuple(shape)
        self.boundary_conditions = list(boundary_conditions)
        self.collision_type = collision_type
        # nse_stepper.py:86-87; the examples pass jnp.array(force): default float dtype == compute dtype of the policy
        self.force_vector = None if force_vector is None else np.asarray(force_vector, dtype=self.cdt)
        if collision_type == "KBC" and not ((lat.d, lat.q) in [(2, 9), (3, 27)]):
            raise NotImplementedError("Velocity set not supported: {}".format(lat))  # kbc.py:65-66

    def prepare_fields(self, initializer=None):
        """nse_stepper.py:99-148 (+ helper/nse_fields.py:45-55, helper/initializers.py:55-72)."""
        lat = self.lat
        f_0 = np.zeros((lat.q,) + self.shape, dtype=self.sdt)
        f_1 = f_0.copy()  # :121-122 -> zeros
        bc_mask = np.zeros((1,) + self.shape, dtype=np.uint8)
        missing_mask = np.zeros((lat.q,) + self.shape, dtype=bool)
        with_idx = [bc for bc in self.boundary_conditions if bc.indices is not None]
        if with_idx:
            bc_mask, missing_mask = indices_boundary_masker(lat, with_idx, bc_mask, missing_mask)
        if initializer is not None:
            f_0 = initializer(bc_mask, f_0)
        else:
            rho = np.full((1,) + self.shape, 1.0, dtype=self.cdt)
            u = np.full((lat.d,) + self.shape, 0.0, dtype=self.cdt)
            f_0 = ops.quadratic_equilibrium(lat, rho, u, dtype=self.cdt)  # not cast to store (JAX returns compute dtype)
        return f_0, f_1, bc_mask, missing_mask

    def __call__(self, f_0, f_1, bc_mask, missing_mask, omega, timestep=0):
        """nse_stepper.py:238-282."""
        lat = self.lat
        f_0 = np.asarray(f_0, dtype=self.cdt)  # :243
        f_post_stream = ops.stream(lat, f_0)  # :247
        for bc in self.boundary_conditions:  # :250-257
            if bc.implementation_step == STREAMING:
                f_post_stream = bc(f_0, f_post_stream, bc_mask, missing_mask)
        rho, u = ops.macroscopic(lat, f_post_stream)  # :260
        feq = ops.quadratic_equilibrium(lat, rho, u, dtype=self.cdt)  # :263
        f_post_collision = ops.collide(lat, self.collision_type, f_post_stream, feq, omega, self.force_vector, dtype=self.cdt)  # :266
        for bc in self.boundary_conditions:  # :269-277
            f_post_collision = bc.assemble_auxiliary_data(f_post_stream, f_post_collision, bc_mask, missing_mask)
            if bc.implementation_step == COLLISION:
                f_post_collision = bc(f_post_stream, f_post_collision, bc_mask, missing_mask)
        f_1 = f_post_collision.astype(self.sdt)  # :280
        return f_0, f_1

    def run(self, f_0, f_1, bc_mask, missing_mask, omega, num_steps):
        """The caller's loop with buffer swap (examples/performance/mlups_3d.py:227-229).
        Returns the latest populations (what the caller holds in ``f_0`` after the swap)."""
        for i in range(num_steps):
            f_0, f_1 = self(f_0, f_1, bc_mask, missing_mask, omega, i)
            f_0, f_1 = f_1, f_0
        return f_0, f_1
