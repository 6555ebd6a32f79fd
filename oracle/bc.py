"""Oracle boundary conditions (TEST INFRASTRUCTURE -- see oracle/__init__.py).

Whole-grid ``np.where(bc_mask == id, ...)`` selects, exactly like the JAX branches of
xlb/operator/boundary_condition/*.py.  Objects are plain descriptors; ids are assigned
in construction order starting at 1 like boundary_condition_registry.py:6-30, but per
``Registry`` instance so that tests are independent.
"""

from collections import Counter

import numpy as np

from . import ops

COLLISION = "COLLISION"  # boundary_condition.py:26-30
STREAMING = "STREAMING"


class Registry:
    """boundary_condition_registry.py:6-30 (ids start at 1; 0 = no BC)."""

    def __init__(self):
        self.next_id = 1

    def register(self):
        i = self.next_id
        self.next_id += 1
        return i


_default_registry = Registry()


def _bmask(lat, bc_mask, id_):
    """``lax.broadcast_in_dim(bc_mask == id, (q, ...))`` e.g. bc_fullway_bounce_back.py:52-55."""
    boundary = bc_mask == id_
    return np.broadcast_to(boundary, (lat.q,) + boundary.shape[1:])


class BoundaryCondition:
    """boundary_condition.py:58-144 (host-side descriptor part)."""

    needs_padding = False

    def __init__(self, lat, policy, step, indices=None, registry=None):
        self.id = (registry or _default_registry).register()
        self.lat = lat
        self.policy = policy
        self.cdt, self.sdt = ops.policy_dtypes(policy)
        self.implementation_step = step
        self.indices = indices

    def pad_indices(self):
        """boundary_condition.py:123-136."""
        bc_indices = np.array(self.indices)
        if self.needs_padding:
            padded = bc_indices[:, :, None] + self.lat.c[:, None, :]
            return np.unique(padded.reshape(self.lat.d, -1), axis=1)
        return bc_indices

    def assemble_auxiliary_data(self, f_pre, f_post, bc_mask, missing_mask):
        """boundary_condition.py:138-144 (identity unless overridden)."""
        return f_post

    def __call__(self, f_pre, f_post, bc_mask, missing_mask):
        raise NotImplementedError


class FullwayBounceBackBC(BoundaryCondition):
    """bc_fullway_bounce_back.py:30-56 (COLLISION step)."""

    def __init__(self, lat, policy, indices=None, registry=None):
        super().__init__(lat, policy, COLLISION, indices, registry)

    def __call__(self, f_pre, f_post, bc_mask, missing_mask):
        boundary = _bmask(self.lat, bc_mask, self.id)
        return np.where(boundary, f_pre[self.lat.opp_indices, ...], f_post)


class HalfwayBounceBackBC(BoundaryCondition):
    """bc_halfway_bounce_back.py:37-134 (STREAMING step; constant wall velocity only)."""

    needs_padding = True  # :60

    def __init__(self, lat, policy, indices=None, prescribed_value=None, registry=None):
        super().__init__(lat, policy, STREAMING, indices, registry)
        self.needs_moving_wall_treatment = prescribed_value is not None  # :66-70
        if prescribed_value is None:
            prescribed_value = [0] * lat.d  # :73-75
        pv = np.asarray(prescribed_value, dtype=np.float64)  # :83-86
        # prescribed_profile_jax :97-100 -> array in the *store* dtype, shape (d, 1)
        self.u_wall = pv.astype(self.sdt).reshape(-1, 1)

    def __call__(self, f_pre, f_post, bc_mask, missing_mask):
        lat = self.lat
        boundary = _bmask(lat, bc_mask, self.id)
        moving_wall_component = 0.0
        if self.needs_moving_wall_treatment:  # :123-128
            u_wall = self.u_wall
            # int32 (x) store dtype -> store dtype; then w (compute dtype) promotes
            cu = lat.w.astype(self.cdt)[:, None] * np.tensordot(lat.c.astype(u_wall.dtype), u_wall, axes=(0, 0))
            cu = cu.reshape((-1,) + (1,) * (f_post.ndim - 1))
            moving_wall_component = 6.0 * cu
        # :131
        return np.where(np.logical_and(missing_mask, boundary), f_pre[lat.opp_indices] + moving_wall_component, f_post)


class EquilibriumBC(BoundaryCondition):
    """bc_equilibrium.py:41-80 (STREAMING step)."""

    def __init__(self, lat, policy, rho, u, indices=None, registry=None):
        super().__init__(lat, policy, STREAMING, indices, registry)
        self.rho = rho
        self.u = u

    def feq_constants(self):
        """bc_equilibrium.py:74-76: feq of jnp.array([rho]), jnp.array(u) in JAX's default float
        dtype (float32, or float64 once jax_enable_x64 is on, i.e. for the FP64 policies)."""
        dflt = np.float64 if self.cdt == np.float64 else np.float32
        rho = np.array([self.rho], dtype=dflt)
        u = np.array(self.u, dtype=dflt)
        return ops.quadratic_equilibrium(self.lat, rho, u, dtype=self.cdt)

    def __call__(self, f_pre, f_post, bc_mask, missing_mask):
        feq = self.feq_constants()
        feq = feq.reshape(feq.shape + (1,) * self.lat.d)
        boundary = bc_mask == self.id
        return np.where(boundary, feq, f_post)


class DoNothingBC(BoundaryCondition):
    """bc_do_nothing.py:30-54 (STREAMING step)."""

    def __init__(self, lat, policy, indices=None, registry=None):
        super().__init__(lat, policy, STREAMING, indices, registry)

    def __call__(self, f_pre, f_post, bc_mask, missing_mask):
        boundary = bc_mask == self.id
        return np.where(boundary, f_pre, f_post)


class ZouHeBC(BoundaryCondition):
    """bc_zouhe.py:48-304 (STREAMING step)."""

    needs_padding = True  # :146

    def __init__(self, lat, policy, bc_type, prescribed_value=None, profile=None, indices=None, registry=None):
        assert bc_type in ["velocity", "pressure"]  # :62
        super().__init__(lat, policy, STREAMING, indices, registry)
        self.bc_type = bc_type
        if prescribed_value is not None:
            if profile is not None:
                raise ValueError("Cannot specify both profile and prescribed_value")  # :84-85
            if bc_type == "velocity":
                prescribed_value = np.asarray(prescribed_value, dtype=np.float64)  # :89-90
            else:
                prescribed_value = float(prescribed_value)  # :94-95
            # prescribed_profile_jax :156-157 -> store dtype, shape (-1, 1)
            self.prescribed_values = np.array(prescribed_value, dtype=self.sdt).reshape(-1, 1)
        else:
            # user profile callable returns the full array (bc_zouhe.py:119-120)
            self.prescribed_values = np.asarray(profile())

    # :165-169
    def _get_known_middle_mask(self, missing_mask):
        known_mask = missing_mask[self.lat.opp_indices]
        middle_mask = ~(missing_mask | known_mask)
        return known_mask, middle_mask

    # :171-176  (int32 result)
    def _get_normal_vec(self, missing_mask):
        main_c = self.lat.c[:, self.lat.main_indices].astype(np.int32)
        m = missing_mask[self.lat.main_indices].astype(np.int32)
        return -np.tensordot(main_c, m, axes=(-1, 0))

    # :178-215
    def _broadcast_prescribed_values(self, pv, target_shape):
        shp = pv.shape
        if len(shp) > len(target_shape):
            raise ValueError("prescribed_values has more dimensions than target_shape")
        if len(shp) < len(target_shape):
            n = len(target_shape) - len(shp)
            if len(shp) == 0:
                shp = (1,) * len(target_shape)
            else:
                shp = (shp[0],) + (1,) * n + tuple(shp[1:])
            pv = pv.reshape(shp)
        for a, b in zip(shp, target_shape):
            if not (a == 1 or a == b):
                raise ValueError(f"Cannot broadcast dimension {a} to {b}")
        return np.broadcast_to(pv, target_shape)

    def _fsum(self, fpop, missing_mask):
        known_mask, middle_mask = self._get_known_middle_mask(missing_mask)
        return np.sum(fpop * middle_mask, axis=0, keepdims=True) + 2.0 * np.sum(fpop * known_mask, axis=0, keepdims=True)

    # :256-266
    def calculate_rho(self, fpop, vel, missing_mask):
        normals = self._get_normal_vec(missing_mask)
        # int32 (x) float -> float dtype of vel (JAX); force it for NumPy
        unormal = np.sum(normals.astype(vel.dtype) * vel, keepdims=True, axis=0)
        fsum = self._fsum(fpop, missing_mask)
        return fsum / (1.0 + unormal)

    # :241-254
    def calculate_vel(self, fpop, rho, missing_mask):
        normals = self._get_normal_vec(missing_mask)
        fsum = self._fsum(fpop, missing_mask)
        unormal = -1.0 + fsum / rho
        return unormal * normals.astype(unormal.dtype)

    # :268-277 with get_rho :217-227, get_vel :229-239
    def calculate_equilibrium(self, f_post, missing_mask):
        lat = self.lat
        if self.bc_type == "velocity":
            target = (lat.d,) + f_post.shape[1:]
            vel = self._broadcast_prescribed_values(self.prescribed_values, target)
            rho = self.calculate_rho(f_post, vel, missing_mask)
        else:
            rho = self.prescribed_values
            vel = self.calculate_vel(f_post, rho, missing_mask)
        return ops.quadratic_equilibrium(lat, rho, vel, dtype=self.cdt)

    # :279-288
    def bounceback_nonequilibrium(self, fpop, feq, missing_mask):
        opp = self.lat.opp_indices
        fknown = fpop[opp] + feq - feq[opp]
        return np.where(missing_mask, fknown, fpop)

    # :290-304
    def __call__(self, f_pre, f_post, bc_mask, missing_mask):
        boundary = _bmask(self.lat, bc_mask, self.id)
        feq = self.calculate_equilibrium(f_post, missing_mask)
        f_post_bd = self.bounceback_nonequilibrium(f_post, feq, missing_mask)
        return np.where(boundary, f_post_bd, f_post).astype(f_post.dtype, copy=False)


class RegularizedBC(ZouHeBC):
    """bc_regularized.py:52-135: Zou-He followed by Latt's regularisation of the boundary populations."""

    # :78-114
    def regularize_fpop(self, fpop, feq):
        lat = self.lat
        weights = lat.w.astype(self.cdt).reshape((-1,) + (1,) * lat.d)  # velocity_set.w is in the compute dtype
        Qi = lat.qi.astype(self.cdt)
        f_neq = fpop - feq
        PiNeq = ops.second_moment(lat, f_neq)
        QiPi1 = np.tensordot(Qi, PiNeq, axes=(1, 0))
        fpop1 = 9.0 / 2.0 * weights * QiPi1
        return feq + fpop1

    # :116-135
    def __call__(self, f_pre, f_post, bc_mask, missing_mask):
        boundary = _bmask(self.lat, bc_mask, self.id)
        feq = self.calculate_equilibrium(f_post, missing_mask)
        f_post_bd = self.bounceback_nonequilibrium(f_post, feq, missing_mask)
        f_post_bd = self.regularize_fpop(f_post_bd, feq)
        return np.where(boundary, f_post_bd, f_post).astype(f_post.dtype, copy=False)


class ExtrapolationOutflowBC(BoundaryCondition):
    """bc_extrapolation_outflow.py:48-145 (STREAMING step + post-collision assemble)."""

    def __init__(self, lat, policy, indices=None, registry=None):
        super().__init__(lat, policy, STREAMING, indices, registry)
        self._get_normal_vectors(indices)

    # :78-92
    def _get_normal_vectors(self, indices):
        freq_counts = [Counter(coord).most_common(1)[0] for coord in indices]
        counts = np.array([count for _, count in freq_counts])
        elements = np.array([element for element, _ in freq_counts])
        self.normal = counts // counts.max()
        if elements[np.argmax(counts)] == 0:
            self.normal *= -1

    # :94-104
    def _roll(self, fld, vec):
        return np.roll(fld, tuple(int(v) for v in vec), axis=tuple(range(1, self.lat.d + 1)))

    # :106-133
    def assemble_auxiliary_data(self, f_pre, f_post, bc_mask, missing_mask):
        lat = self.lat
        # ``1.0 / jnp.sqrt(3.0)`` is evaluated in JAX's default float dtype
        dflt = np.float64 if self.cdt == np.float64 else np.float32
        sound_speed = dflt(1.0) / np.sqrt(dflt(3.0))
        boundary = _bmask(lat, bc_mask, self.id)
        neighbour = self._roll(boundary, -self.normal)
        fpop = np.where(boundary, f_pre, f_post)
        fpop_neighbour = np.where(neighbour, f_pre, f_post)
        fpop_neighbour = self._roll(fpop_neighbour, self.normal)
        fpop_extrapolated = sound_speed * fpop_neighbour + (1.0 - sound_speed) * fpop
        opp = lat.opp_indices
        known_mask = missing_mask[opp]
        return np.where(np.logical_and(boundary, known_mask), fpop_extrapolated[opp], f_post)

    # :135-145
    def __call__(self, f_pre, f_post, bc_mask, missing_mask):
        boundary = _bmask(self.lat, bc_mask, self.id)
        return np.where(np.logical_and(missing_mask, boundary), f_pre[self.lat.opp_indices], f_post)
