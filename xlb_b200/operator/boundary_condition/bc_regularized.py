"""Regularized velocity / pressure boundary condition (xlb/operator/boundary_condition/bc_regularized.py:52-135).

Zou-He (non-equilibrium bounce-back of the missing populations) followed by Latt's regularisation: every population
of the boundary cell is rebuilt as ``feq + 4.5 w_l (Q_l : Pi_neq)`` with ``Pi_neq = sum_l c c (f - feq)`` (Latt et al.
2008, eq. 45).  This is the inlet the reference's shipped sphere / wind-tunnel examples use
(examples/cfd/flow_past_sphere_3d.py:107, windtunnel_3d.py:103).  Same constructor as ``ZouHeBC``.
"""

from ... import _lib
from .bc_zouhe import ZouHeBC


class RegularizedBC(ZouHeBC):
    def __init__(self, bc_type, profile=None, prescribed_value=None, velocity_set=None, precision_policy=None, compute_backend=None,
                 indices=None, mesh_vertices=None, voxelization_method=None):
        super().__init__(bc_type, profile, prescribed_value, velocity_set, precision_policy, compute_backend, indices, mesh_vertices,
                         voxelization_method)
        self.abi_kind = _lib.BC_REGULARIZED_VELOCITY if bc_type == "velocity" else _lib.BC_REGULARIZED_PRESSURE
