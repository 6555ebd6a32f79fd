"""Momentum-exchange force on a no-slip obstacle (xlb/operator/force/momentum_transfer.py:120-211).

``MomentumTransfer(no_slip_bc)(f_0, f_1, bc_mask, missing_mask) -> force[d]`` with ``f_0`` the post-collision
populations the stepper returned (flow_past_sphere_3d.py:139-172).  One launch of ``xlb_momentum_transfer``: the
pull-stream, the no-slip BC and the reduction run in the kernel (FP64 accumulation); nothing is materialised.
On an x-slab grid (``grid=`` given, one process per GPU) the partial sums are all-reduced over the ranks.
"""

from enum import Enum, auto

import torch

from ...compute_backend import ComputeBackend
from ..operator import Operator, byref, check, lib, stream_ptr


class LBMOperationSequence(Enum):
    """momentum_transfer.py:16-26; the dense single-resolution stepper is stream-then-collide."""

    STREAM_THEN_COLLIDE = auto()
    COLLIDE_THEN_STREAM = auto()


class MomentumTransfer(Operator):
    def __init__(
        self,
        no_slip_bc_instance,
        operation_sequence: LBMOperationSequence = LBMOperationSequence.STREAM_THEN_COLLIDE,
        velocity_set=None,
        precision_policy=None,
        compute_backend=None,
        grid=None,
    ):
        self.no_slip_bc_instance = no_slip_bc_instance
        self.operation_sequence = operation_sequence
        if operation_sequence != LBMOperationSequence.STREAM_THEN_COLLIDE:
            raise NotImplementedError("COLLIDE_THEN_STREAM belongs to the multi-resolution stepper, which is outside this backend")
        self.grid = grid
        super().__init__(velocity_set, precision_policy, compute_backend)

    @Operator.register_backend(ComputeBackend.CUDA)
    def cuda_implementation(self, f_0, f_1, bc_mask, missing_mask):
        vs = self.velocity_set
        self._check_field(f_0, "f_0", vs.q)
        self._check_field(bc_mask, "bc_mask", 1)
        self._check_field(missing_mask, "missing_mask", vs.q)
        bc = self.no_slip_bc_instance
        halo = getattr(self.grid, "halo", 0) if self.grid is not None else 0
        dom = self._domain(f_0, x_periodic=0 if halo else 1)
        x_begin, x_end = (halo, dom.nx - halo) if vs.d == 3 else (0, dom.nx)
        table = bc._single_row_table(f_0.shape[1:])
        acc = torch.zeros(3, dtype=torch.float64, device=f_0.device)
        check(
            lib().xlb_momentum_transfer(
                byref(dom), self._code(f_0), f_0.data_ptr(), bc_mask.data_ptr(), missing_mask.data_ptr(), table.data_ptr(), int(bc.id),
                int(x_begin), int(x_end), acc.data_ptr(), stream_ptr(),
            ),
            "MomentumTransfer",
        )
        if halo and torch.distributed.is_available() and torch.distributed.is_initialized():
            torch.distributed.all_reduce(acc)
        return acc[3 - vs.d :].to(self.compute_dtype)  # kernel axes -> the d lattice axes
