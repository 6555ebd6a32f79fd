"""Replay several consecutive steps as one CUDA graph.

On small lattices the fused step kernel is shorter than the host-side cost of issuing it (BASELINE configs[0], 128^3:
49 us of HBM time against a ~60 us Python + ctypes + launch path), so the loop

    for i in range(n):
        f_0, f_1 = stepper(f_0, f_1, bc_mask, missing_mask, omega, i)
        f_0, f_1 = f_1, f_0

(examples/performance/mlups_3d.py:227-229) is launch-bound.  ``GraphedSteps`` captures an even number of such steps --
so the buffers end where they started and the newest state is in ``f_0`` again -- into a ``torch.cuda.CUDAGraph`` once
and replays it with a single launch.  The captured work is exactly the ``xlb_step_ex`` launches the eager loop issues
(same kernels, same arguments), so the result is bit-identical to the eager loop.  ``omega`` and the field pointers are
frozen at capture; the ``timestep`` argument is not read by any kernel (nse_stepper.py:229 carries it for the Warp
path's time-dependent BCs, which the JAX path -- the parity target -- does not have).
"""

import torch


class GraphedSteps:
    def __init__(self, stepper, f_0, f_1, bc_mask, missing_mask, omega, steps=2):
        if steps < 2 or steps % 2:
            raise ValueError("GraphedSteps: steps must be an even number >= 2 (the two buffers swap every step)")
        self.stepper, self.steps = stepper, int(steps)
        self.f_0, self.f_1 = f_0, f_1
        # one eager step pair first: builds the launch plan / BC cell list (host synchronisations are illegal in a capture)
        snap = f_0.clone()
        self._issue(bc_mask, missing_mask, omega, 2)
        f_0.copy_(snap)
        del snap
        torch.cuda.synchronize()
        self.graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(self.graph):
            self._issue(bc_mask, missing_mask, omega, self.steps)

    def _issue(self, bc_mask, missing_mask, omega, n):
        a, b = self.f_0, self.f_1
        for i in range(n):
            a, b = self.stepper(a, b, bc_mask, missing_mask, omega, i)
            a, b = b, a

    def __call__(self):
        """Advance ``steps`` time steps; returns ``(f_0, f_1)`` with the newest state in ``f_0``."""
        self.graph.replay()
        return self.f_0, self.f_1
