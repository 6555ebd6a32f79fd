import itertools

import numpy as np

from .velocity_set import VelocitySet


class D3Q27(VelocitySet):
    """All 27 velocities of {0,-1,1}^3, x slowest (d3q27.py:17-32)."""

    def __init__(self, precision_policy=None, compute_backend=None):
        dirs = list(itertools.product([0, -1, 1], repeat=3))
        c = np.array(dirs).T
        by_speed = {0: 8.0 / 27.0, 1: 2.0 / 27.0, 2: 1.0 / 54.0, 3: 1.0 / 216.0}
        w = np.array([by_speed[sum(map(abs, v))] for v in dirs])
        super().__init__(3, 27, c, w, precision_policy=precision_policy, compute_backend=compute_backend)
