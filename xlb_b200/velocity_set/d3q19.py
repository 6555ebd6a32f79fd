import itertools

import numpy as np

from .velocity_set import VelocitySet


class D3Q19(VelocitySet):
    """19 velocities of the {0,-1,1}^3 enumeration with at most two non-zero components (d3q19.py:17-30)."""

    def __init__(self, precision_policy=None, compute_backend=None):
        dirs = [v for v in itertools.product([0, -1, 1], repeat=3) if sum(map(abs, v)) <= 2]
        c = np.array(dirs).T
        by_speed = {0: 1.0 / 3.0, 1: 1.0 / 18.0, 2: 1.0 / 36.0}
        w = np.array([by_speed[sum(map(abs, v))] for v in dirs])
        super().__init__(3, 19, c, w, precision_policy=precision_policy, compute_backend=compute_backend)
