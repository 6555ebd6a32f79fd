import numpy as np

from .velocity_set import VelocitySet


class D2Q9(VelocitySet):
    """The reference's D2Q9 ordering (d2q9.py:16-24), which is *not* the product enumeration."""

    def __init__(self, precision_policy=None, compute_backend=None):
        c = np.array([[0, 0, 0, 1, -1, 1, -1, 1, -1], [0, 1, -1, 0, 1, -1, 0, 1, -1]])
        w = np.array([4 / 9, 1 / 9, 1 / 9, 1 / 9, 1 / 36, 1 / 36, 1 / 9, 1 / 36, 1 / 36])
        super().__init__(2, 9, c, w, precision_policy=precision_policy, compute_backend=compute_backend)
