"""Oracle lattice constants (TEST INFRASTRUCTURE -- see oracle/__init__.py).

Restates xlb/velocity_set/velocity_set.py:63-83,139-253 and the three concrete sets
(d2q9.py:16-24, d3q19.py:17-30, d3q27.py:17-32).  Direction order is the reference's:
``itertools.product([0, -1, 1], repeat=3)`` (x slowest, z fastest), D3Q19 keeps |c|_1 <= 2.
"""

import itertools
import math

import numpy as np


class Lattice:
    """Plain-data twin of ``VelocitySet`` (velocity_set.py:20-83)."""

    def __init__(self, d, q, c, w):
        self.d = d
        self.q = q
        self.c = np.array(c, dtype=np.int64)  # (d, q)  velocity_set.py:67
        self.w = np.array(w, dtype=np.float64)  # (q,)   velocity_set.py:68
        ct = self.c.T.tolist()
        # velocity_set.py:182-196
        self.opp_indices = np.array([ct.index([-v for v in ci]) for ci in ct])
        # velocity_set.py:155-180: cc columns in order (a, b>=a): xx,xy,xz,yy,yz,zz
        nt = d * (d + 1) // 2
        self.cc = np.zeros((q, nt))
        k = 0
        for a in range(d):
            for b in range(a, d):
                self.cc[:, k] = self.c[a] * self.c[b]
                k += 1
        # velocity_set.py:139-153
        self.qi = self.cc.copy()
        diag, off = ((0, 3, 5), (1, 2, 4)) if d == 3 else ((0, 2), (1,))
        self.qi[:, diag] += -1.0 / 3.0
        self.qi[:, off] *= 2.0
        # velocity_set.py:75-83, 198-253
        self.cs = math.sqrt(3) / 3.0
        self.cs2 = 1.0 / 3.0
        self.inv_cs2 = 3.0
        l1 = np.abs(self.c).sum(axis=0)
        self.main_indices = np.nonzero(l1 == 1)[0]
        self.right_indices = np.nonzero(self.c[0] == 1)[0]
        self.left_indices = np.nonzero(self.c[0] == -1)[0]
        self.center_index = int(np.nonzero(l1 == 0)[0][0])

    def __repr__(self):
        return "D{}Q{}".format(self.d, self.q)


def D3Q19():
    """d3q19.py:17-30"""
    c = np.array([ci for ci in itertools.product([0, -1, 1], repeat=3) if np.sum(np.abs(ci)) <= 2]).T
    w = np.zeros(19)
    for i in range(19):
        n = np.sum(np.abs(c[:, i]))
        w[i] = {0: 1.0 / 3.0, 1: 1.0 / 18.0, 2: 1.0 / 36.0}[int(n)]
    return Lattice(3, 19, c, w)


def D3Q27():
    """d3q27.py:17-32"""
    c = np.array(list(itertools.product([0, -1, 1], repeat=3))).T
    w = np.zeros(27)
    for i in range(27):
        n = np.sum(np.abs(c[:, i]))
        w[i] = {0: 8.0 / 27.0, 1: 2.0 / 27.0, 2: 1.0 / 54.0, 3: 1.0 / 216.0}[int(n)]
    return Lattice(3, 27, c, w)


def D2Q9():
    """d2q9.py:16-24"""
    cx = [0, 0, 0, 1, -1, 1, -1, 1, -1]
    cy = [0, 1, -1, 0, 1, -1, 0, 1, -1]
    c = np.array(tuple(zip(cx, cy))).T
    w = np.array([4 / 9, 1 / 9, 1 / 9, 1 / 9, 1 / 36, 1 / 36, 1 / 9, 1 / 36, 1 / 36])
    return Lattice(2, 9, c, w)


def make_lattice(name):
    return {"D2Q9": D2Q9, "D3Q19": D3Q19, "D3Q27": D3Q27}[name]()
