"""Fetch the reference's own example scripts for the drop-in test (TEST INFRASTRUCTURE -- see oracle/__init__.py).

``tests/test_reference_examples.py`` runs the reference's example SOURCES against ``xlb_b200`` with nothing changed but
the backend they select (north_star: "existing examples run by switching only the backend").  The GPU
box has no ``/root/reference``, and reference sources must not be committed, so -- like ``oracle/_ref`` for a compiled
reference -- the scripts are copied verbatim, where they lie, into the git-ignored ``oracle/_ref/examples/``, which travels
with the repo snapshot.  ``__graft_entry__.build()`` calls ``fetch()`` whenever the reference tree is present.

    python -m oracle.ref_examples          # copy now
"""

import os
import shutil

REFERENCE = "/root/reference"
HERE = os.path.dirname(os.path.abspath(__file__))
DEST = os.path.join(HERE, "_ref", "examples")

# (path under the reference tree, why it is here)
EXAMPLES = (
    ("examples/cfd/lid_driven_cavity_2d.py", "D2Q9 BGK cavity, half-way walls + equilibrium lid; class-based, post-processing through Macroscopic"),
    ("examples/cfd/lid_driven_cavity_2d_distributed.py", "the reference's only shipped use of distribute() (2-D slabs)"),
    ("examples/cfd/flow_past_sphere_3d.py", "D3Q19 BGK, RegularizedBC inlet with a per-index profile, extrapolation outflow, half-way sphere, MomentumTransfer"),
    ("examples/performance/mlups_3d.py", "the MLUPS benchmark driver of BASELINE configs[0]"),
)


def fetch(reference=REFERENCE, dest=DEST):
    """Copy the example scripts; returns the list of files written ([] when the reference tree is absent)."""
    if not os.path.isdir(reference):
        return []
    os.makedirs(dest, exist_ok=True)
    out = []
    for rel, _ in EXAMPLES:
        dst = os.path.join(dest, os.path.basename(rel))
        shutil.copyfile(os.path.join(reference, rel), dst)
        out.append(dst)
    return out


def path_of(name, dest=DEST):
    """Path of a fetched example, fetching first if it is missing and the reference tree is here; None if unavailable."""
    p = os.path.join(dest, name)
    if not os.path.exists(p):
        fetch(dest=dest)
    return p if os.path.exists(p) else None


if __name__ == "__main__":
    for f in fetch():
        print(f)
