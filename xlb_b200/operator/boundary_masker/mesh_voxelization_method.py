"""Names of the mesh voxelisation strategies, API-compatible with the reference's registry
(xlb/operator/boundary_masker/mesh_voxelization_method.py:12-55): ``MeshVoxelizationMethod("RAY")`` returns an object with
``.id``, ``.name`` and ``.options`` carrying the reference's numbering (AABB 1, RAY 2, AABB_CLOSE 3, WINDING 4).

The CUDA backend implements ``RAY`` -- the method the shipped wind-tunnel example selects (examples/cfd/windtunnel_3d.py:79);
a stepper that meets another one raises ``NotImplementedError``.
"""

_ORDER = ("AABB", "RAY", "AABB_CLOSE", "WINDING")
METHODS = {label: number for number, label in enumerate(_ORDER, start=1)}


class VoxelizationMethod:
    """One voxelisation choice: numeric ``id``, ``name`` and free-form ``options`` (e.g. ``close_voxels``)."""

    __slots__ = ("id", "name", "options")

    def __init__(self, id, name, options=None):
        self.id, self.name, self.options = int(id), str(name), dict(options or {})

    def __eq__(self, other):
        return isinstance(other, VoxelizationMethod) and (self.id, self.options) == (other.id, other.options)

    def __hash__(self):
        return hash((self.id, tuple(sorted(self.options.items()))))

    def __repr__(self):
        return f"VoxelizationMethod(id={self.id}, name={self.name!r}, options={self.options!r})"


def MeshVoxelizationMethod(name, **options):
    if name not in METHODS:
        raise AssertionError(f"Unsupported voxelization method: {name}")
    return VoxelizationMethod(METHODS[name], name, options)
