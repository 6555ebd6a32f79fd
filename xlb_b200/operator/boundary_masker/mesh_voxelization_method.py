"""Mesh voxelisation method registry (xlb/operator/boundary_masker/mesh_voxelization_method.py:12-55): the same names
and ids as the reference.  The CUDA backend implements ``RAY`` (the method the shipped wind-tunnel example uses,
examples/cfd/windtunnel_3d.py:79); the others raise ``NotImplementedError`` when a stepper meets them."""

from dataclasses import dataclass

METHODS = {"AABB": 1, "RAY": 2, "AABB_CLOSE": 3, "WINDING": 4}


@dataclass
class VoxelizationMethod:
    id: int
    name: str
    options: dict


def MeshVoxelizationMethod(name: str, **options):
    assert name in METHODS.keys(), f"Unsupported voxelization method: {name}"
    return VoxelizationMethod(METHODS[name], name, options)
