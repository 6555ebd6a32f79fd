"""bc_mask tag constants (xlb/cell_type.py:9-11). Registered boundary conditions use ids 1..253."""

BC_NONE = 0
BC_SFV = 254
BC_SOLID = 255
