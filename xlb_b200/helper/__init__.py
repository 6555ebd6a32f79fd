from .nse_fields import create_nse_fields as create_nse_fields
from .initializers import initialize_eq as initialize_eq
from .check_boundary_overlaps import check_bc_overlaps as check_bc_overlaps
from .mesh import read_stl as read_stl, write_stl as write_stl, fit_to_grid as fit_to_grid, ellipsoid_soup as ellipsoid_soup
