"""placeholder, filled below"""


class SlabDecomposition:
    pass


def distribute(operator, grid, velocity_set, num_results=1, ops="permute"):
    raise NotImplementedError
