from .distribute import distribute as distribute, SlabDecomposition as SlabDecomposition
