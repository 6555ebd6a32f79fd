from .distribute import distribute as distribute, balanced_planes as balanced_planes, SlabDecomposition as SlabDecomposition, HaloExchange as HaloExchange, DistributedStepper as DistributedStepper
