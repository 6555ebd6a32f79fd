from .distribute import distribute as distribute, SlabDecomposition as SlabDecomposition, HaloExchange as HaloExchange, DistributedStepper as DistributedStepper
