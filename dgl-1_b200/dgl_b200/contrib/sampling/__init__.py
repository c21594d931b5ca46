from .sampler import NeighborSampler  # noqa: F401
