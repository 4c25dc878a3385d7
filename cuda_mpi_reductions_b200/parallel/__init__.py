"""Multi-GPU layer: communicators over the symmetric heap and the collectives built on them."""
from .comm import Communicator, host_reduce_threads

__all__ = ["Communicator", "host_reduce_threads"]
