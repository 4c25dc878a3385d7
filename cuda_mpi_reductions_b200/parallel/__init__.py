"""Multi-GPU layer: communicators over the symmetric heap and the collectives built on them (comm.py), and their
multi-process CPU twin over shared memory (host.py)."""
from .comm import Communicator, host_reduce_threads
from .host import HostCommunicator

__all__ = ["Communicator", "HostCommunicator", "host_reduce_threads"]
