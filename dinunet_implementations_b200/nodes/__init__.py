from .local import COINNLocal
from .remote import COINNRemote

__all__ = ["COINNLocal", "COINNRemote"]
