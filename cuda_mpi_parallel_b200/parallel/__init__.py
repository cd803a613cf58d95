from .partition import make_partition, make_layout, local_slices
from .fabric import attach_single_process, attach_torch_distributed

__all__ = ["make_partition", "make_layout", "local_slices", "attach_single_process", "attach_torch_distributed"]
