"""Public API of b200ddp (re-exported by ``b200ddp/__init__.py``): the ``torch.distributed`` surface the reference's trainers use
-- ``init_process_group`` / ``get_rank`` / ``get_world_size`` / ``all_reduce`` with ``reduce_op.SUM`` / DDP wrap
(``notebooks/code/cifar10-distributed-native-cpu.py:88-92,104-109,141``; ``cifar10-distributed-smddp-gpu.py:17-23,116,148``) --
so that a script switches over by changing its import."""
from .parallel.process_group import (ReduceOp, all_gather_object, all_reduce, barrier, broadcast,
                                     broadcast_object, destroy_process_group, get_backend, get_comm,
                                     get_device, get_local_rank, get_rank, get_world_size,
                                     init_process_group, is_initialized, max_over_ranks, reduce_op)
from .parallel.ddp import DistributedDataParallel
from .ops.functional import get_compute_dtype, set_compute_dtype

__version__ = "0.1.0"

__all__ = [
    "init_process_group", "destroy_process_group", "is_initialized", "get_backend", "get_rank",
    "get_world_size", "get_local_rank", "get_device", "get_comm", "all_reduce", "broadcast", "barrier",
    "all_gather_object", "broadcast_object", "max_over_ranks", "ReduceOp", "reduce_op",
    "DistributedDataParallel", "set_compute_dtype", "get_compute_dtype",
]
