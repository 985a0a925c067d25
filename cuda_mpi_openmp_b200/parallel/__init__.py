"""Multi-GPU layer: one process per GPU, peer-mapped symmetric memory over NVLink 5 /
NVSwitch, and the lab workloads sharded across ranks.

``impl="fused"`` is the product path: every cross-GPU byte moves through loads,
stores and atomics issued INSIDE the lab kernels on peer-mapped addresses.
``impl="nccl"`` is the baseline the reference never had: the same local kernels with
the exchange done by NCCL collectives through torch.distributed.
"""
from .symm import (PeerGroup, DistPeerGroup, LocalPeerGroup, SoloPeerGroup, SymmetricBuffer,
                   init_distributed)
from . import vmm
from .sharded import (ShardedVectorSub, ShardedRoberts, ShardedClassifier, ShardedSort, shard_rows,
                      raise_on_error)

__all__ = [
    "PeerGroup", "DistPeerGroup", "LocalPeerGroup", "SoloPeerGroup", "SymmetricBuffer", "init_distributed", "vmm",
    "ShardedVectorSub", "ShardedRoberts", "ShardedClassifier", "ShardedSort", "shard_rows", "raise_on_error",
]
