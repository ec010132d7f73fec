"""nexus_b200 -- B200-native exact dense-retrieval engine behind hanhainebula/Nexus's retrieval surface.

Only what the hot path needs: ``csrc/`` (hand-written sm_100a kernels + the C ABI of ``include/nexus_flat.h``)
and the host-side mirror of the reference interface (``index``/``search``, a Faiss-shaped index object,
pool+normalise, row-sharded multi-GPU search).  There is no CPU fallback anywhere in this package.
"""
from ._lib import NexusFlatError, version  # noqa: F401
from .flat import FlatIndex, merge_packed, merge_topk, positive_rank  # noqa: F401
from .sharded import DistributedFlatIndex, ShardedFlatIndex, shard_bounds  # noqa: F401
from .retrieval import index, search, unpack_results  # noqa: F401
from .pooling import pool_and_normalize, pooling  # noqa: F401

__all__ = ["positive_rank", "unpack_results", "FlatIndex", "ShardedFlatIndex", "DistributedFlatIndex", "index", "search", "pool_and_normalize", "pooling",
           "merge_topk", "merge_packed", "shard_bounds", "NexusFlatError", "version"]
