"""cluster_anywhere_b200 -- B200-native engine for Ray Data's all-to-all exchange
(Dataset.sort / random_shuffle / groupby().aggregate on Arrow blocks).

Only the hot path is here: CUDA kernels + C-ABI (``csrc/``, ``include/xchg_b200.h``)
and the host-side mirror of the reference's operator surface.
"""
from ._cabi import XchgError  # noqa: F401
from .block import BlockExecStats, BlockMetadata  # noqa: F401
from .block_accessor import B200ArrowBlockAccessor  # noqa: F401
from .sort_key import SortKey  # noqa: F401
from .aggregate import AbsMax, AggregateFn, Count, Max, Mean, Min, Std, Sum  # noqa: F401

__all__ = ["B200ArrowBlockAccessor", "SortKey", "BlockMetadata", "BlockExecStats", "XchgError",
           "AggregateFn", "Count", "Sum", "Min", "Max", "Mean", "Std", "AbsMax"]
