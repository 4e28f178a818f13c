"""arrow-ballista_b200: B200-native aggregation + shuffle hot path for Ballista.

The compute lives in libbgpu.so (hand-written sm_100a CUDA behind the C ABI of include/bgpu.h); this package is
the host-side mirror of the reference operator interfaces for that path. Import it via `import ballista_b200`
(the directory name contains a hyphen).
"""
from ._lib import LIB_PATH, SIGNATURES, BgpuError, lib  # noqa: F401
from .runtime import Context, DeviceBatch, create_hashes  # noqa: F401
from .partitioner import BatchPartitioner, Partitioning  # noqa: F401
from .aggregate import AggregateFunction, AggregateMode, GpuAggregateExec  # noqa: F401
from .expr import Column, Literal, BinaryExpr, col, lit  # noqa: F401
from . import expr  # noqa: F401
from .exchange import Exchange, exchange_plan, owner_of  # noqa: F401
from .shuffle_writer import GpuShuffleWriterExec, ShuffleWritePartition, result_schema  # noqa: F401
from . import ipc  # noqa: F401
from .shuffle_reader import (CoalesceBatchesExec, CoalesceTasksExec, ExecutorMetadata, FetchFailed, GpuShuffleReaderExec,  # noqa: F401
                             PartitionLocation, PartitionStats, stats_for_partitions)
from .codec import GpuAggregateExecNode, GpuPhysicalCodec  # noqa: F401
from .execution_engine import (DefaultExecutionEngine, DefaultQueryStageExec, GpuExecutionEngine, GpuQueryStageExec,  # noqa: F401
                               QueryStageExecutor, match_aggregate_stage)
from .task_group import OptimizeTaskGroup  # noqa: F401
