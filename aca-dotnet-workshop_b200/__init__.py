"""B200-native overdue-task sweep and creator filter behind the reference's ITasksManager.

The directory name carries the reference's hyphens, so import it with
``importlib.import_module("aca-dotnet-workshop_b200")`` (see tests/conftest.py).

Everything here is a thin host layer over ``libtasktable.so`` (hand-written sm_100a CUDA,
C ABI in ``include/tasktable.h``).  There is no CPU fallback: importing works without a GPU,
but every operation raises if the library or a CUDA device is missing.
"""
from .tasktable import (  # noqa: F401
    MODE_FAKE,
    MODE_STORE,
    SWEEP_LIST_ONLY,
    SWEEP_DUE_ON_OR_BEFORE,
    SWEEP_MARK,
    TICKS_PER_DAY,
    DEFAULT_TODAY,
    DEFAULT_RUN_AT,
    DeviceView,
    Exchange,
    GResult,
    ShardedTable,
    JSON_DOCUMENT,
    JSON_RESPONSE,
    JsonText,
    NotImplementedInReference,
    Result,
    SweepStats,
    SynthCfg,
    TaskTable,
    TaskTableError,
    default_cfg,
    lib,
    lib_path,
)
from .tasks_manager import GpuTasksManager, TaskModel  # noqa: F401
from .sharding import HitExchange, ShardExchange, ShardPlan, exchange_hits, push_slot_layout, shard_rows  # noqa: F401
