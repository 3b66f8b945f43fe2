"""nano-gpu-scheduler_b200 — B200-native dealer (filter / prioritize / bind evaluation of
nano-gpu-scheduler's pkg/dealer) behind a C ABI.  See DESIGN.md and include/ngs_dealer.h."""
from . import _abi  # noqa: F401
from ._abi import (ERR_APPLY, ERR_CUDA, ERR_INFEASIBLE, ERR_INVALID, ERR_RANGE, ERR_UNKNOWN_NODE,  # noqa: F401
                   ERR_UNSUPPORTED, INT32_MIN, MAX_CONTAINERS, MAX_FREE, MAX_GPUS, MAX_GPUS_WIDE, MAX_PERIODS, METRIC_CORE,
                   METRIC_MEMORY, OK, POLICY_BINPACK, POLICY_RANDOM, POLICY_SPREAD, REASON_INFEASIBLE, REASON_OK,
                   REASON_UNKNOWN_NODE)
from .dealer import (GpuDealer, NgsError, make_load_ctx, parse_update_time, parse_usage,  # noqa: F401
                     schedule_batch_noalloc)
