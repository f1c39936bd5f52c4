"""Timers, loggers, meters, schedules, checkpoints, mixed-precision helpers."""
from .timing import Timer, DeviceTimer, PhaseTimer, localtime  # noqa: F401
from .logging import (TableLogger, TSVLogger, StatsLogger, FileLogger, ScalarLogger,  # noqa: F401
                      TensorboardLogger, NoOp)
from .meter import AverageMeter, TimeMeter, NetworkMeter, network_bytes  # noqa: F401
from .schedule import PiecewiseLinear, PhaseScheduler, listify  # noqa: F401
from .checkpoint import save_checkpoint, load_checkpoint  # noqa: F401
from .hostread import LaggedReader  # noqa: F401
from .misc import union, path_iter, text_pickle, text_unpickle, is_set, random_id  # noqa: F401
