"""Parallel runtime: flat arenas, shard map, NVLink peer groups, asynchronous / synchronous optimizers,
the parameter-server role and process launchers."""
from .arena import FlatArena, ShardMap, ensure_arena
from .async_optim import Asynchronous, DownpourSGD, Listener, staleness_scale
from .fused_sgd import ArenaOptimizer, FusedSGD
from .server import ParameterServer, WorkerLost
from .sync_optim import Synchronous

__all__ = ["FlatArena", "ShardMap", "ensure_arena", "Asynchronous", "DownpourSGD", "Listener", "staleness_scale", "ArenaOptimizer",
           "FusedSGD", "ParameterServer", "WorkerLost", "Synchronous"]
