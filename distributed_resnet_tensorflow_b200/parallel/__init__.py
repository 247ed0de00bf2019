"""Synchronous data parallelism: symmetric memory, in-kernel collectives, PS / allreduce reducers."""
from .reducers import AllreduceReducer, ParameterServerReducer, make_reducer  # noqa: F401
