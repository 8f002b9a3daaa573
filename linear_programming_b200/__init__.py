"""linear_programming_b200 -- B200-native dense LP engine (drop-in for the hot path of
aakarsh/linear_programming: week-2 simplex and Gauss-Jordan).  See DESIGN.md.

All numerics run in liblp_b200.so (hand-written sm_100a CUDA behind the C ABI of
include/lp_b200.h); importing this package does not load it, the first solve does, and it
raises if the library or a CUDA device is missing -- there is no CPU fallback.
"""
from .simplex import (BatchResult, IterationLimitError, Simplex, allocate_ads, solve, solve_batch,
                      solve_diet_problem, verify_bounds_batch)
from .gauss_jordan import solve_equation, solve_equation_batch
from ._lib import EngineError
from .sharded import ShardedComm, column_shard, problem_shard, solve_sharded_emulated

__all__ = ["Simplex", "solve", "solve_batch", "BatchResult", "allocate_ads", "solve_diet_problem",
           "solve_equation", "solve_equation_batch", "EngineError", "IterationLimitError", "ShardedComm", "column_shard", "problem_shard",
           "solve_sharded_emulated", "verify_bounds_batch"]
