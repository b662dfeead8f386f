"""Partition of the prior sample over ranks (one process per GPU).

The objective is a weighted sum over prior draws (src/designproblem.jl:205-215), so every rank
evaluates all particles on a contiguous shard of ceil(N / world) draws and the S partial sums are
added with one all-reduce per swarm evaluation.  The reference has no counterpart (it is serial).
"""


def shard_range(N, rank, world):
    """Half-open draw range [j0, j1) owned by `rank`; shards are contiguous, sizes differ by <= 1
    chunk and the last shards may be empty only when world > N."""
    if world < 1 or rank < 0 or rank >= world:
        raise ValueError("need 0 <= rank < world")
    per = -(-N // world)
    j0 = min(N, rank * per)
    return j0, min(N, j0 + per)
