"""Partition rules of the two multi-GPU modes; the same integer arithmetic as
pls_b200/host/engine.cpp (DeviceMirror::make_shards) and model.cpp (cv_on_device).

 * row sharding of one large fit: rank g owns rows [g*N//G, (g+1)*N//G)
 * fold sharding of cross-validation: rank g owns folds [g*F//G, (g+1)*F//G)
Both are contiguous, cover the range exactly once and differ in size by at most one unit.
"""


def bounds(total, world):
    if world < 1 or total < 0:
        raise ValueError("world must be >= 1 and total >= 0")
    return [g * total // world for g in range(world + 1)]


def row_shard(N, rank, world):
    b = bounds(N, world)
    return b[rank], b[rank + 1]


def fold_shard(F, rank, world):
    b = bounds(F, world)
    return b[rank], b[rank + 1]
