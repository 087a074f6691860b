"""Replica sharding for batch data generation: instance
i runs on GPU ``i mod G``; no collective on the data path."""


def shard(n_instances: int, world: int, rank: int):
    """Instance ids owned by `rank` out of `world` ranks (static, i mod G)."""
    if world <= 0 or not (0 <= rank < world):
        raise ValueError(f"bad rank/world {rank}/{world}")
    return list(range(rank, n_instances, world))


def frames_per_second(total_frames: int, per_rank_ms):
    """Whole-job throughput: all frames of all ranks / the slowest rank's time."""
    worst = max(per_rank_ms)
    return total_frames / (worst * 1e-3) if worst > 0 else float("inf")
