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


def cluster_for_share(n_garments: int, n_sms: int = 148, max_cluster: int = 16) -> int:
    """CTAs per garment for a shared-launch batch of n_garments on one GPU: the largest power of two that still lets
    every garment of the share be resident at once (512 garments on one B200 -> 1, i.e. 3.5 waves of one SM each;
    64 per GPU on eight -> 2; 16 -> 8)."""
    c = 1
    while c * 2 <= max_cluster and n_garments * c * 2 <= n_sms:
        c *= 2
    return c
