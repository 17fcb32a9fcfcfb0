"""Multi-GPU plumbing of the path: environments shard contiguously over ranks, no data-path collective; the only
exchange is the all-reduce of the episode-statistics vector (what the curriculum gate of the reference,
setup_training.py:281-284, consumes).  Backend: NCCL on GPUs, gloo in the CPU tests."""
import torch
import torch.distributed as dist

STAT_KEYS = ("episodes", "successes", "collisions", "timeouts", "sum_return", "sum_length", "reset_failures", "reserved")


def env_shard(rank, envs_per_rank):
    """[first, last) global environment indices owned by `rank`."""
    return rank * envs_per_rank, (rank + 1) * envs_per_rank


def first_reset_seeds(base_seed, rank, envs_per_rank):
    """Seed of each local environment's first episode: SB3's VecEnv.seed convention seed + global index."""
    lo, hi = env_shard(rank, envs_per_rank)
    return [base_seed + g for g in range(lo, hi)]


def allreduce_stats(stats, group=None):
    """Sum the 8-entry statistics vector over all ranks (in place on a clone); returns a dict."""
    s = stats.detach().clone().to(torch.float64)
    if dist.is_available() and dist.is_initialized():
        dist.all_reduce(s, op=dist.ReduceOp.SUM, group=group)
    v = s.cpu().tolist()
    out = {k: float(v[i]) for i, k in enumerate(STAT_KEYS[:7])}
    ep = max(out["episodes"], 1.0)
    out["success_rate"] = out["successes"] / ep
    out["collision_rate"] = out["collisions"] / ep
    out["mean_return"] = out["sum_return"] / ep
    out["mean_length"] = out["sum_length"] / ep
    return out
