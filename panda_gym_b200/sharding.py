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


def bind_to_local_numa(device_index):
    """Pin the calling process to the CPUs NVML reports as local to CUDA device ``device_index`` (one process per GPU).
    Page-locked host buffers allocated afterwards come from that NUMA node, so the per-step result copy of every rank
    stays on its own socket / PCIe root instead of all ranks sharing the launcher's node.  Returns a small report dict
    (or the reason nothing was done); never raises."""
    import os
    try:
        import pynvml
        props = torch.cuda.get_device_properties(device_index)
        pynvml.nvmlInit()
        bus = "%08x:%02x:%02x.0" % (props.pci_domain_id, props.pci_bus_id, props.pci_device_id)
        h = pynvml.nvmlDeviceGetHandleByPciBusId(bus.encode())
        words = (os.cpu_count() + 63) // 64
        mask = pynvml.nvmlDeviceGetCpuAffinity(h, words)
        cpus = [64 * w + b for w in range(words) for b in range(64) if (int(mask[w]) >> b) & 1]
        allowed = sorted(set(cpus) & set(os.sched_getaffinity(0)))
        if not allowed:
            return {"bound": False, "why": "no overlap between the GPU's local CPUs and this process's allowed set"}
        os.sched_setaffinity(0, allowed)
        return {"bound": True, "pci": bus, "cpus": "%d-%d (%d)" % (allowed[0], allowed[-1], len(allowed))}
    except Exception as ex:  # NVML missing, container without topology, ...
        return {"bound": False, "why": repr(ex)[:200]}


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


class AsyncStats:
    """The episode-statistics all-reduce off the step stream: a snapshot of the 8-entry vector is taken on the current stream,
    the collective runs on a side stream behind an event, and the simulation carries on -- a rank that reaches the collective
    early no longer stalls its step kernels until the slowest rank arrives (at 8 ranks that wait, not the 64-byte transfer, was
    the whole cost of the collective).  ``result()`` waits and returns the summed dict of the LAST ``start()``."""

    def __init__(self, device, group=None):
        self.side = torch.cuda.Stream(device=device)
        self.group = group
        self.pending = None

    def start(self, stats):
        s = stats.detach().clone().to(torch.float64)      # snapshot on the caller's stream
        ev = torch.cuda.Event()
        ev.record()
        work = None
        if dist.is_available() and dist.is_initialized():
            with torch.cuda.stream(self.side):
                self.side.wait_event(ev)
                work = dist.all_reduce(s, op=dist.ReduceOp.SUM, group=self.group, async_op=True)
        self.pending = (s, work, ev)

    def result(self):
        if self.pending is None:
            return None
        s, work, ev = self.pending
        if work is not None:
            work.wait()                                    # orders the current stream after the collective
            self.side.synchronize()
        else:
            ev.synchronize()
        v = s.cpu().tolist()
        out = {k: float(v[i]) for i, k in enumerate(STAT_KEYS[:7])}
        ep = max(out["episodes"], 1.0)
        out.update(success_rate=out["successes"] / ep, collision_rate=out["collisions"] / ep,
                   mean_return=out["sum_return"] / ep, mean_length=out["sum_length"] / ep)
        return out


def allgather_packed(packed, obs_dim, act_dim, group=None):
    """Replay all-gather of a batch that already IS one ``[B, W]`` block -- ``HerReplayBuffer.last_packed()``, rows in the
    stored record's field order -- with ONE ``all_gather_into_tensor`` and no packing pass; the nine fields come back as
    column ranges of the gathered ``[world * B, W]`` block (rank-major)."""
    from .her import DictReplayBufferSamples
    out = packed
    if dist.is_available() and dist.is_initialized():
        world = dist.get_world_size(group)
        out = torch.empty((world * packed.shape[0], packed.shape[1]), dtype=packed.dtype, device=packed.device)
        dist.all_gather_into_tensor(out, packed.contiguous(), group=group)
    od, ad = int(obs_dim), int(act_dim)
    c = 2 * od
    obs = {"observation": out[:, :od], "achieved_goal": out[:, c:c + 3], "desired_goal": out[:, c + 3:c + 6]}
    nobs = {"observation": out[:, od:c], "achieved_goal": out[:, c + 6:c + 9], "desired_goal": out[:, c + 9:c + 12]}
    return DictReplayBufferSamples(obs, out[:, c + 12:c + 12 + ad], nobs, out[:, c + 13 + ad:c + 14 + ad], out[:, c + 12 + ad:c + 13 + ad])


def allgather_samples(samples, group=None):
    """Replay all-gather: every rank contributes the minibatch it sampled from its own replay shard
    (``HerReplayBuffer.sample``: the shard holds the transitions of the rank's environments) and receives the
    concatenation over ranks, rank-major -- what a single learner that trains on all shards consumes
    (SURVEY 8e).  The nine fields of a sample are packed into one ``[B, W]`` row block so the exchange is ONE
    ``all_gather_into_tensor`` (NCCL over NVLink on GPUs; gloo in the CPU tests), then split back into views of the
    gathered block.  Every rank must pass the same batch size.  Without an initialised process group it returns
    the input unchanged."""
    from .her import DictReplayBufferSamples
    if not (dist.is_available() and dist.is_initialized()):
        return samples
    world = dist.get_world_size(group)
    o, no = samples.observations, samples.next_observations
    keys = ("observation", "achieved_goal", "desired_goal")
    fields = [o[k] for k in keys] + [samples.actions] + [no[k] for k in keys] + [samples.dones, samples.rewards]
    B = fields[0].shape[0]
    widths = [int(f.numel() // B) for f in fields]
    packed = torch.cat([f.reshape(B, w).to(torch.float32) for f, w in zip(fields, widths)], dim=1).contiguous()
    out = torch.empty((world * B, packed.shape[1]), dtype=packed.dtype, device=packed.device)
    dist.all_gather_into_tensor(out, packed, group=group)
    parts = torch.split(out, widths, dim=1)
    obs = {k: parts[i] for i, k in enumerate(keys)}
    nobs = {k: parts[4 + i] for i, k in enumerate(keys)}
    return DictReplayBufferSamples(obs, parts[3], nobs, parts[7], parts[8])


def benchmark_results(raw, n_substeps=20, group=None):
    """The ``results`` dict of evaluation/evaluate.py:perform_benchmark (:286-300) from the 12 accumulators of
    ``pg_get_eval_metrics`` (summed over ranks first when a process group is up).  Same keys, same rounding."""
    import numpy as np
    v = torch.as_tensor(np.asarray(raw, dtype=np.float64)).clone()
    if dist.is_available() and dist.is_initialized():
        if dist.get_backend(group) == "nccl":
            v = v.cuda()
        dist.all_reduce(v, op=dist.ReduceOp.SUM, group=group)
    ns, ls, eff, man, jerk, spd, ne, le, ncol, ret, nto = [float(x) for x in v.cpu().tolist()[:11]]
    div = lambda a, b: float("nan") if b == 0 else a / b   # np.mean of an empty list is nan in the reference too
    r4 = lambda x: float(np.round(x, 4))
    return {"mean_reward": r4(div(ret, ne)), "success_rate": r4(div(ns, ne)), "collision_rate": r4(div(ncol, ne)),
            "timeout_rate": r4(div(nto, ne)), "num_episodes": r4(ne), "mean_ep_length": r4(div(le, ne)),
            "mean_ep_length_success": r4(div(ls, ns)), "mean_num_sim_steps": r4(div(le * n_substeps, ne)),
            "mean_num_sim_steps_success": r4(div(ls * n_substeps, ns)), "mean_effort": r4(div(eff, ns)),
            "mean_manipulability": r4(div(man, ns)), "mean_norm_jerk": r4(div(jerk, ls)), "mean_ee_speed": r4(div(spd, ls))}
