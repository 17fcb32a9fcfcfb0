"""N > 1 path on CPU: two gloo ranks shard the environments and all-reduce the episode statistics."""
import os
import sys

import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _worker(rank, world, port, out):
    sys.path.insert(0, ROOT)
    os.environ["MASTER_ADDR"] = "127.0.0.1"; os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from panda_gym_b200.sharding import allreduce_stats, env_shard, first_reset_seeds
    n = 1024
    lo, hi = env_shard(rank, n)
    seeds = first_reset_seeds(100, rank, n)
    # local statistics as k_step accumulates them: [episodes, successes, collisions, timeouts, sum_ret, sum_len, fails, -]
    local = torch.tensor([10.0 + rank, 2.0 * (rank + 1), 5.0, 3.0 + rank - 2 * (rank + 1) + 2 * (rank + 1), -50.0 * (rank + 1), 400.0, 0.0, 0.0], dtype=torch.float64)
    red = allreduce_stats(local)
    out[rank] = (lo, hi, seeds[0], seeds[-1], red)
    dist.barrier()
    dist.destroy_process_group()


def test_two_rank_sharding_and_stats_allreduce():
    world = 2
    mgr = mp.Manager(); out = mgr.dict()
    port = 29650 + os.getpid() % 200
    mp.spawn(_worker, args=(world, port, out), nprocs=world, join=True)
    assert out[0][:2] == (0, 1024) and out[1][:2] == (1024, 2048)          # contiguous, disjoint shards
    assert out[0][2:4] == (100, 1123) and out[1][2:4] == (1124, 2147)      # seed + global env index
    r0, r1 = out[0][4], out[1][4]
    assert r0 == r1                                                         # every rank sees the global statistics
    assert r0["episodes"] == 21.0 and r0["successes"] == 6.0 and r0["collisions"] == 10.0
    assert abs(r0["success_rate"] - 6.0 / 21.0) < 1e-12 and abs(r0["mean_return"] + 150.0 / 21.0) < 1e-12


def _gather_worker(rank, world, port, out):
    sys.path.insert(0, ROOT)
    os.environ["MASTER_ADDR"] = "127.0.0.1"; os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from panda_gym_b200.her import DictReplayBufferSamples
    from panda_gym_b200.sharding import allgather_samples
    B, od, ad = 5, 56, 7
    g = torch.Generator().manual_seed(100 + rank)
    mk = lambda w: torch.rand((B, w), generator=g) + 10.0 * rank
    local = DictReplayBufferSamples({"observation": mk(od), "achieved_goal": mk(3), "desired_goal": mk(3)}, mk(ad),
                                    {"observation": mk(od), "achieved_goal": mk(3), "desired_goal": mk(3)}, mk(1), mk(1))
    full = allgather_samples(local)
    ok = full.observations["observation"].shape == (world * B, od) and full.actions.shape == (world * B, ad)
    sl = slice(rank * B, (rank + 1) * B)   # my own block sits at my rank's offset, bit-identical
    ok &= torch.equal(full.observations["observation"][sl], local.observations["observation"])
    ok &= torch.equal(full.next_observations["desired_goal"][sl], local.next_observations["desired_goal"])
    ok &= torch.equal(full.actions[sl], local.actions) and torch.equal(full.rewards[sl], local.rewards) and torch.equal(full.dones[sl], local.dones)
    other = 1 - rank                        # the other rank's block carries its rank offset
    ok &= bool((full.rewards[other * B:(other + 1) * B] >= 10.0 * other).all() and (full.rewards[other * B:(other + 1) * B] < 10.0 * other + 1).all())
    # the packed form (HerReplayBuffer.last_packed(): one [B, 144] block in record order) goes out without a packing pass
    from panda_gym_b200.sharding import allgather_packed
    W = 144
    blk = torch.arange(B * W, dtype=torch.float32).reshape(B, W) + 1000.0 * rank
    fp = allgather_packed(blk, od, ad)
    ok &= fp.observations["observation"].shape == (world * B, od) and fp.rewards.shape == (world * B, 1)
    ok &= torch.equal(fp.observations["observation"][sl], blk[:, :od]) and torch.equal(fp.next_observations["observation"][sl], blk[:, od:2 * od])
    ok &= torch.equal(fp.observations["desired_goal"][sl], blk[:, 2 * od + 3:2 * od + 6]) and torch.equal(fp.actions[sl], blk[:, 2 * od + 12:2 * od + 19])
    ok &= torch.equal(fp.rewards[sl], blk[:, 2 * od + 19:2 * od + 20]) and torch.equal(fp.dones[sl], blk[:, 2 * od + 20:2 * od + 21])
    ok &= bool((fp.actions[other * B:(other + 1) * B] >= 1000.0 * other).all())
    out[rank] = (bool(ok), float(full.actions.sum()))
    dist.barrier()
    dist.destroy_process_group()


def test_two_rank_replay_allgather():
    world = 2
    mgr = mp.Manager(); out = mgr.dict()
    port = 29870 + os.getpid() % 100
    mp.spawn(_gather_worker, args=(world, port, out), nprocs=world, join=True)
    assert out[0][0] and out[1][0]
    assert out[0][1] == out[1][1]           # both ranks hold the same gathered batch
