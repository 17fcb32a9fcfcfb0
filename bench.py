#!/usr/bin/env python3
"""bench.py -- env-steps/s of the batched Panda simulator on N B200s of one node.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--scenario reachao3]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P bench.py --gpus N ...

Metric (BASELINE.json): env-steps/sec, whole box.  Workload = BASELINE.json configs[1]: collision-avoidance reach
(PandaReachAO-v3, joints control, sparse reward, collision_reward -100, truncate on collision, terminate on
success, TimeLimit 100, 20 substeps of 1/500 s, SB3-style auto-reset) with 65 536 environments per GPU.
One "step" = one vectorised env.step over all environments of a rank.  Prints ONE JSON line on rank 0.

  value     : device-timed (CUDA events, per step, L2 flushed between steps), actions pre-generated in HBM
  e2e       : the same metric through the public host API (PandaVecEnv.step with pinned numpy actions, results
              copied back to host numpy every step, finished episodes' Monitor data consumed every step)
  roofline  : binding roofline of the step kernel is FP32 FMA issue (no dense contraction => tensor cores unused,
              HBM traffic ~1 KB/env-step is <1% of peak); achieved = algorithmic FLOPs/launch / kernel time,
              peak = FP32 FMA micro-benchmark measured live in this process (MEASURED_PEAKS.json has no FP32 figure).
              The HBM view of the same launch is reported alongside against MEASURED_PEAKS.json's hbm_gbs.
  cpu_baseline : the CPU oracle (kind "port": C restatement of the reference path, oracle/) on the host cores,
              bounded sample, rank 0 / N=1 only.  --impl reference times that same CPU implementation with all host
              threads (the reference's PyBullet backend is not installable here: no wheel, no network).
  extra     : (N=1 only, after the main timed region, each bounded) the other BASELINE configs measured in the same run:
              extra.reachao_rand (configs[1], box + sphere obstacles), extra.push (configs[3]: PandaPush, ee control through
              IK, 16 384 envs), extra.her (configs[4]: 1 M-transition on-device HER buffer, sample + relabel).
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
# Rank 0 must print exactly ONE JSON line on stdout.  Native libraries write there too (NCCL's "NCCL version ..." banner at
# communicator creation, at any NCCL_DEBUG level >= VERSION), so file descriptor 1 points at stderr for the whole run and is
# switched back only for the final line.
_REAL_STDOUT = os.dup(1)
os.dup2(2, 1)


def emit(line):
    sys.stdout.flush()
    os.dup2(_REAL_STDOUT, 1)
    print(json.dumps(line), flush=True)
    os.dup2(2, 1)


sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

# Algorithmic FLOPs per env-step of the step kernel AS IMPLEMENTED: 20 substeps of
# RNEA + CRBA + 7x7 Cholesky inverse + 50-sweep PGS + collider frames + link-obstacle tests, + one observation pass.
# Frozen from ncu per-opcode predicated-on thread-instruction counts (2*FFMA + FADD + FMUL) of one k_step launch at the stationary
# mix of episode phases the bench times (the 156th step; profiles/r02_k_step_traffic.json, r02_k_step_summary.md) -- the round-1
# figure, 0.191 MFLOP, was counted a few steps after a cold start, when every arm still stands near the neutral pose.
# A scenario WITHOUT a frozen count reports frac = null and says so (flops_source).
FLOPS_PER_ENV_STEP = {"reachao3": 0.1604e6}
BYTES_PER_ENV_STEP = 750.0       # ~308 B read + ~440 B written
HER_BYTES_PER_SAMPLE = 1080.0    # ~544 B read + ~532 B written per sampled transition


def clocks_sampler(stop, samples, index):
    q = "clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown," \
        "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap,power.draw"
    while not stop.is_set():
        try:
            out = subprocess.run(["nvidia-smi", "-i", str(index), "--query-gpu=" + q, "--format=csv,noheader,nounits"],
                                 capture_output=True, text=True, timeout=5).stdout.strip()
            if out:
                samples.append([x.strip() for x in out.split(",")])
        except Exception:
            pass
        stop.wait(0.2)


def summarize_clocks(samples):
    if not samples:
        return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
    sm = sorted(int(s[0]) for s in samples if s[0].isdigit())
    reasons = set()
    for s in samples:
        for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), s[2:6]):
            if v.lower().startswith("active"):
                reasons.add(name)
    power = []
    for s in samples:
        try:
            power.append(float(s[6]))
        except Exception:
            pass
    return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": int(samples[0][1]) if samples[0][1].isdigit() else None,
            "reasons": sorted(reasons), "power_w_max": max(power) if power else None, "samples": len(samples)}


class Clocks:
    """nvidia-smi sampler around a timed region."""

    def __init__(self, index, enabled=True):
        self.stop, self.samples, self.enabled = threading.Event(), [], enabled
        self.th = threading.Thread(target=clocks_sampler, args=(self.stop, self.samples, index), daemon=True)

    def __enter__(self):
        if self.enabled:
            self.th.start()
        return self

    def __exit__(self, *a):
        self.stop.set()
        if self.enabled:
            self.th.join(timeout=6)

    def summary(self):
        return summarize_clocks(self.samples)


def workload_name(scenario, n):
    return ("BASELINE configs[1]: PandaReachAO-v3 collision-avoidance reach, scenario %s, %d envs/GPU, joints control, "
            "sparse reward + collision penalty -100, TimeLimit 100, 20 substeps, auto-reset" % (scenario, n))


STAGES = ["reachao1", "reachao2", "reachao3", "wangexp-3"]


def config_block(args, stage_switches=0):
    """The `config` object of the JSON line -- built by BOTH arms from the same arguments, so the driver's same_config check
    compares like with like (what differs between the arms is stated in cpu_baseline.sample, not here)."""
    n = args.envs_per_gpu
    return {"workload": (workload_name(args.scenario, n) if not args.curriculum else
                         workload_name("reachao1 -> reachao2 -> reachao3 -> wangexp-3 (curriculum)", n).replace("configs[1]", "configs[2]")),
            "envs_per_gpu": n, "scenario": args.scenario, "sharding": "envs partitioned contiguously, no data-path collective; statistics all-reduce every 10 steps on a side stream",
            "l2": "flushed between timed steps (256 MiB write); per-step CUDA-event timing on the launch stream",
            "actions": "U(-1,1), counter-based, pre-generated in HBM",
            "preroll_steps": args.preroll,
            "curriculum": ({"stages": STAGES, "every": args.stage_every,
                            "note": "BASELINE configs[2]: stage parameter block swapped (and every env reset) inside the timed region"}
                           if args.curriculum else None)}


def cpu_reference_line(args, scenario):
    import oracle_lib as O
    O.build()
    cores = os.cpu_count() or 1
    O.cpu_rollout(20, cores, scenario=scenario, max_episode_steps=100)  # warm-up
    total_steps = 0
    t_all = 0.0
    for k in range(max(1, args.warmup)):
        O.cpu_rollout(40, cores, scenario=scenario, max_episode_steps=100, seed=k)
    for k in range(args.steps):
        rate, dt, _ = O.cpu_rollout(args.ref_steps_per_env, cores, scenario=scenario, max_episode_steps=100, seed=100 + k)
        total_steps += args.ref_steps_per_env * cores
        t_all += dt
    value = total_steps / t_all
    sample = ("reference arm = CPU restatement of the reference path (PyBullet itself is not installable here); per timed step a "
              "bounded sample of the workload: %d envs (one per host thread) x %d env-steps, %d timed steps, random actions, "
              "auto-reset" % (cores, args.ref_steps_per_env, args.steps))
    return {
        "impl": "reference", "metric": "env-steps/sec (whole box)", "value": value, "unit": "env-steps/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * t_all / args.steps, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": config_block(args),
        "cpu_baseline": {"value": value, "unit": "env-steps/s", "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": value, "unit": "env-steps/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }


def timed_steps(torch, b, actions, W, K, flush, extra_per_step=None):
    """W warm-up + K timed vector steps, one CUDA-event pair per step on the launch stream, L2 flushed between steps.
    Returns the summed device milliseconds."""
    for k in range(W):
        b.step(actions[k])
    torch.cuda.synchronize()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(K)]
    for k in range(K):
        flush.fill_(k & 0xFF)  # L2 flush, outside the per-step event pair
        ev[k][0].record()
        if extra_per_step is not None:
            extra_per_step(k)
        b.step(actions[W + k])
        ev[k][1].record()
    torch.cuda.synchronize()
    return sum(a.elapsed_time(bb) for a, bb in ev)


def extra_env_record(torch, local, env_id, scenario, n, K, W, flush, fp32_peak, hbm_peak, label, preroll=0, **kw):
    """One bounded extra record: another BASELINE config timed like the main one (device events, L2 flush, resident actions)."""
    from panda_gym_b200.vec_env import PandaVecEnv
    env = PandaVecEnv(env_id, n_envs=n, seed=4321, scenario=scenario, device=local, max_episode_steps=kw.pop("max_episode_steps", 100), **kw)
    b = env.batch
    b.reset()
    for k in range(preroll):  # untimed: to the stationary mix of episode phases (Push: arms in contact with the object)
        b.step(b.random_actions(555, k))
    actions = [b.random_actions(977, k) for k in range(K + W)]
    l0 = b.launch_count()
    with Clocks(local) as ck:
        ms = timed_steps(torch, b, actions, W, K, flush)
    launches = b.launch_count() - l0
    flops = FLOPS_PER_ENV_STEP.get(label)
    ach = (flops * n / (ms / K * 1e-3) / 1e12) if flops else None
    hbm = BYTES_PER_ENV_STEP * n / (ms / K * 1e-3) / 1e9
    rec = {"workload": label, "env_id": env_id, "scenario": scenario, "envs": n, "steps": K, "warmup": W, "preroll_steps": preroll, "ms_per_step": ms / K,
           "value": n * K / (ms * 1e-3), "unit": "env-steps/s", "gpu_launches": launches * K // (K + W),
           "roofline": {"bound": "fp32", "achieved": ach, "peak": fp32_peak, "unit": "TFLOP/s",
                        "frac": (ach / fp32_peak) if (ach and fp32_peak) else None,
                        "flops_source": ("frozen ncu opcode count" if flops else
                                         "none: no frozen FLOP count for this workload, frac not claimed"),
                        "hbm": {"achieved": hbm, "peak": hbm_peak, "unit": "GB/s", "frac": hbm / hbm_peak}},
           "episode_stats": env.episode_stats(), "clocks": ck.summary()}
    env.close()
    return rec


def extra_her_record(torch, local, hbm_peak, reps=20):
    """BASELINE configs[4]: 1 048 576-transition buffer x 1 024 env columns filled with synthetic episodes U{10..100} (seed 7),
    sample + future-goal relabel + reward recompute on the device.  HBM-bound: 1.08 KB per sampled transition."""
    import numpy as np
    from panda_gym_b200.her import HerReplayBuffer
    from panda_gym_b200.vec_env import PandaVecEnv
    E, cap, od, ad = 1024, 1 << 20, 56, 7
    env = PandaVecEnv("PandaReachAO-v3", n_envs=E, seed=3, scenario="reachao3", device=local, max_episode_steps=100)
    env.reset()
    her = HerReplayBuffer(cap, env, n_sampled_goal=4, max_batch_size=1 << 20, rng="device", device_seed=5)
    dev = env.batch.device
    g = torch.Generator(device=dev).manual_seed(7)
    rng = np.random.default_rng(7)
    remaining = rng.integers(10, 101, E)
    for _ in range(1024 + 64):
        blk = torch.rand((E, 2 * od + 12 + ad + 1), generator=g, device=dev)
        remaining -= 1
        done = remaining == 0
        remaining[done] = rng.integers(10, 101, int(done.sum()))
        d = torch.from_numpy(done).to(dev)
        her.add({"observation": blk[:, :od], "achieved_goal": blk[:, 2 * od:2 * od + 3], "desired_goal": blk[:, 2 * od + 3:2 * od + 6]},
                {"observation": blk[:, od:2 * od], "achieved_goal": blk[:, 2 * od + 6:2 * od + 9], "desired_goal": blk[:, 2 * od + 9:2 * od + 12]},
                blk[:, 2 * od + 12:2 * od + 12 + ad], blk[:, -1], d, timeouts=torch.zeros_like(d))
    out = {"workload": "BASELINE configs[4]: HER relabel+sample (k=4 future), 1 048 576-transition buffer x 1 024 env columns, obs 56",
           "buffer_transitions": cap, "env_columns": E, "batches": {}}
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    with Clocks(local) as ck:
        for B in (256, 65536, 1 << 20):
            for _ in range(3):
                her.sample(B)
            torch.cuda.synchronize()
            ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(reps)]
            for k in range(reps):
                flush.fill_(k & 0xFF)
                ev[k][0].record()
                her.sample(B)
                ev[k][1].record()
            torch.cuda.synchronize()
            ms = sum(a.elapsed_time(b) for a, b in ev) / reps
            gbs = HER_BYTES_PER_SAMPLE * B / (ms * 1e-3) / 1e9
            out["batches"][str(B)] = {"ms_per_call": ms, "samples_per_s": B / (ms * 1e-3),
                                      "roofline": {"bound": "hbm", "achieved": gbs, "peak": hbm_peak, "unit": "GB/s", "frac": gbs / hbm_peak,
                                                   "bytes_per_sample": HER_BYTES_PER_SAMPLE}}
    # numpy-exact index mode (bit-identical to SB3's np.random stream), the mode the parity tests pin; host-timed (it syncs)
    her.rng_mode = "numpy"
    np.random.seed(0)
    for B in (256, 65536):
        her.sample(B)
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for _ in range(5):
            her.sample(B)
        torch.cuda.synchronize()
        out["batches"][str(B)]["numpy_exact_ms_per_call"] = (time.perf_counter() - t0) / 5 * 1e3
    out["clocks"] = ck.summary()
    out["l2"] = "flushed between timed calls (256 MiB write)"
    her.close(); env.close()
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--scenario", default="reachao3")
    ap.add_argument("--envs-per-gpu", type=int, default=65536)
    ap.add_argument("--ref-steps-per-env", type=int, default=100)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--preroll", type=int, default=150,
                    help="untimed vector steps before the warm-up: freshly reset arms all stand at the neutral pose and step ~25 %% faster "
                         "than the stationary mix of episode phases the workload settles into after ~1.5 mean episode lengths")
    ap.add_argument("--no-extra", action="store_true", help="skip the bounded extra records (other BASELINE configs)")
    ap.add_argument("--curriculum", action="store_true",
                    help="BASELINE configs[2]: cycle the stage block reachao1 -> reachao2 -> reachao3 -> wangexp-3 every "
                         "--stage-every vector steps inside the timed region (pg_set_stage; use --steps >= 4 x --stage-every)")
    ap.add_argument("--stage-every", type=int, default=50)
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if args.curriculum:
        args.scenario = "reachao1"  # first stage of the cycle

    if args.impl == "reference":
        if rank == 0:
            emit(cpu_reference_line(args, args.scenario))
        return 0

    import numpy as np
    import torch
    import torch.distributed as dist
    import __graft_entry__ as ge
    if not os.path.exists(ge.LIB):
        ge.build_cuda()
    from panda_gym_b200 import _lib, sharding
    from panda_gym_b200.vec_env import PandaVecEnv

    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device (no CPU fallback)")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    # one process per GPU: run on (and take page-locked memory from) the NUMA node the GPU hangs off, BEFORE any pinned
    # allocation -- eight ranks pushing 17 MB per step through the far socket is what held the 8-GPU end-to-end number back
    numa = sharding.bind_to_local_numa(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    N = args.envs_per_gpu
    W = max(3, args.warmup)
    K = args.steps

    env = PandaVecEnv("PandaReachAO-v3", n_envs=N, seed=1234, scenario=args.scenario, device=local, max_episode_steps=100,
                      first_env_index=rank * N)
    b = env.batch
    lib = _lib.load()
    b.reset()
    # inputs resident in HBM before the timed region: K + W pre-generated action batches
    actions = [b.random_actions(1234 + rank, k) for k in range(K + W)]
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)  # > 126 MB L2
    stats_buf = b.view("stats")
    for k in range(args.preroll):  # to the stationary mix of episode phases (untimed, own action batches)
        b.step(b.random_actions(99 + rank, k))
    for k in range(W):
        b.step(actions[k])
    astats = sharding.AsyncStats(dev) if world > 1 else None
    if world > 1:  # warm the one collective of the path up too (communicator / first-call cost stays out of the timed region)
        for _ in range(2):
            astats.start(stats_buf); astats.result()
    torch.cuda.synchronize()

    ck = Clocks(local, enabled=(rank == 0))
    ck.__enter__()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(K)]
    launches0 = b.launch_count()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    t_wall0 = time.perf_counter()
    stage_switches = 0
    for k in range(K):
        flush.fill_(k & 0xFF)  # L2 flush, outside the per-step event pair
        ev[k][0].record()
        if args.curriculum and k > 0 and k % args.stage_every == 0:  # inside the timed region: the switch (+ its reset) is part of the cost
            env.batch.set_stage(__import__("panda_gym_b200.sim", fromlist=["x"]).stage_from_scenario(
                STAGES[(k // args.stage_every) % len(STAGES)], 100))
            stage_switches += 1
        b.step(actions[W + k])
        if world > 1 and (k % 10) == 9:  # episode statistics all-reduce: the one collective of the path, on a side stream
            if k > 9:
                astats.result()                # the previous one finished long ago: a host-side wait of nothing
            astats.start(stats_buf)
        ev[k][1].record()
    if world > 1:
        astats.result()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    t_wall = time.perf_counter() - t_wall0
    launches = b.launch_count() - launches0
    step_ms = [a.elapsed_time(bb) for a, bb in ev]
    dev_ms = sum(step_ms)
    t = torch.tensor([dev_ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    dev_ms = float(t.item())
    ms_per_step = dev_ms / K
    value = world * N * K / (dev_ms * 1e-3)
    if args.curriculum:  # back to a fixed stage for the end-to-end loop
        env.set_stage("reachao3", max_episode_steps=100)

    # ---- e2e: public host API, pinned host actions in, host numpy results out, every step
    # a DIFFERENT action batch every step, like the device-timed loop: cycling a handful of batches is a periodic drive that
    # walks the arms into their joint limits (the solver's slow path) and would time a different workload
    host_actions = [a.cpu().numpy() for a in actions]
    for k in range(3):  # warm-up of the whole host path, the consumers of `infos` included
        obs, rew, dones, infos = env.step(host_actions[k % len(host_actions)])
        if len(infos.finished):
            infos.terminal_observations(); infos.episode_returns.sum()
    torch.cuda.synchronize()
    Ke = max(10, K // 2)
    if world > 1:
        dist.barrier()
    consumed = 0.0
    t0 = time.perf_counter()
    for k in range(Ke):
        # the policy's output lands in the environment's page-locked action buffer (PandaVecEnv.action_buffer); the copy into
        # it stands in for the policy writing there and is part of the timed step
        abuf = env.action_buffer()
        np.copyto(abuf, host_actions[(k + 3) % len(host_actions)])
        obs, rew, dones, infos = env.step(abuf)
        # what a Monitor + replay buffer take from `infos` every step, in vectorised form: which episodes ended, their return and
        # length, whether the ending was a truncation, and the terminal observations (a device gather + copy of those rows)
        fin = infos.finished
        if len(fin):
            term = infos.terminal_observations()
            consumed += float(infos.episode_returns.sum()) + float(infos.episode_lengths.sum()) + float(term["observation"][0, 0]) \
                + float(infos.time_limit_truncated[fin].sum())
    torch.cuda.synchronize()
    e2e_s = time.perf_counter() - t0
    te = torch.tensor([e2e_s], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(te, op=dist.ReduceOp.MAX)
    e2e_value = world * N * Ke / float(te.item())
    h2d = N * b.act_dim * 4
    d2h = b.host_layout()[1]
    ck.__exit__()

    # ---- roofline of the step kernel (the only kernel that matters: k_step; k_reset rides along for done envs)
    tf = _lib.ctypes.c_double(0.0)
    _lib.check(lib.pg_debug_fma_peak(local, 148 * 8, 1 << 15, _lib.ctypes.byref(tf)))
    fp32_peak = float(tf.value)
    flops = None if args.curriculum else FLOPS_PER_ENV_STEP.get(args.scenario)
    achieved_tflops = (flops * N / (ms_per_step * 1e-3) / 1e12) if flops else None
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    hbm_peak = float(peaks.get("hbm_gbs", 6650.0))
    hbm_achieved = BYTES_PER_ENV_STEP * N / (ms_per_step * 1e-3) / 1e9
    stats = env.episode_stats()
    traffic = None  # dram__bytes_read + dram__bytes_write of one k_step launch, from the committed ncu --set full capture
    try:
        with open(os.path.join(ROOT, "profiles", "r02_k_step_traffic.json")) as f:
            tj = json.load(f)
        if args.scenario == "reachao3" and N == 65536 and not args.curriculum:
            traffic = tj["traffic_bytes_per_launch"]
    except Exception:
        pass

    if rank == 0:
        cfg = config_block(args, stage_switches)
        if args.curriculum:
            cfg["curriculum"]["switches"] = stage_switches
        line = {
            "metric": "env-steps/sec (whole box)", "value": value, "unit": "env-steps/s", "n_gpus": world, "steps": K, "warmup": W,
            "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
            "data": "synthetic",
            "config": cfg,
            "e2e": {"value": e2e_value, "unit": "env-steps/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                    "api": "PandaVecEnv.step(numpy actions) -> numpy obs/reward/done (pg_step_host: one H2D + ONE D2H copy of the "
                           "result slab into pinned memory) + infos consumed vectorised (finished episodes' r / l / truncation "
                           "flags / terminal observations fetched every step)",
                    "steps": Ke, "d2h_gbs_per_rank": d2h * Ke / float(te.item()) / 1e9},
            "gpu_launches": launches,
            "clocks": ck.summary(),
            "roofline": {"bound": "fp32", "achieved": achieved_tflops, "peak": fp32_peak, "unit": "TFLOP/s",
                         "frac": (achieved_tflops / fp32_peak) if (achieved_tflops and fp32_peak) else None, "traffic": traffic,
                         "kernel": "k_step", "flops_per_env_step": flops,
                         "flops_source": ("frozen ncu opcode count for %s at steady state (profiles/r02_k_step_traffic.json)" % args.scenario) if flops else
                                         "none: no frozen FLOP count for this scenario, frac not claimed",
                         "peak_source": "FP32 FMA micro-benchmark measured in this run (pg_debug_fma_peak); MEASURED_PEAKS.json has no FP32 figure",
                         "hbm": {"achieved": hbm_achieved, "peak": hbm_peak, "unit": "GB/s", "frac": hbm_achieved / hbm_peak,
                                 "peak_source": "MEASURED_PEAKS.json" if peaks else "fallback 6650"}},
            "episode_stats": stats,
            "step_ms": {"min": min(step_ms), "median": sorted(step_ms)[len(step_ms) // 2], "max": max(step_ms),
                        "first_tenth_mean": sum(step_ms[:max(1, K // 10)]) / max(1, K // 10),
                        "last_tenth_mean": sum(step_ms[-max(1, K // 10):]) / max(1, K // 10)},
            "wall_s_timed_region": t_wall,
            "numa": numa,
        }
    env.close()
    del actions, host_actions
    if rank == 0:
        if world == 1 and not args.no_extra:
            extra = {}
            Kx, Wx = 10, 3
            try:
                extra["reachao_rand"] = extra_env_record(torch, local, "PandaReachAO-v3", "reachao_rand", N, Kx, Wx, flush, fp32_peak, hbm_peak,
                                                         "reachao_rand", preroll=min(args.preroll, 100))
                extra["push"] = extra_env_record(torch, local, "PandaPush-v3", None, 16384, Kx, Wx, flush, fp32_peak, hbm_peak, "push",
                                                 preroll=min(args.preroll, 75), max_episode_steps=50)
                del flush
                extra["her"] = extra_her_record(torch, local, hbm_peak)
            except Exception as ex:  # an extra record must never cost the main line
                extra["error"] = repr(ex)
            line["extra"] = extra
        if world == 1 and not args.no_cpu_baseline:
            import oracle_lib as O
            O.build()
            cores = os.cpu_count() or 1
            O.cpu_rollout(20, cores, scenario=args.scenario, max_episode_steps=100)
            target = 12.0  # seconds of CPU work
            rate, dt, _ = O.cpu_rollout(200, cores, scenario=args.scenario, max_episode_steps=100, seed=1)
            per_env = max(200, int(200 * target / max(dt, 1e-3)))
            rate, dt, _ = O.cpu_rollout(per_env, cores, scenario=args.scenario, max_episode_steps=100, seed=2)
            line["cpu_baseline"] = {"value": rate, "unit": "env-steps/s", "cores": cores, "kind": "port",
                                    "sample": "%d envs (one per host thread) x %d env-steps of the same workload, %.1f s" % (cores, per_env, dt)}
        emit(line)
    if world > 1:
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
