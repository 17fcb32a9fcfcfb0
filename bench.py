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
              copied back to host numpy every step)
  roofline  : binding roofline of the step kernel is FP32 FMA issue (no dense contraction => tensor cores unused,
              HBM traffic ~1 KB/env-step is <1% of peak); achieved = algorithmic FLOPs/launch / kernel time,
              peak = FP32 FMA micro-benchmark measured live in this process (MEASURED_PEAKS.json has no FP32 figure).
              The HBM view of the same launch is reported alongside against MEASURED_PEAKS.json's hbm_gbs.
  cpu_baseline : the CPU oracle (kind "port": C restatement of the reference path, oracle/) on the host cores,
              bounded sample, rank 0 / N=1 only.  --impl reference times that same CPU implementation with all host
              threads (the reference's PyBullet backend is not installable here: no wheel, no network).
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

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

# algorithmic FLOPs per env-step of the step kernel AS IMPLEMENTED: 20 substeps of
# RNEA + CRBA + 7x7 Cholesky inverse + 50-sweep PGS + collider frames + link-obstacle tests, + one observation pass.
# Frozen from ncu per-opcode thread-instruction counts of a real rollout launch (2*FFMA + FADD + FMUL), see
# profiles/r01_k_step_summary.md.  Scenarios without a frozen count fall back to the reachao3 figure.
FLOPS_PER_ENV_STEP = {"reachao3": 0.191e6}
BYTES_PER_ENV_STEP = 750.0  # ~308 B read + ~440 B written


def clocks_sampler(stop, samples, index):
    q = "clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown," \
        "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"
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
    return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": int(samples[0][1]) if samples[0][1].isdigit() else None,
            "reasons": sorted(reasons)}


def workload_name(scenario, n):
    return ("BASELINE configs[1]: PandaReachAO-v3 collision-avoidance reach, scenario %s, %d envs/GPU, joints control, "
            "sparse reward + collision penalty -100, TimeLimit 100, 20 substeps, auto-reset" % (scenario, n))


def cpu_reference_line(args, scenario):
    import oracle_lib as O
    O.build()
    cores = os.cpu_count() or 1
    O.cpu_rollout(20, cores, scenario=scenario, max_episode_steps=100)  # warm-up
    per_step = []
    total_steps = 0
    t_all = 0.0
    for k in range(max(1, args.warmup)):
        O.cpu_rollout(40, cores, scenario=scenario, max_episode_steps=100, seed=k)
    for k in range(args.steps):
        rate, dt, _ = O.cpu_rollout(args.ref_steps_per_env, cores, scenario=scenario, max_episode_steps=100, seed=100 + k)
        per_step.append(dt)
        total_steps += args.ref_steps_per_env * cores
        t_all += dt
    value = total_steps / t_all
    sample = "%d envs (one per host thread) x %d env-steps per timed step, %d timed steps, random actions, auto-reset" % (
        cores, args.ref_steps_per_env, args.steps)
    return {
        "impl": "reference", "metric": "env-steps/sec (whole box)", "value": value, "unit": "env-steps/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * t_all / args.steps, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": workload_name(scenario, args.envs_per_gpu), "envs_per_gpu": args.envs_per_gpu, "scenario": scenario,
                   "note": "reference arm: CPU restatement of the reference path (PyBullet itself is not installable here), "
                           "a bounded sample of the workload per timed step"},
        "cpu_baseline": {"value": value, "unit": "env-steps/s", "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": value, "unit": "env-steps/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }


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
    ap.add_argument("--curriculum", action="store_true",
                    help="BASELINE configs[2]: cycle the stage block reachao1 -> reachao2 -> reachao3 -> wangexp-3 every "
                         "--stage-every vector steps inside the timed region (pg_set_stage; use --steps >= 4 x --stage-every)")
    ap.add_argument("--stage-every", type=int, default=50)
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))

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
    from panda_gym_b200 import _lib, sim
    from panda_gym_b200.vec_env import PandaVecEnv

    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device (no CPU fallback)")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    if args.curriculum:
        args.scenario = "reachao1"  # first stage of the cycle
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
    for k in range(W):
        b.step(actions[k])
    torch.cuda.synchronize()

    stop = threading.Event(); samples = []
    th = threading.Thread(target=clocks_sampler, args=(stop, samples, local), daemon=True)
    if rank == 0:
        th.start()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(K)]
    launches0 = b.launch_count()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    t_wall0 = time.perf_counter()
    stages = ["reachao1", "reachao2", "reachao3", "wangexp-3"]
    stage_switches = 0
    for k in range(K):
        flush.fill_(k & 0xFF)  # L2 flush, outside the per-step event pair
        ev[k][0].record()
        if args.curriculum and k > 0 and k % args.stage_every == 0:  # inside the timed region: the switch is part of the cost
            env.set_stage(stages[(k // args.stage_every) % len(stages)], max_episode_steps=100)
            stage_switches += 1
        b.step(actions[W + k])
        if world > 1 and (k % 10) == 9:  # episode statistics all-reduce: the one collective of the path
            s = stats_buf.clone()
            dist.all_reduce(s)
        ev[k][1].record()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    t_wall = time.perf_counter() - t_wall0
    launches = b.launch_count() - launches0
    dev_ms = sum(a.elapsed_time(bb) for a, bb in ev)
    t = torch.tensor([dev_ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    dev_ms = float(t.item())
    ms_per_step = dev_ms / K
    value = world * N * K / (dev_ms * 1e-3)

    # ---- e2e: public host API, pinned host actions in, host numpy results out, every step
    # a DIFFERENT action batch every step, like the device-timed loop: cycling a handful of batches is a periodic drive that
    # walks the arms into their joint limits (the solver's slow path) and would time a different workload
    host_actions = [a.cpu().numpy() for a in actions]
    for k in range(3):
        env.step(host_actions[k % len(host_actions)])
    torch.cuda.synchronize()
    Ke = max(10, K // 2)
    if world > 1:
        dist.barrier()
    t0 = time.perf_counter()
    for k in range(Ke):
        obs, rew, dones, infos = env.step(host_actions[(k + 3) % len(host_actions)])
    torch.cuda.synchronize()
    e2e_s = time.perf_counter() - t0
    te = torch.tensor([e2e_s], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(te, op=dist.ReduceOp.MAX)
    e2e_value = world * N * Ke / float(te.item())
    h2d = N * b.act_dim * 4
    d2h = N * (b.obs_dim + 3 + 3 + 1) * 4 + 2 * N
    stop.set()

    # ---- roofline of the step kernel (the only kernel that matters: k_step; k_reset rides along for done envs)
    tf = _lib.ctypes.c_double(0.0)
    _lib.check(lib.pg_debug_fma_peak(local, 148 * 8, 1 << 15, _lib.ctypes.byref(tf)))
    fp32_peak = float(tf.value)
    flops = FLOPS_PER_ENV_STEP.get(args.scenario.split("-")[0], 0.191e6)
    achieved_tflops = flops * N / (ms_per_step * 1e-3) / 1e12
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
        with open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "profiles", "r01_k_step_traffic.json")) as f:
            tj = json.load(f)
        if args.scenario == "reachao3" and N == 65536:
            traffic = tj["traffic_bytes_per_launch"]
    except Exception:
        pass

    if rank == 0:
        line = {
            "metric": "env-steps/sec (whole box)", "value": value, "unit": "env-steps/s", "n_gpus": world, "steps": K, "warmup": W,
            "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
            "data": "synthetic",
            "config": {"workload": (workload_name(args.scenario, N) if not args.curriculum else
                                    workload_name("reachao1 -> reachao2 -> reachao3 -> wangexp-3 (curriculum)", N).replace("configs[1]", "configs[2]")),
                       "envs_per_gpu": N, "scenario": args.scenario, "sharding": "envs partitioned contiguously, no data-path collective",
                       "l2": "flushed between timed steps (256 MiB write); per-step CUDA-event timing on the launch stream",
                       "actions": "U(-1,1), counter-based, pre-generated in HBM",
                       "curriculum": ({"stages": stages, "every": args.stage_every, "switches": stage_switches,
                                       "note": "BASELINE configs[2]: stage parameter block swapped inside the timed region"}
                                      if args.curriculum else None)},
            "e2e": {"value": e2e_value, "unit": "env-steps/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                    "api": "PandaVecEnv.step(numpy actions) -> numpy obs/reward/done (pg_step_host, pinned buffers)", "steps": Ke},
            "gpu_launches": launches,
            "clocks": summarize_clocks(samples),
            "roofline": {"bound": "fp32", "achieved": achieved_tflops, "peak": fp32_peak, "unit": "TFLOP/s",
                         "frac": achieved_tflops / fp32_peak if fp32_peak else None, "traffic": traffic, "kernel": "k_step",
                         "flops_per_env_step": flops,
                         "peak_source": "FP32 FMA micro-benchmark measured in this run (pg_debug_fma_peak); MEASURED_PEAKS.json has no FP32 figure",
                         "hbm": {"achieved": hbm_achieved, "peak": hbm_peak, "unit": "GB/s", "frac": hbm_achieved / hbm_peak,
                                 "peak_source": "MEASURED_PEAKS.json" if peaks else "fallback 6650"}},
            "episode_stats": stats,
            "wall_s_timed_region": t_wall,
        }
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
