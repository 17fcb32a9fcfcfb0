"""Device-to-host bandwidth of the step's 16.6 MB result copy when N ranks copy at once (torchrun, one rank per GPU):
each rank alone in turn, then all together, then the first half / every second rank together.  Prints `nvidia-smi topo -m`."""
import os, subprocess, time
import torch, torch.distributed as dist
rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(local)
if world > 1:
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
NB = 16646144
src = torch.empty(NB, dtype=torch.uint8, device="cuda")
dst = torch.empty(NB, dtype=torch.uint8).pin_memory()
def bar():
    if world > 1: dist.barrier()
def run(active, reps=60):
    bar()
    ms = 0.0
    if active:
        for _ in range(5): dst.copy_(src, non_blocking=True)
        torch.cuda.synchronize()
    bar()
    if active:
        t0 = time.perf_counter()
        for _ in range(reps):
            dst.copy_(src, non_blocking=True); torch.cuda.synchronize()
        ms = (time.perf_counter() - t0) / reps * 1e3
    bar()
    t = torch.tensor([ms], device="cuda")
    if world > 1:
        out = [torch.zeros(1, device="cuda") for _ in range(world)]
        dist.all_gather(out, t)
        return [float(o.item()) for o in out]
    return [ms]
def show(name, res):
    if rank == 0:
        print(name, " ".join("%.1f" % (NB / (m * 1e-3) / 1e9) if m > 0 else "-" for m in res), "GB/s per rank", flush=True)
if rank == 0:
    print(subprocess.run(["nvidia-smi", "topo", "-m"], capture_output=True, text=True).stdout[:3000], flush=True)
for r in range(world):
    show("alone rank %d:" % r, run(rank == r))
show("all together:", run(True))
if world >= 4:
    show("first half  :", run(rank < world // 2))
    show("even ranks  :", run(rank % 2 == 0))
    show("ranks 0,1   :", run(rank < 2))
if world > 1:
    dist.destroy_process_group()
