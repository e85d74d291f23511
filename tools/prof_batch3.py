"""Replica scaling of the native batch sweep: every rank runs the same sweep on its own GPU at the same time."""
import time, os, sys
import numpy as np, torch
sys.path.insert(0, os.getcwd())
import torch.distributed as dist
rank = int(os.environ.get("RANK", "0")); world = int(os.environ.get("WORLD_SIZE", "1"))
torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", "0")))
if world > 1:
    dist.init_process_group("nccl", device_id=torch.device("cuda", torch.cuda.current_device()))
import calib_b200
from calib_b200 import ba, scene
scenes = []
for d in range(16):
    sc = scene.cfg1(seed=7000 + d, N=1000)
    scenes.append((sc.camid, sc.K, sc.R_init, sc.t_init) + tuple(scene.ba_inputs(sc)))
probs = [scenes[k % 16] for k in range(768)]
ba.ba_batch(probs[:48], 1.0, 1.0, n_streams=12)
if rank == 0:
    print("cpus", os.cpu_count(), "affinity", len(os.sched_getaffinity(0)), flush=True)
for nw in [int(x) for x in os.environ.get("WORKERS", "4,8,12").split(",")]:
    if world > 1:
        dist.barrier()
    os.environ["BAX_BATCH_FORCE_WORKERS"] = str(nw)
    t0 = time.perf_counter(); ba.ba_batch(probs, 1.0, 1.0, n_streams=nw); dt = time.perf_counter() - t0
    print(f"rank {rank} workers {nw}: {len(probs) / dt:.0f} problems/s", flush=True)
if world > 1:
    dist.destroy_process_group()
