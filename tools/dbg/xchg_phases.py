"""torchrun --nproc-per-node G: finalize phase clocks (SM cycles) of the sharded step with ranks aligned by an all-reduce."""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch, torch.distributed as dist
import motionplanning_b200 as m
from motionplanning_b200 import dist as mdist, scenarios as sc
capi = m.capi
rank, world, lr = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(lr)
dist.init_process_group("nccl", device_id=torch.device("cuda", lr))
K = 65536 * world
args, kw, circles, blocks = sc.config(1, K=K)
off, kl = mdist.shard_range(K, rank, world)
kw = dict(kw, device=lr, precision=capi.FP32, noise_mode=capi.NOISE_PHILOX, k_offset=off, K_local=kl, FeasibilityCount=K)
mp = m.defineMPPI(*args, **kw); m.defineMPPIobs(mp, circles)
if world > 1: mdist.init_peer_exchange(mp)
flush = torch.empty(256 * 1024 * 1024, dtype=torch.uint8, device="cuda")
st = torch.cuda.ExternalStream(mp.stream(), device=torch.device("cuda", lr))
sync_t = torch.zeros(1, device="cuda")
for _ in range(10): mp.step_async(adopt=False)
mp.sync()
ph, dev = [], []
for i in range(60):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with torch.cuda.stream(st):
        flush.zero_()
        if world > 1: dist.all_reduce(sync_t)
    e0.record(st); mp.step_async(adopt=False); e1.record(st)
    mp.sync()
    c = mp.phase_clocks()
    ph.append([c[j + 1] - c[j] for j in range(6)]); dev.append(e0.elapsed_time(e1) * 1e3)
ph = np.median(np.array(ph[10:]), axis=0); 
print(f"rank {rank}/{world}: step {np.median(dev[10:]):.1f} us; finalize phases (cycles @1.965GHz) partial {ph[0]:.0f} xchg+combine {ph[1]:.0f} U_new {ph[2]:.0f} states {ph[3]:.0f} costing {ph[4]:.0f} sum/write {ph[5]:.0f} total {ph.sum()/1965:.2f} us", flush=True)
dist.barrier(); dist.destroy_process_group()
