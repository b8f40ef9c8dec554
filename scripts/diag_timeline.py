"""torchrun -n N scripts/diag_timeline.py: per-batch dispatch / completion timeline of a 20-step window on the mailbox cluster
(diagnosis aid for the ramp of short timed windows)."""
import os, sys, time, json
import numpy as np, torch, torch.distributed as dist
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from distributedvolunteercomputing_b200.config import Config
from distributedvolunteercomputing_b200.engine import NativeEngine
from distributedvolunteercomputing_b200.models.mobilenet_ssd import SSDGraph
from distributedvolunteercomputing_b200.parallel.cluster import BoxCluster
rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local); dev = f"cuda:{local}"
dist.init_process_group("nccl", device_id=torch.device(dev))
B, H, W, lanes, steps = 64, 300, 300, 4, 20
eng = NativeEngine(SSDGraph.random(seed=1, conf_gain=2.5), dev, Config(batch=B), requester_id="c")
cl = BoxCluster(eng, B, H, W, lanes=lanes, arena_slots=8 * world + world * lanes + 8, done_timeout_s=10.0, fetch=False)
ev = []
if rank == 0:
    slots = [cl.client.alloc_slot() for _ in range(8 * world)]
    for s in slots: cl.client.slot_tensor(s).random_(0, 255)
    torch.cuda.synchronize()
    orig = cl.client.dispatch
    def disp(w, l, *a, **k):
        ev.append(("d", time.perf_counter(), w, l)); return orig(w, l, *a, **k)
    cl.client.dispatch = disp
    cl.box.on_result = lambda b, v, w: ev.append(("c", time.perf_counter(), w, -1))
def run(n):
    sub = done = 0
    while done < n:
        while sub < n and cl.box.sched.pending() < 2 * world * lanes:
            cl.box.submit_slot(slots[sub % len(slots)], range(B)); sub += 1
        done += cl.box.pump()
for phase in range(3):
    cl.begin()
    if rank == 0:
        cl.box.fetch = False
        ev.clear(); torch.cuda.synchronize(); t0 = time.perf_counter()
        run(steps * world)
        t1 = time.perf_counter()
        cl.end()
        if phase == 2:
            d = [e for e in ev if e[0] == "d"]; c = [e for e in ev if e[0] == "c"]
            print(f"window {1e3*(t1-t0):.2f} ms for {steps} steps = {1e3*(t1-t0)/steps:.3f} ms/step; first wave ({world*lanes} dispatches) took {1e3*(d[world*lanes-1][1]-t0):.3f} ms; "
                  f"first completion at {1e3*(c[0][1]-t0):.3f} ms, last dispatch at {1e3*(d[-1][1]-t0):.3f} ms, last completion {1e3*(c[-1][1]-t0):.3f} ms")
            per = {}
            for _, t, w, _ in c: per.setdefault(w, []).append(1e3 * (t - t0))
            for w in sorted(per): print(f"  gpu{w}: {len(per[w])} batches, completions at", " ".join(f"{x:.2f}" for x in per[w]))
    else:
        cl.serve(); cl.end()
dist.barrier(); cl.close(); dist.destroy_process_group()
