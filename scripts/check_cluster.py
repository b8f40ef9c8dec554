"""torchrun --nproc-per-node N scripts/check_cluster.py: the torchrun front-end (BoxCluster) of the mailbox data plane.
Every rank's results must equal the single-GPU engine bit for bit, every rank must have served, the non-client ranks
must run on the weights they pulled from the client (K13) and no collective may be needed inside a phase."""
import os, sys
import numpy as np, torch
import torch.distributed as dist
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from distributedvolunteercomputing_b200.config import Config
from distributedvolunteercomputing_b200.engine import NativeEngine
from distributedvolunteercomputing_b200.models.mobilenet_ssd import SSDGraph
from distributedvolunteercomputing_b200.parallel.cluster import BoxCluster

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dev = f"cuda:{local}"
dist.init_process_group("nccl", device_id=torch.device(dev))
B, H, W = 8, 240, 320
graph = SSDGraph.random(seed=1, conf_gain=2.5)
for use_graph in (False, True):
    eng = NativeEngine(graph, dev, Config(batch=B, sub_batch=4, use_cuda_graph=use_graph), requester_id="127.0.0.1:5554")
    if rank != 0:
        for t in eng.w.values():          # K13: scrambled here, adopted from the client's blob over NVLink at join
            t.zero_()
    got = {}
    cl = BoxCluster(eng, B, H, W, client_rank=0, lanes=2, done_timeout_s=5.0,
                    on_result=lambda b, v, w: got.__setitem__(b.batch_id, ({k: a.copy() for k, a in v.items()}, w)))
    rng = np.random.default_rng(7)
    pool = [torch.from_numpy(rng.integers(0, 256, (B, H, W, 3), dtype=np.uint8)).pin_memory() for _ in range(5)]
    for phase in range(2):                # two phases: leaving and re-joining members works under torchrun too
        cl.begin()
        if rank == 0:
            ids = [cl.box.submit_host(pool[k % 5], range(B)).batch_id for k in range(10 * world)]
            assert cl.box.drain(60.0), cl.box.stats
        else:
            n = cl.serve()
            assert n > 0, f"rank {rank} served nothing"
        cl.end()
        if rank == 0:
            ref = NativeEngine(graph, dev, Config(batch=B, sub_batch=4, use_cuda_graph=False), requester_id="127.0.0.1:5554")
            refs = [ref.process(f).numpy() for f in pool]
            per = {}
            for k, bid in enumerate(ids):
                g, w = got[bid]
                r = refs[k % 5]
                assert np.array_equal(g["det"], r.det), (use_graph, bid, w, "det")
                assert np.array_equal(g["det_count"], r.det_count) and np.array_equal(g["box_count"], r.box_count), (use_graph, bid, w)
                for i in range(B):
                    assert np.array_equal(g["boxes_i"][i, :r.box_count[i]], r.boxes_i[i, :r.box_count[i]]), (use_graph, bid, w, "boxes")
                per[w] = per.get(w, 0) + 1
            assert set(per) == set(range(world)), per
            assert cl.box.stats["timeouts"] == 0, cl.box.stats
            print(f"cluster check ok: world={world} graph={use_graph} phase={phase} batches per rank {dict(sorted(per.items()))} "
                  f"detections={sum(int(r.det_count.sum()) for r in refs)}", flush=True)
    dist.barrier()
    cl.close()
dist.destroy_process_group()
