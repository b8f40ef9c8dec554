"""Eager (no CUDA graph) steps of the flagship batch for ncu launch lists: python scripts/profile_step.py [steps] [batch]"""
import sys, numpy as np, torch
sys.path.insert(0, '.')
from distributedvolunteercomputing_b200.config import Config
from distributedvolunteercomputing_b200.engine import NativeEngine
from distributedvolunteercomputing_b200.models.mobilenet_ssd import SSDGraph
steps = int(sys.argv[1]) if len(sys.argv) > 1 else 3
B = int(sys.argv[2]) if len(sys.argv) > 2 else 64
sub = int(sys.argv[3]) if len(sys.argv) > 3 else 0
g = SSDGraph.random(seed=1, conf_gain=2.5)
e = NativeEngine(g, "cuda:0", Config(batch=B, sub_batch=sub, use_cuda_graph=False), requester_id="127.0.0.1:5554")
st = e.state(B, 300, 300)
st.raw.copy_(torch.from_numpy(np.random.default_rng(0).integers(0, 256, (B, 300, 300, 3), dtype=np.uint8)))
for i in range(steps):
    e.run(st)
torch.cuda.synchronize()
print("ops per step", e.launches_per_batch(st))

if "mbx" in sys.argv:
    # the same batch through the mailbox data plane (client GPU as its own worker, eager): post -> wait -> K1a -> plan
    # -> K10 -> done-wait, so the launch list shows the flow-control kernels in place
    from distributedvolunteercomputing_b200.parallel.mailbox import ClientBox, ElasticBox, WorkerLanes
    client = ClientBox("cuda:0", B, 300, 300, max_workers=1, lanes=1, engine=e, done_timeout_s=1.0)
    client.attach_local(0, WorkerLanes(e, client.layout, *client.bases(), worker=0))
    box = ElasticBox(client)
    box.add_worker(0)
    slot = client.alloc_slot()
    client.slot_tensor(slot).copy_(st.raw)
    for i in range(2):
        box.submit_slot(slot, range(B))
        assert box.drain(10.0)
    torch.cuda.synchronize()
    print("mailbox steps ok", dict(box.stats))
