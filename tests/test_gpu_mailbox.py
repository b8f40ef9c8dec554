"""Device side of the mailbox protocol (csrc/mailbox.cu, parallel/mailbox.py) on ONE GPU: the client's own GPU serves as
worker 0 through the inline path (post -> wait -> K1a -> network -> K9 + done word -> wait -> copy-out on one stream), so
no kernel ever spins on something a later launch of the same GPU has to produce.  Oracle: the same engine's plain
`process()` on the same frames.  Multi-GPU and process-kill coverage is scripts/check_mailbox.py (tests/test_gpu_multi.py)."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

if not torch.cuda.is_available():
    pytest.skip("needs a GPU", allow_module_level=True)

from distributedvolunteercomputing_b200.config import Config
from distributedvolunteercomputing_b200.engine import NativeEngine
from distributedvolunteercomputing_b200.models.mobilenet_ssd import SSDGraph
from distributedvolunteercomputing_b200.parallel.mailbox import ClientBox, ElasticBox, WorkerLanes

B, H, W = 4, 120, 160


@pytest.fixture(scope="module")
def setup():
    graph = SSDGraph.random(seed=1, conf_gain=2.5)
    eng = NativeEngine(graph, "cuda:0", Config(batch=B, use_cuda_graph=True), requester_id="127.0.0.1:5554")
    client = ClientBox("cuda:0", B, H, W, max_workers=2, lanes=2, engine=eng, done_timeout_s=0.05)
    lanes = WorkerLanes(eng, client.layout, *client.bases(), worker=0)
    client.attach_local(0, lanes)
    yield eng, client, lanes
    client.close()


def frames(n, seed):
    rng = np.random.default_rng(seed)
    return torch.from_numpy(rng.integers(0, 256, (n, H, W, 3), dtype=np.uint8)).pin_memory()


def test_dispatch_equals_plain_engine(setup):
    eng, client, lanes = setup
    got = {}

    def on_result(batch, v, worker):
        got[batch.batch_id] = {k: a.copy() for k, a in v.items()}

    box = ElasticBox(client, on_result=on_result)
    box.add_worker(0)
    batches = [frames(B, 10 + i) for i in range(5)]
    ids = [box.submit_host(f, range(i * B, (i + 1) * B)).batch_id for i, f in enumerate(batches)]
    assert box.drain(20.0), box.stats
    assert sorted(got) == sorted(ids) and box.stats["timeouts"] == 0
    total = 0
    for bid, f in zip(ids, batches):
        ref = eng.process(f).numpy()
        g = got[bid]
        assert np.array_equal(g["det"], ref.det) and np.array_equal(g["det_count"], ref.det_count)
        assert np.array_equal(g["box_count"], ref.box_count)
        for k in range(B):
            assert np.array_equal(g["boxes_i"][k, :ref.box_count[k]], ref.boxes_i[k, :ref.box_count[k]])
        total += int(ref.det_count.sum())
    assert total > 0, "the comparison must see detections"
    # the last batch's annotated frames stay on the worker: lane of the last dispatch
    ref = eng.process(batches[-1]).numpy()
    assert any(np.array_equal(lanes.frames400(l).cpu().numpy(), ref.frames400) for l in range(2))
    box.remove_worker(0, stop=False)


def test_silent_worker_times_out_without_hanging(setup):
    eng, client, lanes = setup
    box = ElasticBox(client)
    box.add_worker(0)
    box.add_worker(1)                      # nobody serves worker 1: its done word never comes
    for i in range(6):
        box.submit_host(frames(B, 50 + i), range(B))
    assert box.drain(20.0), box.stats
    assert box.stats["completed"] == 6 and box.failed_workers == [1] and box.stats["redispatched"] >= 1
    k = 1 * client.layout.lanes
    assert int(client.hs[k]["wait_ns"]) >= 0.04e9, "the wait kernel ran into its deadline, it did not spin for ever"
    box.remove_worker(0, stop=False)


def test_stale_epoch_is_rejected_by_k9(setup):
    eng, client, lanes = setup
    box = ElasticBox(client)
    box.add_worker(0)
    stale0 = lanes.stale_count()
    client.epoch[0] -= 1                   # the descriptor will carry an epoch older than the live word
    slot = client.alloc_slot()
    client.slot_tensor(slot).copy_(frames(B, 99))
    dr = client.device_result(0, 0)
    dr["det_count"].fill_(-7)
    client.dispatch(0, 0, slot, B, 1234)
    torch.cuda.synchronize()
    assert client.poll(0, 0) == "timeout", "a batch whose frames were rejected never publishes its done word"
    assert lanes.stale_count() - stale0 == B
    assert (dr["det_count"] == -7).all(), "stale rows were delivered"
    client.epoch[0] += 1
    client.dispatch(0, 0, slot, B, 1235)   # the lane is usable again (its arrival counter went back to zero)
    torch.cuda.synchronize()
    assert client.poll(0, 0) == "ok" and (dr["det_count"] >= 0).all()
    client.free_slot(slot)
    box.remove_worker(0, stop=False)
