import numpy as np, torch, sys, cv2
sys.path.insert(0, '.')
from distributedvolunteercomputing_b200.config import Config
from distributedvolunteercomputing_b200.engine import NativeEngine, CpuEngine
from distributedvolunteercomputing_b200.models.mobilenet_ssd import SSDGraph, TorchSSD
from distributedvolunteercomputing_b200.models import preprocess as P
g = SSDGraph.random(seed=1, conf_gain=2.5)
cfg = Config(sub_batch=4, use_cuda_graph=False)
e = NativeEngine(g, "cuda:0", cfg)
cpu = CpuEngine(g, cfg)
B = 4
rng = np.random.default_rng(9)
frames = np.stack([cv2.resize(rng.integers(0, 256, (300 // 16 + 1, 300 // 16 + 1, 3), dtype=np.uint8), (300, 300), interpolation=cv2.INTER_CUBIC) for _ in range(B)])
st = e.state(B, 300, 300)
st.raw.copy_(torch.from_numpy(frames))
e.run(st, annotate=False)
torch.cuda.synchronize()
f400_gpu = st.frames400.cpu().numpy()
f400_cpu = np.stack([P.resize_width_400(f) for f in frames])
print("f400 diff max", np.abs(f400_gpu.astype(int) - f400_cpu.astype(int)).max(), "frac", (f400_gpu != f400_cpu).mean())
acts = e.debug_layerwise(torch.from_numpy(f400_cpu).cuda())
blobs = np.concatenate([P.blob_from_image(P.resize_linear_u8(f, 300, 300), cfg.mean) for f in f400_cpu])
cap = {}
with torch.no_grad():
    rl, rc = TorchSSD(g)(torch.from_numpy(blobs), capture=cap)
print("plan loc vs layerwise(cpu f400)", (st.loc.view(B, -1, 4) - acts['loc'].view(B, -1, 4)).abs().max().item())
print("layerwise loc vs torch", (acts['loc'].view(B, -1, 4).cpu() - rl).abs().max().item(), "ref max", rl.abs().max().item())
print("plan loc vs torch", (st.loc.view(B, -1, 4).cpu() - rl).abs().max().item())
for c in g.body:
    ref = cap[c.name].permute(0, 2, 3, 1)
    got = acts[c.name].float().cpu()
    print(c.name, "err", round((got - ref).abs().max().item(), 4), "refmax", round(ref.abs().max().item(), 3))
