"""Kernel timeline of one graphed training step (CUPTI through torch.profiler; no nsys on this image).

    python tests/timeline.py [batch] [out.json]

Prints, for the last profiled step: wall span, union-busy time over all streams, per-stream busy time, the idle gaps of
the union (histogram), time per kernel name, and how much of the busy time had >= 2 kernels in flight.  Numbers taken
under the profiler are for attribution only, never bench values.
"""
import collections
import json
import os
import sys

import torch
from torch.profiler import ProfilerActivity, profile

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from distributed_3d_diffusion_pytorch_b200 import synthetic  # noqa: E402
from distributed_3d_diffusion_pytorch_b200.trainer import Trainer  # noqa: E402

B = int(sys.argv[1]) if len(sys.argv) > 1 else 16
out = sys.argv[2] if len(sys.argv) > 2 else None
rank, world, local = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1")), int(os.environ.get("LOCAL_RANK", "0"))
dev = torch.device("cuda", local)
torch.cuda.set_device(local)
if world > 1:          # under torchrun: the data-parallel step (B pairs per rank); rank 0 writes the trace
    import torch.distributed as dist
    dist.init_process_group("nccl", device_id=dev)
kw = dict(H=64, W=64, ch=128)          # bench.py's CFG-64
tr = Trainer(kw, B, dev, world=world, seed=1234, use_graph=True)
res = [tuple(t.to(dev) for t in synthetic.make_views(B, kw["H"], seed=i)) for i in range(2)]
for i in range(4):
    tr.step_resident(res[i % 2])
torch.cuda.synchronize()
with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
    for i in range(2):
        tr.step_resident(res[i % 2])
        torch.cuda.synchronize()
if rank != 0:
    if world > 1:
        dist.barrier()
    sys.exit(0)
tmp = out or "/tmp/b3d_timeline.json"
prof.export_chrome_trace(tmp)
ev = json.load(open(tmp))["traceEvents"]
ks = [e for e in ev if e.get("cat") in ("kernel", "gpu_memcpy", "gpu_memset") and "dur" in e]
ks.sort(key=lambda e: e["ts"])
# split the two steps at the largest gap
gaps = [(ks[i + 1]["ts"] - (ks[i]["ts"] + ks[i]["dur"]), i) for i in range(len(ks) - 1)]
cut = max(gaps)[1] + 1
ks = ks[cut:]
t0 = ks[0]["ts"]
t1 = max(e["ts"] + e["dur"] for e in ks)
print(f"step: {len(ks)} device activities, span {(t1 - t0) / 1e3:.3f} ms")
# union busy and overlap
pts = []
for e in ks:
    pts.append((e["ts"], 1)); pts.append((e["ts"] + e["dur"], -1))
pts.sort()
busy = over = 0.0
depth = 0
last = t0
idle = []
for t, d in pts:
    if depth >= 1:
        busy += t - last
    if depth >= 2:
        over += t - last
    if depth == 0 and t > last:
        idle.append(t - last)
    depth += d
    last = t
print(f"union busy {busy / 1e3:.3f} ms, >=2 kernels in flight {over / 1e3:.3f} ms, idle {sum(idle) / 1e3:.3f} ms in {len(idle)} gaps")
h = collections.Counter()
for g in idle:
    h["<1us" if g < 1 else "<2us" if g < 2 else "<4us" if g < 4 else "<8us" if g < 8 else "<16us" if g < 16 else ">=16us"] += g
print("idle by gap length (ms):", {k: round(v / 1e3, 3) for k, v in sorted(h.items())})
st = collections.defaultdict(float)
for e in ks:
    st[e["args"].get("stream", e.get("tid"))] += e["dur"]
print("busy per stream (ms):", {k: round(v / 1e3, 3) for k, v in st.items()})
by = collections.defaultdict(lambda: [0.0, 0])
for e in ks:
    n = e["name"].split("(")[0].split("<")[0]
    by[n][0] += e["dur"]; by[n][1] += 1
for n, (d, c) in sorted(by.items(), key=lambda kv: -kv[1][0])[:30]:
    print(f"{d / 1e3:8.3f} ms  n={c:4d}  {d / c:7.1f} us  {n}")
# kernels on the main stream shorter than 8 us: how many, how much time, and the gap that follows each
main = max(st, key=st.get)
mk = [e for e in ks if e["args"].get("stream", e.get("tid")) == main]
short = [e for e in mk if e["dur"] < 8]
print(f"main stream {main}: {len(mk)} kernels, {len(short)} shorter than 8 us ({sum(e['dur'] for e in short) / 1e3:.3f} ms)")
g2 = [mk[i + 1]["ts"] - (mk[i]["ts"] + mk[i]["dur"]) for i in range(len(mk) - 1)]
if world > 1:
    dist.barrier()
print(f"main-stream gaps: total {sum(g for g in g2 if g > 0) / 1e3:.3f} ms, median {sorted(g2)[len(g2) // 2]:.2f} us")
