"""Summarise a --profile-json dump: time per kernel class and per GEMM shape family."""
import collections
import json
import sys

rows = json.load(open(sys.argv[1]))
tot = sum(r["ms"] for r in rows)
by = collections.defaultdict(lambda: [0.0, 0, 0.0])
for r in rows:
    key = r["name"]
    if r["name"] == "b3d_gemm":
        i = r["info"]
        key = "gemm " + " ".join(x for x in i.split() if x.split("=")[0] in ("maj00", "maj01", "maj11", "taps", "s", "split", "fp32", "batched") or x.startswith("maj"))
    by[key][0] += r["ms"]; by[key][1] += 1; by[key][2] += r["flop"]
print(f"total {tot:.2f} ms over {len(rows)} launches")
for k, (ms, n, fl) in sorted(by.items(), key=lambda kv: -kv[1][0])[:40]:
    tf = fl / ms / 1e9 if fl else 0
    print(f"{ms:8.3f} ms {100*ms/tot:5.1f}%  n={n:4d}  {tf:7.1f} TF/s  {k}")
if len(sys.argv) > 2:
    print("---- slowest gemm launches")
    for r in sorted((r for r in rows if r["name"] == "b3d_gemm"), key=lambda r: -r["ms"])[:int(sys.argv[2])]:
        print(f"{r['ms']:7.3f} ms {r['flop']/r['ms']/1e9:7.1f} TF/s  {r['info']}")
