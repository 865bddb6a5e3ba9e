"""Turns an `ncu --set full` report (brought back from the GPU box under gpurun_out/) into the small text / JSON summaries
committed under profiles/:   python tests/ncu_summary.py raw <report.ncu-rep> <out.json>
                             python tests/ncu_summary.py stalls <report.ncu-rep> <kernel substring> <out.txt>
                             python tests/ncu_summary.py sass <libb3d.so> <out.txt>
(runs in the build container: `ncu -i` and `cuobjdump` need no GPU)."""
import csv
import io
import json
import re
import subprocess
import sys
from collections import Counter

RAW = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
       "sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed",
       "sm__warps_active.avg.pct_of_peak_sustained_active", "lts__t_sector_hit_rate.pct", "lts__throughput.avg.pct_of_peak_sustained_elapsed",
       "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum.pct_of_peak_sustained_elapsed", "smsp__inst_executed.sum", "launch__grid_size",
       "launch__cluster_size", "launch__registers_per_thread", "sm__cycles_elapsed.max"]


def raw(report, out):
    txt = subprocess.run(["ncu", "-i", report, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(txt)))
    hdr, units = rows[0], rows[1]
    ix = {n: i for i, n in enumerate(hdr)}
    res = []
    for r in rows[2:]:
        d = {"kernel": r[ix["Kernel Name"]][:100]}
        for k in RAW:
            if k in ix:
                d[f"{k} [{units[ix[k]]}]"] = r[ix[k]]
        res.append(d)
    json.dump({"report": report, "kernels": res}, open(out, "w"), indent=1)
    print(out, len(res), "kernels")


def load_source(report, pat):
    txt = subprocess.run(["ncu", "-i", report, "--page", "source", "--csv"], capture_output=True, text=True).stdout
    kernels, cur = [], None
    for r in csv.reader(io.StringIO(txt)):
        if r and r[0] == "Kernel Name":
            cur = {"name": r[1], "rows": [], "hdr": None}
            kernels.append(cur)
        elif cur is not None and cur["hdr"] is None:
            cur["hdr"] = r
        elif cur is not None:
            cur["rows"].append(r)
    return [k for k in kernels if pat in k["name"]]


def _i(r, ix, k):
    try:
        return int(float(r[ix[k]] or 0))
    except Exception:  # noqa: BLE001
        return 0


def stalls(report, pat, out):
    k = load_source(report, pat)[0]
    ix = {n: i for i, n in enumerate(k["hdr"])}
    rows = [r for r in k["rows"] if len(r) > ix["Instructions Executed"]]
    tot = sum(_i(r, ix, "# Samples") for r in rows)
    mark = re.compile(r"UTMALDG|UTCHMMA|UTCBAR|LDTM|UTMASTG|UTMAREDG|SYNCS\.PHASECHK|SYNCS\.ARRIVE|BAR\.SYNC|ERRBAR|MEMBAR|EXIT")
    with open(out, "w") as fh:
        fh.write(f"{k['name'][:120]}\nwarp-stall samples {tot} over {len(rows)} SASS instructions; listed: >= 0.6 % of the samples, or a "
                 "synchronisation / tensor / TMA instruction that executed\n  line  executed  samples  instruction\n")
        for i, r in enumerate(rows):
            s, ex, src = _i(r, ix, "# Samples"), _i(r, ix, "Instructions Executed"), r[ix["Source"]]
            if s >= tot * 0.006 or (mark.search(src) and ex > 0):
                fh.write(f"{i:6d} {ex:9d} {s:8d}  {src[:110]}\n")
        c = Counter()
        for r in rows:
            t = r[ix["Source"]].split()
            if t:
                op = t[1] if t[0].startswith("@") and len(t) > 1 else t[0]
                c[op.split(".")[0]] += _i(r, ix, "Instructions Executed")
        fh.write("executed warp instructions by opcode: " + ", ".join(f"{k_} {v}" for k_, v in c.most_common(16)) + "\n")
    print(out, tot, "samples")


def sass(lib, out):
    txt = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True).stdout
    cur, per = None, {}
    ops = ["UTCHMMA", "UTCHMMA.2CTA", "LDTM", "UTMALDG", "UTMASTG", "UTMAREDG", "UTCBAR", "SYNCS", "MUFU.EX2", "ERRBAR"]
    for line in txt.splitlines():
        m = re.search(r"Function : (\S+)", line)
        if m:
            cur = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip()[:110]
            per[cur] = Counter()
            continue
        if cur:
            for o in ops:
                if re.search(r"\b" + re.escape(o) + r"\b", line) or (o == "UTCHMMA.2CTA" and "UTCHMMA.2CTA" in line):
                    per[cur][o] += 1
    with open(out, "w") as fh:
        fh.write("cuobjdump -sass libb3d.so: tcgen05 / TMEM / TMA instruction counts per kernel (sm_100a)\n")
        for k, c in per.items():
            if any(c[o] for o in ("UTCHMMA", "LDTM", "UTMALDG", "UTMASTG", "UTMAREDG")):
                fh.write(f"{k}\n    " + "  ".join(f"{o}={c[o]}" for o in ops if c[o]) + "\n")
    print(out)


def launches(csv_path, out, title=""):
    """ncu --metrics gpu__time_duration.sum --csv launch list -> per-kernel totals and shares"""
    rows = [r for r in csv.reader(open(csv_path)) if len(r) > 10]
    hdr = rows[0]
    ix = {n: i for i, n in enumerate(hdr)}
    agg = {}
    for r in rows[1:]:
        if r[ix["Metric Name"]] != "gpu__time_duration.sum":
            continue
        name = re.sub(r"\(.*", "", r[ix["Kernel Name"]]).replace("void ", "").replace("<unnamed>::", "")
        v = float(r[ix["Metric Value"]].replace(",", ""))
        unit = r[ix["Metric Unit"]]
        ms = v / 1e6 if unit in ("ns", "nsecond") else v / 1e3 if unit in ("us", "usecond") else v
        a = agg.setdefault(name, [0, 0.0])
        a[0] += 1
        a[1] += ms
    tot = sum(v[1] for v in agg.values())
    with open(out, "w") as fh:
        fh.write(f"{title}\n{sum(v[0] for v in agg.values())} launches, {tot:.1f} ms total (serialised, cold caches: shares, not absolutes, compare with bench.py)\n")
        for k, v in sorted(agg.items(), key=lambda kv: -kv[1][1]):
            fh.write(f"{v[1]:10.3f} ms  {100 * v[1] / tot:5.1f}%  n={v[0]:4d}  {k}\n")
    print(out, tot)


if __name__ == "__main__":
    {"raw": raw, "stalls": stalls, "sass": sass, "launches": launches}[sys.argv[1]](*sys.argv[2:])

