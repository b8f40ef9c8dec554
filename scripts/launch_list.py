"""gpurun_out/launches_warm.csv (ncu --metrics gpu__time_duration.sum over scripts/profile_step.py) ->
profiles/r2/launch_list_step.md: every launch of the LAST eager step, in order, plus per-kernel totals."""
import collections, csv, io, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
src = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "gpurun_out", "launches_warm.csv")
txt = [ln for ln in open(src) if not ln.startswith("==")]
rows = list(csv.DictReader(io.StringIO("".join(txt))))
mine = [r for r in rows if "dvc::" in r["Kernel Name"] or any(k in r["Kernel Name"] for k in ("_kernel",))]
mine = [r for r in mine if "at::" not in r["Kernel Name"]]
starts = [i for i, r in enumerate(mine) if "resize_u8_kernel" in r["Kernel Name"]]
first = starts[-1]
while first > 0 and any(k in mine[first - 1]["Kernel Name"] for k in ("mbx_post", "mbx_wait_desc")):      # mailbox path: the post / wait kernels open the step
    first -= 1
step = mine[first:]
def us(r):
    v = float(r["Metric Value"].replace(",", "")); u = r["Metric Unit"]
    return v / 1e3 if u.startswith("n") else (v * 1e3 if u.startswith("m") else v)
lines = ["# One flagship batch (64 frames, eager, dispatched through the mailbox data plane on one GPU) under `ncu --metrics gpu__time_duration.sum --cache-control none`", "",
         "Warm-L2 durations of every launch of one step, in launch order (serialised by ncu: shares, not absolutes).", "",
         "| # | kernel | grid | us |", "|---|---|---|---|"]
tot = collections.OrderedDict()
for i, r in enumerate(step):
    name = r["Kernel Name"].split("(")[0].replace("void ", "")
    lines.append(f"| {i} | {name} | {r['Grid Size']} | {us(r):.1f} |")
    c = tot.setdefault(name, [0, 0.0]); c[0] += 1; c[1] += us(r)
total = sum(v[1] for v in tot.values())
lines += ["", f"Total {total:.1f} us over {len(step)} launches.", "", "| kernel | launches | us | share |", "|---|---|---|---|"]
for k, (n, t) in sorted(tot.items(), key=lambda kv: -kv[1][1]):
    lines.append(f"| {k} | {n} | {t:.1f} | {t / total:.1%} |")
open(os.path.join(ROOT, "profiles", "r2", "launch_list_step.md"), "w").write("\n".join(lines) + "\n")
print("\n".join(lines[-12:]))
