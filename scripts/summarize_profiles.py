"""Turn gpurun_out/prof_ops.ncu-rep (+ the built .so) into the tracked summaries under profiles/.
Runs on the CPU box: `ncu -i ... --page raw --csv` and `cuobjdump -sass`."""
import csv, io, json, os, re, subprocess, sys, collections
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
rep = os.path.join(ROOT, "gpurun_out", "prof_ops.ncu-rep")
so = os.path.join(ROOT, "distributedvolunteercomputing_b200", "ops", "libdvc_b200.so")
out_dir = os.path.join(ROOT, "profiles")
os.makedirs(out_dir, exist_ok=True)
peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))) if os.path.exists(os.path.join(ROOT, "MEASURED_PEAKS.json")) else {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0}

LABELS = ["K1a resize 64x300x300->400x400", "K1b preproc+conv0, 16 frames", "K3 dw conv1/dw 16x150x150x32", "K2 gemm conv1 M=360000 N=64 K=32",
          "K3 dw conv7/dw 64x19x19x512", "K2 gemm conv7 M=23104 N=512 K=512", "K2 gemm conv13 M=6400 N=1024 K=1024",
          "K5 head gemm conv11_mbox M=23104 N=80 K=512", "K7 softmax 64x1917x21", "K8 per-class NMS 64 frames x 20 classes (dense random scores)", "K9 merge+deliver 64 frames", "K10 annotate 64 frames"]
FLOPS = [None, 16 * 19.44e6 * 2, None, 2.0 * 360000 * 64 * 32, None, 2.0 * 23104 * 512 * 512, 2.0 * 6400 * 1024 * 1024, 2.0 * 23104 * 80 * 512, None, None, None, None]
BYTES = [64 * (270000 + 480000), 16 * (480000 + 1440000), 16 * 22500 * 32 * 2 * 2, 360000 * (32 + 64) * 2, 64 * 361 * 512 * 2 * 2,
         23104 * 1024 * 2 + 512 * 512 * 2, 6400 * 2048 * 2 + 1024 * 1024 * 2, 23104 * 512 * 2 + 23104 * 75 * 4, 64 * 1917 * 21 * 4 * 2, None, None, None]

if os.path.exists(rep):
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    hdr, units, data = rows[0], rows[1], rows[2:]
    idx = {h: i for i, h in enumerate(hdr)}
    cols = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
            "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "sm__warps_active.avg.pct_of_peak_sustained_active",
            "launch__registers_per_thread", "launch__grid_size", "sm__throughput.avg.pct_of_peak_sustained_elapsed"]
    lines = ["# ncu --set full, one launch per hot kernel (scripts/profile_ops.py; cold L2, serialised - use for shares and stall analysis)",
             "", f"Roofline denominators (measured, MEASURED_PEAKS.json): HBM copy {peaks['hbm_gbs']} GB/s, cuBLAS bf16 {peaks['bf16_tflops']} TFLOP/s.", "",
             "| kernel | shape | time us | dram rd MB | dram % | tensor pipe % | warps active % | regs | grid | algorithmic GB/s (of measured) | TFLOP/s (of measured) |", "|---|---|---|---|---|---|---|---|---|---|---|"]
    def val(r, c):
        v = r[idx[c]].replace(",", "")
        return float(v) if v not in ("", "n/a") else float("nan")
    for k, r in enumerate(data):
        name = r[idx["Kernel Name"]].split("(")[0].replace("void ", "")
        t = val(r, "gpu__time_duration.sum")
        tu = units[idx["gpu__time_duration.sum"]]
        t_us = t / 1e3 if tu in ("ns", "nsecond") else (t * 1e3 if tu in ("ms", "msecond") else t)
        rd = val(r, "dram__bytes_read.sum"); ru = units[idx["dram__bytes_read.sum"]]
        rd_mb = rd / 1e6 if ru == "byte" else (rd / 1e3 if ru == "Kbyte" else (rd * 1e3 if ru == "Gbyte" else rd))
        lab = LABELS[k] if k < len(LABELS) else ""
        gbs = (BYTES[k] / t_us / 1e3) if k < len(BYTES) and BYTES[k] else None
        tf = (FLOPS[k] / t_us / 1e6) if k < len(FLOPS) and FLOPS[k] else None
        lines.append(f"| {name} | {lab} | {t_us:.1f} | {rd_mb:.1f} | {val(r, cols[3]):.1f} | {val(r, cols[4]):.1f} | {val(r, cols[5]):.1f} | "
                     f"{int(val(r, cols[6]))} | {int(val(r, cols[7]))} | " + (f"{gbs:.0f} ({gbs / peaks['hbm_gbs']:.0%})" if gbs else "-") + " | " +
                     (f"{tf:.0f} ({tf / peaks['bf16_tflops']:.0%})" if tf else "-") + " |")
    open(os.path.join(out_dir, "ncu_ops_summary.md"), "w").write("\n".join(lines) + "\n")
    print("\n".join(lines))

if os.path.exists(so):
    sass = subprocess.run(["cuobjdump", "-sass", so], capture_output=True, text=True).stdout
    cur, table = None, collections.OrderedDict()
    for ln in sass.splitlines():
        m = re.search(r"Function : (\S+)", ln)
        if m:
            cur = m.group(1); table[cur] = collections.Counter(); continue
        m = re.match(r"\s+/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z][A-Z0-9_]*)", ln)
        if m and cur:
            table[cur][m.group(1)] += 1
    keys = ["UTCHMMA", "UTCBAR", "LDTM", "UTMALDG", "UTMASTG", "UBLKCP", "SYNCS", "ACQBULK", "PREEXIT", "FHFMA", "HMMA", "FFMA", "LDS", "STS", "LDG", "STG", "ST", "LD", "BAR"]
    lines = ["# SASS mnemonic counts per kernel (cuobjdump -sass of ops/libdvc_b200.so, sm_100a)", "",
             "UTCHMMA = tcgen05.mma, LDTM = tcgen05.ld, UTMALDG/UTMASTG = TMA load/store, SYNCS = mbarrier ops, ACQBULK / PREEXIT = griddepcontrol.wait / launch_dependents (programmatic dependent launch), FHFMA = fp16xfp16+fp32 FMA.", "",
             "| kernel | instrs | " + " | ".join(keys) + " |", "|---|---|" + "---|" * len(keys)]
    for fn, cnt in table.items():
        short = subprocess.run(["c++filt", fn], capture_output=True, text=True).stdout.strip().split("(")[0]
        lines.append(f"| {short} | {sum(cnt.values())} | " + " | ".join(str(cnt.get(k, 0)) for k in keys) + " |")
    open(os.path.join(out_dir, "sass_mnemonics.md"), "w").write("\n".join(lines) + "\n")
    # full listing of the two flagship kernels
    for pat, fname in (("gemm_f16_kernel", "sass_gemm_f16_kernel.txt"), ("dw3x3_tma_kernelILi1", "sass_dw3x3_tma_kernel.txt")):
        keep, on = [], False
        for ln in sass.splitlines():
            if "Function :" in ln:
                on = pat in ln
            if on:
                keep.append(ln)
        open(os.path.join(out_dir, fname), "w").write("\n".join(keep) + "\n")
    print("\n".join(lines[:4] + lines[4:]))
