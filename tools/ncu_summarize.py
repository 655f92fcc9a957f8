"""Turns the raw ncu exports of tools/ncu_round.sh (gpurun_out/prof_*_raw.csv, launches_bench.csv)
into the tracked summaries under profiles/: one small per-launch CSV per workload, a markdown table,
traffic.json (dram bytes per launch, read by bench.py for roofline.traffic) and the kernel shares of
the bench launch list.   usage: python tools/ncu_summarize.py <round tag, e.g. r01b>"""
import csv
import glob
import json
import os
import re
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SRC = os.path.join(ROOT, "gpurun_out")
DST = os.path.join(ROOT, "profiles")
tag = sys.argv[1] if len(sys.argv) > 1 else "r01b"

COLS = {
    "time_us": "gpu__time_duration.sum",
    "dram_read": "dram__bytes_read.sum",
    "dram_write": "dram__bytes_write.sum",
    "grid": "launch__grid_size",
    "block": "launch__block_size",
    "regs": "launch__registers_per_thread",
    "dyn_smem": "launch__shared_mem_per_block_dynamic",
    "dram_pct": "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
    "l2_hit_pct": "lts__t_sector_hit_rate.pct",
    "l1_hit_pct": "l1tex__t_sector_hit_rate.pct",
    "warps_active_pct": "sm__warps_active.avg.pct_of_peak_sustained_active",
}
UNIT_SCALE = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "ns": 1e-3, "us": 1.0, "ms": 1e3,
              "usecond": 1.0, "msecond": 1e3, "nsecond": 1e-3, "second": 1e6}


def short(name):
    m = re.search(r"\b(k_[a-z0-9_]+)", name)  # (also inside cgb::<unnamed>::)
    return m.group(1) if m else name[:40]


def num(v):
    try:
        return float(v.replace(",", ""))
    except ValueError:
        return None


def read_raw(path):
    rows = list(csv.reader(open(path)))
    hdr, units = rows[0], rows[1]
    idx = {k: hdr.index(v) for k, v in COLS.items() if v in hdr}
    out = []
    for r in rows[2:]:
        d = {"kernel": short(r[hdr.index("Kernel Name")])}
        for k, i in idx.items():
            v = num(r[i])
            if v is not None and k in ("time_us", "dram_read", "dram_write", "dyn_smem"):
                v *= UNIT_SCALE.get(units[i].split("/")[0], 1.0)
            d[k] = v
        out.append(d)
    return out


ALG = {  # algorithmic bytes per launch
    "dense": {"n": 32768, "nnz": 32768 ** 2, "dense": True},
    "lap2d": {"n": 4096 ** 2, "nnz": 5 * 4096 ** 2 - 4 * 4096},
    "lap3d512": {"n": 512 ** 3, "nnz": 7 * 512 ** 3 - 6 * 512 ** 2},
    "banded": {"n": 52269, "nnz": 18003049},   # course-sized long-row system (bench.py: banded52269)
    "bandeddia": {"n": 52269, "nnz": 18003049, "dia": 345 * 52288},  # the same as 345 diagonals of ld = 52288 values
}
DENSE_ITERS_PER_LAUNCH = 5   # tools/ncu_round.sh profiles one persistent launch: loop entry + 5 iterations


def alg_bytes(w, kernel):
    a = ALG[w]
    n, nnz = a["n"], a["nnz"]
    if kernel.startswith("k_cg_dense_solve"):  # loop entry (32 n) + 5 whole iterations (A + 88 n each)
        return DENSE_ITERS_PER_LAUNCH * (8.0 * nnz + 88.0 * n) + 32.0 * n
    if kernel.startswith("k_cg_dense"):  # the whole iteration: contract bytes A + 88 n
        return 8.0 * nnz + 88.0 * n
    if kernel.startswith("k_gemv"):
        return 8.0 * nnz + 16.0 * n
    if kernel.startswith("k_spmv_dia"):  # every stored diagonal once (padding included) + p in + s out
        return 8.0 * a["dia"] + 16.0 * n
    if kernel.startswith("k_spmv_pattern"):  # one byte per row instead of the CSR arrays
        return 1.0 * n + 16.0 * n
    if kernel.startswith("k_spmv"):
        return 12.0 * nnz + 4.0 * (n + 1) + 16.0 * n
    return {"k_update_xr": 48.0 * n, "k_update_p": 24.0 * n, "k_update_sr": 72.0 * n}.get(kernel)


traffic = {"_by_workload": {}, "_note": "dram__bytes_read.sum + dram__bytes_write.sum per launch, ncu --set full --clock-control none, round " + tag}
md = ["# ncu `--set full --clock-control none` captures, round " + tag + " (one B200; cold-cache, serialised replays)", "",
      "Command per workload: `tools/ncu_round.sh` (each profiled command first exits 0 without ncu). "
      "`alg` = algorithmic bytes per launch (DESIGN.md section 4). Times under ncu are not bench values.", "",
      "| workload | variant | kernel | grid x block | regs | dyn smem KB | time us | DRAM read GB | DRAM write GB | traffic GB | alg GB | traffic/alg | GB/s | DRAM % of peak | L2 hit % | L1 hit % | warps active % |",
      "|---|---|---|---|---|---|---|---|---|---|---|---|---|---|---|---|---|"]
ONE = {"one_pat512_raw.csv": ("lap3d512", "pattern")}  # single-kernel captures (tools/ncu_one.sh)
paths = sorted(glob.glob(os.path.join(SRC, "prof_*_raw.csv"))) + [os.path.join(SRC, k) for k in ONE if os.path.exists(os.path.join(SRC, k))]
for path in paths:
    if os.path.basename(path) in ONE:
        w, v = ONE[os.path.basename(path)]
    else:
        m = re.match(r"prof_(\w+?)_(classic|sr)_raw\.csv", os.path.basename(path))
        w, v = m.group(1), m.group(2)
    rows = read_raw(path)
    keep = os.path.join(DST, f"{tag}_ncu_{w}_{v}.csv")
    with open(keep, "w", newline="") as f:
        wr = csv.writer(f)
        wr.writerow(["kernel"] + list(COLS))
        for r in rows:
            wr.writerow([r["kernel"]] + [r.get(k) for k in COLS])
    seen = {}
    for r in rows:  # average the captured launches of each kernel
        seen.setdefault(r["kernel"], []).append(r)
    for kname, rs in seen.items():
        avg = lambda k: sum(x[k] for x in rs if x.get(k) is not None) / max(1, sum(1 for x in rs if x.get(k) is not None))
        rd, wrt, t = avg("dram_read"), avg("dram_write"), avg("time_us")
        alg = alg_bytes(w, kname)
        md.append(f"| {w} | {v} | {kname} | {int(avg('grid'))} x {int(avg('block'))} | {int(avg('regs'))} | {avg('dyn_smem') / 1e3:.1f} | {t:.1f} | "
                  f"{rd / 1e9:.4f} | {wrt / 1e9:.4f} | {(rd + wrt) / 1e9:.4f} | {alg / 1e9 if alg else 0:.4f} | {(rd + wrt) / alg if alg else 0:.3f} | "
                  f"{(rd + wrt) / t / 1e3:.0f} | {avg('dram_pct'):.1f} | {avg('l2_hit_pct'):.1f} | {avg('l1_hit_pct'):.1f} | {avg('warps_active_pct'):.1f} |")
        if v == "classic" or kname == "k_update_sr" or kname.startswith("k_spmv_pattern"):
            key = {"banded": "banded52269", "bandeddia": "banded52269_dia"}.get(w, w)
            traffic["_by_workload"].setdefault(key, {})[kname] = round(rd + wrt)
            if kname == "k_cg_dense_solve":
                traffic["_by_workload"][key]["_iterations_per_launch"] = DENSE_ITERS_PER_LAUNCH
with open(os.path.join(DST, f"{tag}_ncu_summary.md"), "w") as f:
    f.write("\n".join(md) + "\n")
json.dump(traffic, open(os.path.join(DST, "traffic.json"), "w"), indent=1)

# launch list of the bench command: per-kernel count, total time and share
lp = os.path.join(SRC, "launches_bench.csv")
if os.path.exists(lp):
    lines = [l for l in open(lp) if not l.startswith("==")]
    rows = list(csv.reader(lines))
    hdr = rows[0]
    ki, mi, vi, ui = hdr.index("Kernel Name"), hdr.index("Metric Name"), hdr.index("Metric Value"), hdr.index("Metric Unit")
    agg = {}
    for r in rows[1:]:
        if len(r) <= vi or r[mi] != "gpu__time_duration.sum":
            continue
        t = num(r[vi]) * UNIT_SCALE.get(r[ui], 1.0)
        k = short(r[ki])
        a = agg.setdefault(k, [0, 0.0])
        a[0] += 1
        a[1] += t
    tot = sum(a[1] for a in agg.values())
    with open(os.path.join(DST, f"{tag}_launches_bench_summary.md"), "w") as f:
        f.write(f"# Launch list of `python bench.py --steps 20 --warmup 3 --no-cpu --no-tts` under `ncu --metrics gpu__time_duration.sum --clock-control none` ({tag})\n\n"
                "Per-launch times are cold-cache and serialised; the kernel's SHARE is what must agree with the bench.\n"
                "The run covers the headline workload (dense 32768), the end-to-end leg and the other two BASELINE workloads, set-up kernels included.\n\n"
                "| kernel | launches | total ms | share |\n|---|---|---|---|\n")
        for k, a in sorted(agg.items(), key=lambda kv: -kv[1][1]):
            f.write(f"| {k} | {a[0]} | {a[1] / 1e3:.3f} | {a[1] / tot:.4f} |\n")
    # keep the raw list too (compact: id, kernel, us)
    with open(os.path.join(DST, f"{tag}_launches_bench.csv"), "w", newline="") as f:
        wr = csv.writer(f)
        wr.writerow(["id", "kernel", "gpu__time_duration_us"])
        for r in rows[1:]:
            if len(r) > vi and r[mi] == "gpu__time_duration.sum":
                wr.writerow([r[0], short(r[ki]), round(num(r[vi]) * UNIT_SCALE.get(r[ui], 1.0), 3)])
print(open(os.path.join(DST, f"{tag}_ncu_summary.md")).read())
print(open(os.path.join(DST, f"{tag}_launches_bench_summary.md")).read())
