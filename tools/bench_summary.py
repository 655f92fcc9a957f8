"""Prints one line per bench JSON file: headline value, ms/step, dominant kernel, e2e, workloads."""
import json
import sys

for f in sys.argv[1:]:
    try:
        lines = [l for l in open(f).read().splitlines() if l.startswith("{")]
        d = json.loads(lines[-1])
    except Exception as e:  # noqa: BLE001
        print(f, "unreadable:", e)
        continue
    r = d.get("roofline", {})
    print(f"{f}: N={d.get('n_gpus')} {d.get('value')} it/s {d.get('ms_per_step')} ms/step | {r.get('kernel')} "
          f"{r.get('avg_launch_ms')} ms frac {r.get('frac')} | iter frac {d.get('iteration', {}).get('frac_of_peak')} "
          f"| e2e {d.get('e2e', {}).get('value')} | launches {d.get('gpu_launches')} | other {d.get('other_variant', {}).get('value')}")
    for k, w in d.get("workloads", {}).items():
        if "value" not in w:
            print("    ", k, w)
            continue
        print(f"     {k}: {w['value']} it/s {w['ms_per_step']} ms frac {w.get('frac_of_peak')} K1 {w['roofline']['avg_launch_ms']} ms "
              f"({w['roofline']['frac']}) other {w.get('other_variant', {}).get('value')}")
        pf = w.get("pattern_format")
        if pf and "value" in pf:
            print(f"        pattern format: {pf['value']} it/s {pf['ms_per_step']} ms x{pf['speedup_vs_csr']} vs CSR | "
                  f"{pf['achieved_gbs_per_gpu']} GB/s of its own {pf['bytes_per_iteration'] / 1e9:.2f} GB/iter "
                  f"({pf['frac_of_peak']}) | K1 {pf['matvec']['avg_launch_ms']} ms {pf['matvec']['achieved']} GB/s")
        elif pf:
            print("        pattern format:", pf)
