"""Prints BASELINE.md's results table from bench JSON lines. Usage: python tools/results_table.py profiles/r1l_bench_*.json"""
import json, sys

rows = []
for f in sys.argv[1:]:
    try:
        d = json.loads(open(f).read().strip().splitlines()[-1])
    except Exception:
        continue
    cfg = d["config"]["workload"].split(":")[0]
    r = d.get("roofline", {})
    a = r.get("all", {})
    cpu = d.get("cpu_baseline") or {}
    e2e = d["e2e"]["value"]
    rows.append((cfg, d["n_gpus"], d["ms_per_step"] / 1e3, d["value"], d.get("mrays_per_s", 0.0), d.get("rays_per_path", 0.0), e2e,
                 r.get("share_of_step"), a.get("l2", {}).get("frac"), a.get("fp32", {}).get("frac"), a.get("hbm", {}).get("frac"),
                 cpu.get("value"), cpu.get("cores")))
rows.sort(key=lambda t: (t[0], t[1]))
print("| config | GPUs | s/frame | Mpaths/s | Mrays/s | rays/path | e2e Mpaths/s | extend share | L2 frac | FP32 frac | HBM frac (step) | CPU port Mpaths/s (threads) | e2e / CPU |")
print("|---|---|---|---|---|---|---|---|---|---|---|---|---|")
f2 = lambda v, n=2: "—" if v is None else f"{v:.{n}f}"
for c, n, s, v, mr, rp, e, sh, l2, fp, hb, cv, cc in rows:
    cpu = f"{cv:.2f} ({cc})" if cv and n == 1 else "—"
    ratio = f"{e / cv:.0f}x" if cv and n == 1 else "—"
    print(f"| {c} | {n} | {s:.3f} | {v:.0f} | {mr:.0f} | {rp:.2f} | {e:.0f} | {f2(sh)} | {f2(l2)} | {f2(fp, 3)} | {f2(hb)} | {cpu} | {ratio} |")
