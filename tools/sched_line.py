"""Prints wf_extend's warp-scheduling counters (instrumented build) of bench.py JSON lines read from stdin."""
import json
import sys

for l in sys.stdin:
    if l.startswith("{"):
        d = json.loads(l)
        r = d.get("roofline", {})
        s = r.get("extend_warp_scheduling") or {}
        pr = r.get("per_ray", {})
        out = f"Mpaths/s {d['value']:.1f} Mrays/s {d.get('mrays_per_s', 0):.1f} rays/path {d.get('rays_per_path', 0):.2f} | per ray: node {pr.get('node', 0):.2f} sph {pr.get('sphere', 0):.2f} quad {pr.get('quad', 0):.2f} inst {pr.get('instance', 0):.2f} |"
        for k in ("node", "prim", "event", "fetch"):
            if k in s:
                out += f" {k}: {s[k]['lanes_per_step']:.1f} lanes x {s[k]['steps'] / 1e6:.1f}M steps, {s[k]['cycles_per_step']:.0f} cyc;"
        if s:
            tot = sum(s[k]["cycles"] for k in ("node", "prim", "event", "fetch")) + s["loop_cycles"]
            out += " share of cycles: " + " ".join(f"{k} {s[k]['cycles'] / tot:.2f}" for k in ("node", "prim", "event", "fetch")) + f" loop {s['loop_cycles'] / tot:.2f}"
        print(out)
