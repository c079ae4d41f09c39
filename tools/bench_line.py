"""Prints the key numbers of bench.py JSON lines read from stdin (tuning helper)."""
import json
import sys

for l in sys.stdin:
    if l.startswith("{"):
        d = json.loads(l)
        r = d.get("roofline", {})
        print(f"Mpaths/s {d['value']:.1f}  Mrays/s {d.get('mrays_per_s', 0):.1f}  ms/step {d['ms_per_step']:.2f}  "
              f"e2e {d['e2e']['value']:.1f}  mean {d.get('image_mean')}  launches {d.get('gpu_launches')}  "
              f"extend_ms {r.get('kernel_ms_per_launch', 0):.4f} shade_ms {r.get('shade_ms_per_launch', 0):.4f} iters {r.get('launches_per_step')}")
    else:
        print(l.rstrip()[:300])
