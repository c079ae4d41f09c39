"""Merges the JSON lines written by tests/test_gpu_parity.py::test_first_hit_parity under RTW_L1_RECORD=<file> into
tests/golden/l1_measured.json (one object: case key -> measured figures).   python tools/l1_collect.py in.jsonl out.json"""
import json
import sys

rows = {}
for ln in open(sys.argv[1]):
    if ln.strip():
        rows.update(json.loads(ln))
json.dump(dict(sorted(rows.items())), open(sys.argv[2], "w"), indent=0, sort_keys=True)
print(f"{len(rows)} cases -> {sys.argv[2]}")
