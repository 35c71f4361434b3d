#!/usr/bin/env python
"""Reads bench.py output on stdin, prints the JSON line's key numbers on one line (for gpurun logs)."""
import json
import sys

for line in sys.stdin:
    if not line.startswith("{"):
        continue
    d = json.loads(line)
    if "unavailable" in d:
        print(d)
        continue
    e = d.get("e2e") or {}
    c = d.get("config") or {}
    print(f"impl={d.get('impl')} n={d.get('n_gpus')} value={d['value']:.0f} us/step={1e3 * d['ms_per_step']:.2f} "
          f"e2e={e.get('value', 0):.0f} e2e_us/step={1e3 * e.get('ms_per_step', 0):.2f} spe={c.get('steps_per_execution')} "
          f"reps={[round(x, 3) for x in c.get('ms_per_repeat', c.get('ms_per_rep', []))]} launches={d.get('gpu_launches')} "
          f"check={d.get('correctness')} clocks={d.get('clocks')}")
