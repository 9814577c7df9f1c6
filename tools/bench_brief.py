"""stdin: one bench.py JSON line -> the handful of numbers looked at while tuning."""
import json, sys
d = json.loads(sys.stdin.read())
print(sys.argv[1] if len(sys.argv) > 1 else "", d["value"], d["ms_per_step"], d["roofline"].get("kernel_ms"), d["e2e"]["value"])
