import os, sys
root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))  # repo root
sys.path.insert(0, root); sys.path.insert(0, os.path.join(root, "profiles"))
import perf_probe
for lo, hi in ((100, 200), (20, 40), (10, 20), (5, 10), (2, 4)):
    perf_probe.probe(f"cfg3-like, THROTTLE pair every {lo}-{hi} samples", 3, 64, 8, 30000, max_bytes_per_cpu=16 << 20,
                     throttle=(lo, hi), lost=(0, 0))
