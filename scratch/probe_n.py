import os, sys
root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, root); sys.path.insert(0, os.path.join(root, "profiles"))
import perf_probe
for n in (1, 2, 6, 12, 16, 32, 64):
    perf_probe.probe(f"n={n} {40+8*n}B records 64cpu x 16MiB", 2, 64, n, 50000, max_bytes_per_cpu=16 << 20, n_events=n)
