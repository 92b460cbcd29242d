import sys, os
root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, root); sys.path.insert(0, os.path.join(root, "profiles"))
from dynolog_b200 import hbtgpu
import perf_probe
perf_probe.probe("cfg3 force_general", 3, 64, 8, 70000, max_bytes_per_cpu=16 << 20, flags=hbtgpu.F_FORCE_GENERAL)
