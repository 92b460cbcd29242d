import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from dynolog_b200 import hbtgpu
if len(sys.argv) > 1:
    hbtgpu.LIB_PATH = os.path.abspath(sys.argv[1])
sys.argv = sys.argv[:1]
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "profiles"))
import perf_probe
perf_probe.probe("cfg4 IP|TID 88B 64cpu x 32MiB", 4, 64, 4, 50000, fmt=1, max_bytes_per_cpu=32 << 20)
