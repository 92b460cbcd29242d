import sys, os
root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, root); sys.path.insert(0, os.path.join(root, "profiles"))
from dynolog_b200 import hbtgpu
hbtgpu.LIB_PATH = os.path.abspath(sys.argv[1])
import perf_probe
perf_probe.probe("cfg2 " + os.path.basename(sys.argv[1]), 2, 192, 8, 10100, n_samples=100_000)
perf_probe.probe("n=4 " + os.path.basename(sys.argv[1]), 1, 64, 4, 50000, max_bytes_per_cpu=32 << 20)
