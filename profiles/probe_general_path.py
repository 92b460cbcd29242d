"""One configuration of profiles/perf_probe.py alone, for ncu: the forced general path (classify every
word, pointer jumping, listed samples) on a cfg3-like capture.
  ncu --set full -k regex:k_decode -s 2 -c 1 ... python profiles/probe_general_path.py [general|mixed|cfg4]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import perf_probe
from dynolog_b200 import hbtgpu

what = sys.argv[1] if len(sys.argv) > 1 else "general"
if what == "cfg4":
    perf_probe.probe("cfg4 IP|TID 88B 64cpu x 32MiB", 4, 64, 4, 50000, fmt=1, max_bytes_per_cpu=32 << 20)
elif what == "general":
    perf_probe.probe("cfg3 force_general", 3, 64, 8, 70000, max_bytes_per_cpu=16 << 20, flags=hbtgpu.F_FORCE_GENERAL)
else:
    perf_probe.probe("cfg3-like, THROTTLE pair every 20-40", 3, 64, 8, 70000, max_bytes_per_cpu=16 << 20,
                     throttle=(20, 40), lost=(200, 2000))
