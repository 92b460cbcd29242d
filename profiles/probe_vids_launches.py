import sys, os
sys.path.insert(0, os.getcwd())
import numpy as np
from dynolog_b200 import hbtgpu_cs, synth
ncpu, per = 192, 10 << 20
streams = synth.cs_capture(ncpu, per)
total = sum(len(s) for s in streams)
g = hbtgpu_cs.CsContext(ncpu, total + (1 << 20), tid_level=1)
for c, s in enumerate(streams):
    g.submit(c, s)
for rep in range(2):
    g.run(); g.sync()
    dev = hbtgpu_cs.DeviceThreads()
    dev.assign(g)
    print("ms", dev.last_ms())
    dev.close()
