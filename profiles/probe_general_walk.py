"""The Slicer's general form alone (for ncu -k regex:k_gen_walk): python profiles/probe_general_walk.py"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import numpy as np
from dynolog_b200 import hbtgpu_slices as gs
from oracle import pyoracle as po
from test_slicer_oracle import random_phase_events

r = np.random.default_rng(5)
gev = random_phase_events(r, 16, 100_000, 300, p_phase=0.35, p_we=0.0005, p_create=0.01, p_bad=0.0)
gev = gev[np.lexsort((gev["comp_unit_id"], gev["tstamp"]))]
k = len(gev)
while po.slicer_run("port_general", gev[:k], 16)["rc"] == -95:
    k = k * 9 // 10
gev = gev[:k]
sl = gs.Slicer(0)
for it in range(2):
    got = sl.run([gev[gev["comp_unit_id"] == u] for u in range(16)], general=True)
    print(len(gev), sl.last_ms)
