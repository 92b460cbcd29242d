import sys, os, subprocess
root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
code = '''
import sys, os
sys.path.insert(0, %r); sys.path.insert(0, os.path.join(%r, "profiles"))
import perf_probe
perf_probe.probe("cfg2 chunk_kb=" + os.environ.get("HBTGPU_CHUNK_KB", "-"), 2, 192, 8, 10100, n_samples=100_000)
''' % (root, root)
for kb in sys.argv[1:]:
    env = dict(os.environ, HBTGPU_CHUNK_KB=kb)
    r = subprocess.run([sys.executable, "-c", code], env=env, capture_output=True, text=True)
    print((r.stdout.strip().splitlines() or [r.stderr[-300:]])[-1], flush=True)
