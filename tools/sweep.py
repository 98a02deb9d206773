"""Tuning harness: times every stencil-kernel variant on one grid (default 512^3) and prints a table.
Usage (on the GPU box):  python tools/sweep.py [n] [steps]"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as ge  # noqa: E402

h3 = ge.load_package()
n = int(sys.argv[1]) if len(sys.argv) > 1 else 512
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 200
sustained = os.environ.get("SWEEP_SUSTAINED") == "1"      # long runs under the power cap: report the LAST repetition
warm, reps = (steps, 2) if sustained else (20, 3)
peak = 6538.6
try:
    peak = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"]
except Exception:
    pass
TUNES = [(128, 8, 2, 4, 3), (128, 8, 4, 4, 4), (128, 16, 2, 4, 1), (128, 16, 4, 4, 2), (128, 32, 4, 4, 1), (64, 16, 4, 4, 4)]
cells = (n - 2) ** 3
rows = []
extra = [dict(kernel=2, tune=t + (c,)) for t in [(128, 8, 2, 4, 3), (128, 8, 4, 4, 4), (128, 16, 4, 4, 2)] for c in (12, 24)]
if os.environ.get("SWEEP_ONLY_EXTRA"):
    TUNES = []
for arith in (0, 1):
    for kw in ([dict(kernel=1)] if TUNES else []) + [dict(kernel=2, tune=t) for t in TUNES] + (extra if arith == 0 else []):
        try:
            with h3.Heat3D(n, n, n, arith=arith, **kw) as sim:
                sim.init_reference_field()
                sim.step(warm)
                sim.sync()
                best = 1e30
                for _ in range(reps):
                    sim.step(steps)
                    last = sim.timing()["last_call_ms"]
                    best = last if sustained else min(best, last)
                name = sim.kernel_name()
            glups = cells * steps / (best * 1e-3) / 1e9
            rows.append((name, str(kw.get("tune", ""))[-5:], arith, best / steps, glups, glups * 16, glups * 16 / peak))
            print("%-24s %-5s arith=%d  %.4f ms/step  %7.1f GLUPS  %7.1f GB/s  %.3f of measured HBM peak" % rows[-1], flush=True)
        except Exception as e:  # keep sweeping
            print("FAILED", kw, e, flush=True)
