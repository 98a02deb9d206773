"""Runs one stencil variant for a few steps (for ncu).  python tools/prof_one.py <n> <steps> <kernel:1|2> [tx ty ry st cps] [arith]"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as ge  # noqa: E402

h3 = ge.load_package()
n, steps, kernel = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3])
tune = tuple(int(v) for v in sys.argv[4:9]) if len(sys.argv) >= 9 else None
arith = int(sys.argv[9]) if len(sys.argv) > 9 else 0
with h3.Heat3D(n, n, n, kernel=kernel, tune=tune, arith=arith) as sim:
    sim.init_reference_field()
    sim.step(steps)
    sim.sync()
    t = sim.timing()
    print(sim.kernel_name(), "ms/step", t["last_call_ms"] / steps)
