"""Runs the depth-2 kernel for a few updates (for ncu).  python tools/prof_t2.py <n> <updates> [arith]   (HEAT3D_T2_TUNE selects the variant)"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as ge  # noqa: E402

h3 = ge.load_package()
n, steps = int(sys.argv[1]), int(sys.argv[2])
arith = int(sys.argv[3]) if len(sys.argv) > 3 else 0
shape = tuple(int(v) for v in os.environ["PROF_SHAPE"].split(",")) if os.environ.get("PROF_SHAPE") else (n, n, n)
with h3.Heat3D(*shape, temporal_k=2, arith=arith) as sim:
    sim.init_reference_field()
    sim.step(steps)
    sim.sync()
    print(sim.kernel_name(), "ms/update", sim.timing()["last_call_ms"] / steps)
