"""A/B of CUDA-graph replay vs plain launches on launch-bound (tiny) grids.  HEAT3D_NO_GRAPHS=1 disables graphs."""
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as ge  # noqa: E402

h3 = ge.load_package()
for n in (27, 64, 128):
    for k in (1, 2):
        with h3.Heat3D(n, n, n, temporal_k=k) as sim:
            sim.init_reference_field()
            sim.step(640)
            sim.sync()
            t0 = time.perf_counter()
            sim.step(64000)
            sim.sync()
            wall = time.perf_counter() - t0
            print(f"n={n} k={k} graphs={'off' if os.environ.get('HEAT3D_NO_GRAPHS') else 'on'}: {wall / 64000 * 1e6:.2f} us/update wall, "
                  f"{sim.timing()['last_call_ms'] / 64000 * 1e3:.2f} us/update device", flush=True)
