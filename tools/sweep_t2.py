"""Times the depth-2 (temporal_k=2) kernel variants.  python tools/sweep_t2.py [n] [steps]   (SWEEP_SUSTAINED=1 for long runs)"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as ge  # noqa: E402

h3 = ge.load_package()
n = int(sys.argv[1]) if len(sys.argv) > 1 else 512
shape = tuple(int(v) for v in os.environ["SWEEP_SHAPE"].split(",")) if os.environ.get("SWEEP_SHAPE") else (n, n, n)
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 200
sustained = os.environ.get("SWEEP_SUSTAINED") == "1"
warm, reps = (steps, 2) if sustained else (20, 3)
peak = 6538.6
try:
    peak = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"]
except Exception:
    pass
# a tune is "cx,cy,ry,stages,ctas" + the shared-memory access mode letter: L plain loads, S swapped x-neighbour loads,
# P parity-split mid plane, X both (kernels_t2.cuh)
TUNES = [t for t in os.environ.get("SWEEP_TUNES", "64,32,4,4,2P;64,32,4,4,2L;64,32,4,4,2X;64,16,2,4,3P;128,32,4,4,1P;64,64,4,4,1P").split(";")]
# chunk length 0 = the guided plan (long chunks first, short ones last); > 0 = uniform chunks of that length
CHUNKS = [int(c) for c in os.environ.get("SWEEP_CHUNKS", "0,64").split(",")]
cells = (shape[0] - 2) * (shape[1] - 2) * (shape[2] - 2)
for arith in ((0,) if os.environ.get("SWEEP_STRICT_ONLY") else (0, 1)):
    for tune in TUNES:
        for ch in CHUNKS:
            base, xmode = tune[:-1], {"L": 0, "S": 1, "P": 2, "X": 3, "Q": 6}[tune[-1]]
            os.environ["HEAT3D_T2_TUNE"] = f"{base},{ch},{xmode}"
            try:
                with h3.Heat3D(*shape, arith=arith, temporal_k=2) as sim:
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
                print("%-22s chunk=%-3d arith=%d  %.4f ms/update  %7.1f GLUPS  alg %7.1f GB/s (8 B/update)  %.3f of measured HBM peak"
                      % (name, ch, arith, best / steps, glups, glups * 8, glups * 8 / peak), flush=True)
            except Exception as e:
                print("FAILED", tune, ch, e, flush=True)
