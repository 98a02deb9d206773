#!/usr/bin/env python
"""bench.py — GLUPS (fp64 lattice updates per second) of the heat3D hot path on N B200s of one box.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload auto|c512|weak|strong]
  python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P \
         bench.py --gpus N --steps K --warmup W

A bench "step" is one pass of the hot path over the workload: TSTEPS consecutive FTCS updates of the whole
grid (default 1000, so the default K=10 is BASELINE config 2: 512^3 x 10000 updates on one B200).
  value  GLUPS with the field resident in HBM (device-timed with CUDA events on the library's compute stream,
         max over ranks).
  e2e    the same metric through the C ABI with HOST buffers: every step uploads the field from pinned host
         memory, advances TSTEPS updates and downloads the result — copies inside the timed region.
Workloads: N=1 -> 512^3 (BASELINE configs[1]).  N>1 -> weak scaling, 1024x1024x512 points per GPU, z-slabs
(configs[3]); --workload strong gives 1024^3 over N GPUs (configs[2]).
--impl reference times the reference's own host program (oracle/_ref/heat3D_ref, built from the reference's
src/heat3D.cpp) on the box's CPU: one rank = one core, which is all the parallelism that program has without an
MPI launcher.  torch / torch.distributed are used for rendezvous, barriers and the max over ranks only.
"""
import argparse
import json
import os
import subprocess
import sys
import tempfile
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="auto", choices=["auto", "c512", "weak", "strong"])
    ap.add_argument("--tsteps", type=int, default=0, help="FTCS updates per bench step (0 = workload default)")
    ap.add_argument("--arith", default="strict", choices=["strict", "fma"])
    ap.add_argument("--kernel", default="auto", choices=["auto", "tma", "baseline"])
    ap.add_argument("--temporal-k", type=int, default=2, choices=[1, 2],
                    help="FTCS updates fused per HBM pass (2 = temporal blocking, bit-identical results)")
    ap.add_argument("--grid", default="", help="nx,ny,nz: explicit global grid (experiments; overrides --workload)")
    ap.add_argument("--snapshot-every", type=int, default=0,
                    help="async snapshot of the field to pinned host memory every M bench steps inside the timed loop")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    return ap.parse_args()


def workload(args, n):
    w = args.workload
    if w == "auto":
        w = "c512" if n == 1 else "weak"
    if w == "c512":
        shape, name, scaling, ts = (512, 512, 512 * n), "512^3 fp64 heat3D, Dirichlet BC" if n == 1 else f"512x512x{512 * n} fp64 (512^3 per GPU)", "weak", 1000
    elif w == "weak":
        shape, name, scaling, ts = (1024, 1024, 512 * n), f"weak scaling 1024x1024x512-per-GPU fp64, z-slabs ({1024}x{1024}x{512 * n})", "weak", 250
    else:
        shape, name, scaling, ts = (1024, 1024, 1024), "1024^3 fp64 strong scaling, z-slab decomposition", "strong", 125 * n
    if args.grid:
        shape = tuple(int(v) for v in args.grid.split(","))
        name, scaling = "custom grid %dx%dx%d" % shape, "strong"
    if args.tsteps:
        ts = args.tsteps
    return shape, name, scaling, ts


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled while the timed region runs."""
    Q = "clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown," \
        "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, index):
        self.samples, self.proc, self.index = [], None, index

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "100"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.th = threading.Thread(target=self._read, daemon=True)
            self.th.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.samples.append([c.strip() for c in line.split(",")])

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        self.th.join(timeout=2)
        sm = sorted(float(s[0]) for s in self.samples if s and s[0].replace(".", "").isdigit())
        mx = [float(s[1]) for s in self.samples if len(s) > 1 and s[1].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [nm for i, nm in enumerate(names) if any(len(s) > 3 + i and s[3 + i] == "Active" for s in self.samples)]
        pw = [float(s[2]) for s in self.samples if len(s) > 2 and s[2].replace(".", "").isdigit()]
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None, "reasons": reasons,
                "power_w_max": max(pw) if pw else None, "samples": len(self.samples)}


def cpu_reference_sample(shape, tsteps_sample):
    """Time the reference's own host loop (oracle/_ref) — or, if that binary is absent, the C port — on a bounded sample."""
    from oracle import oracle
    nx, ny, nz = shape
    cells = (nx - 2) * (ny - 2) * (nz - 2)
    if oracle.have_reference_binary():
        with tempfile.TemporaryDirectory() as wd:
            r = oracle.run_reference_binary(nx, ny, nz, tsteps_sample, "1e-30", wd, raw=False, dump=False)
        secs = r["loop_seconds"]
        return {"value": cells * tsteps_sample / secs / 1e9, "unit": "GLUPS", "cores": 1, "kind": "reference",
                "sample": f"{nx}x{ny}x{nz} x {tsteps_sample} updates by the reference's host program (its own loop timer: {secs:.3f} s)",
                "seconds": secs}
    import numpy as np  # noqa: F401
    _, _, D = oracle.coefficients(nx, ny, nz)
    T = oracle.init_field(nx, ny, nz)
    t0 = time.perf_counter()
    oracle.steps(T, D, tsteps_sample, threads=1)
    secs = time.perf_counter() - t0
    return {"value": cells * tsteps_sample / secs / 1e9, "unit": "GLUPS", "cores": 1, "kind": "port",
            "sample": f"{nx}x{ny}x{nz} x {tsteps_sample} updates by the C restatement of the reference loop", "seconds": secs}


def sample_steps_for(shape, seconds=15.0, glups_guess=0.15):
    cells = (shape[0] - 2) * (shape[1] - 2) * (shape[2] - 2)
    return max(1, int(seconds * glups_guess * 1e9 / cells))


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    n = args.gpus
    full_shape, name, scaling, _ = workload(args, n)
    # bounded sample: one GPU's z-slab of the workload (the whole grid at N=1), a few updates per bench step
    shape = (full_shape[0], full_shape[1], max(3, full_shape[2] // n))
    per_step = max(1, sample_steps_for(shape, seconds=4.0))
    if args.warmup > 0:
        cpu_reference_sample(shape, per_step)                 # untimed warm-up run (page cache, allocator, clocks)
    r = cpu_reference_sample(shape, args.steps * per_step)
    ms = r["seconds"] / args.steps * 1e3
    line = {"impl": "reference", "metric": "GLUPS (fp64 lattice updates/s)", "value": r["value"], "unit": "GLUPS", "n_gpus": n,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": scaling,
            "vs_baseline": None, "dtype": "f64", "data": "synthetic (reference IC/BC: zero interior, linear-in-y Dirichlet shell)",
            "config": {"workload": name, "grid": list(full_shape), "sample_grid": list(shape), "updates_per_step": per_step,
                       "note": "reference host program on 1 CPU core; each step is a bounded sample: one GPU's slab of the grid, "
                               "a few updates; warm-up is one separate untimed run of the program"},
            "cpu_baseline": {k: r[k] for k in ("value", "unit", "cores", "kind", "sample")},
            "e2e": {"value": r["value"], "unit": "GLUPS", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)


def main():
    args = parse()
    if args.impl == "reference":
        return run_reference(args)

    import numpy as np
    import torch
    import torch.distributed as dist
    import __graft_entry__ as ge
    h3 = ge.load_package()

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not os.path.exists(h3.LIB_PATH):
        # not built yet (fresh checkout): compile the native library once - there is nothing else this could run on
        if rank == 0:
            ge.build()
        else:
            for _ in range(900):
                if os.path.exists(h3.LIB_PATH):
                    break
                time.sleep(1.0)
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device (the product has no CPU path); use --impl reference for the CPU baseline")
    torch.cuda.set_device(local)
    n = args.gpus
    if world > 1:
        assert world == n, f"--gpus {n} but WORLD_SIZE={world}"
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    elif n > 1:
        raise SystemExit("for --gpus N>1 launch under torch.distributed.run with one rank per GPU")

    shape, wname, scaling, tsteps = workload(args, n)
    nx, ny, nz = shape
    cells = (nx - 2) * (ny - 2) * (nz - 2)

    nccl_id = None
    if world > 1:
        box = [h3.nccl_unique_id() if rank == 0 else None]
        dist.broadcast_object_list(box, src=0)
        nccl_id = box[0]
    sim = h3.Heat3D(nx, ny, nz, arith=h3.ARITH_FMA if args.arith == "fma" else h3.ARITH_STRICT,
                    kernel={"auto": 0, "baseline": 1, "tma": 2}[args.kernel], temporal_k=1 if args.kernel == "baseline" else args.temporal_k,
                    nslabs=world, slab_first=rank, nlocal=1,
                    devices=[local], nccl_id=nccl_id, flags=h3.FLAG_TIME_HALO if world > 1 else 0)
    z0, z1 = sim.local_range()

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        sim.sync()

    def max_over_ranks(x, op=None):
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=op or dist.ReduceOp.MAX)
        return float(t.item())

    # ---------------- device-resident throughput ----------------
    sim.init_reference_field()
    for _ in range(args.warmup):
        sim.step(tsteps)
    barrier()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    dev_ms, launches, stencil_launches, halo_ms = 0.0, 0, 0, 0.0
    snap_buf, nsnap = (h3.pinned_empty((z1 - z0, ny, nx)), 0) if args.snapshot_every else (None, 0)
    t0 = time.perf_counter()
    for it in range(args.steps):
        if snap_buf is not None and it % args.snapshot_every == 0:
            sim.snapshot_async_local(snap_buf)      # D2H on the copy stream while the next updates run
            nsnap += 1
        sim.step(tsteps)
        tm = sim.timing()                     # waits for this step's events; device time of the call
        dev_ms += tm["last_call_ms"]
        launches += tm["last_call_launches"]
        stencil_launches += tm["last_call_stencil_launches"]
        halo_ms += tm["exposed_halo_ms"]
    if snap_buf is not None:
        sim.snapshot_wait()
    barrier()
    wall_ms = (time.perf_counter() - t0) * 1e3
    clocks = sampler.stop() if rank == 0 else None
    dev_ms = max_over_ranks(dev_ms)
    wall_ms = max_over_ranks(wall_ms)
    halo_ms_min = max_over_ranks(halo_ms, dist.ReduceOp.MIN) if world > 1 else halo_ms
    halo_ms = max_over_ranks(halo_ms)
    total_updates = cells * tsteps * args.steps
    value = total_updates / (dev_ms * 1e-3) / 1e9
    kname = sim.kernel_name()

    # roofline of the dominant kernel: algorithmic bytes per launch / average launch duration
    local_cells = (nx - 2) * (ny - 2) * (min(z1, nz - 1) - max(z0, 1))
    k = 1 if args.kernel == "baseline" else args.temporal_k
    # 16 B per cell-update / k (one fp64 read + one write per k fused updates) x the cell-updates one launch performs
    bytes_per_launch = (16.0 / k) * local_cells * tsteps * args.steps / max(1, stencil_launches)
    avg_launch_s = dev_ms * 1e-3 / max(1, stencil_launches)
    achieved = bytes_per_launch / avg_launch_s / 1e9
    peak, peak_src = 6650.0, "fallback"
    try:
        peak, peak_src = float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"]), "measured"
    except Exception:
        pass
    traffic = None
    try:
        tr = json.load(open(os.path.join(ROOT, "profiles", "traffic.json")))
        ent = tr.get(kname)
        if ent and ent.get("grid") == [nx, ny, z1 - z0]:
            traffic = ent["dram_bytes_per_launch"]
    except Exception:
        pass
    roofline = {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak, "traffic": traffic,
                "peak_source": f"{peak_src} (MEASURED_PEAKS.json hbm_gbs)" if peak_src == "measured" else "fallback 6.65 TB/s",
                "kernel": kname, "temporal_k": k, "algorithmic_bytes_per_cell_update": 16.0 / k,
                "algorithmic_bytes_per_launch": bytes_per_launch, "avg_launch_ms": avg_launch_s * 1e3,
                "launches_timed": stencil_launches}

    # ---------------- end to end through the C ABI with host buffers ----------------
    e2e = None
    if not args.no_e2e:
        planes = z1 - z0
        host_in = h3.pinned_empty((planes, ny, nx))
        host_out = h3.pinned_empty((planes, ny, nx))
        sim.init_reference_field()
        sim.download_planes(z0, planes, out=host_in)
        e2e_steps = max(1, min(args.steps, 5))
        barrier()
        t0 = time.perf_counter()
        for _ in range(e2e_steps):
            sim.upload_planes(host_in, z0)              # H2D (pinned) + ghost refresh
            sim.step(tsteps)
            sim.download_planes(z0, planes, out=host_out)   # D2H, blocking
        barrier()
        e2e_s = max_over_ranks(time.perf_counter() - t0)
        e2e = {"value": cells * tsteps * e2e_steps / e2e_s / 1e9, "unit": "GLUPS", "h2d_bytes_per_step": int(host_in.nbytes),
               "d2h_bytes_per_step": int(host_out.nbytes), "steps": e2e_steps, "ms_per_step": e2e_s / e2e_steps * 1e3,
               "path": "heat3d_upload_planes -> heat3d_step -> heat3d_download_planes, pinned host buffers, wall clock"}
        del host_in, host_out

    # ---------------- CPU baseline beside it (rank 0, N=1 only) ----------------
    cpu = None
    if rank == 0 and n == 1 and not args.no_cpu_baseline:
        cpu = {k: v for k, v in cpu_reference_sample(shape, sample_steps_for(shape, seconds=15.0)).items() if k != "seconds"}

    if rank == 0:
        line = {"metric": "GLUPS (fp64 lattice updates/s)", "value": value, "unit": "GLUPS", "n_gpus": n, "steps": args.steps,
                "warmup": args.warmup, "ms_per_step": dev_ms / args.steps, "higher_is_better": True, "scaling": scaling,
                "vs_baseline": None, "dtype": "f64", "data": "synthetic (reference IC/BC: zero interior, linear-in-y Dirichlet shell)",
                "config": {"workload": wname, "grid": [nx, ny, nz], "updates_per_step": tsteps, "arith": args.arith, "temporal_k": k,
                           "decomposition": f"{n} z-slab(s), one process per GPU" + (f", ghost planes by {sim.exchange_name()} overlapped with interior compute" if n > 1 else ""),
                           "l2": "inputs exceed L2 (two %.1f GiB buffers per GPU vs 126 MB)" % ((z1 - z0) * ny * nx * 8 / 2**30)},
                "roofline": roofline, "cpu_baseline": cpu, "e2e": e2e, "gpu_launches": int(launches), "clocks": clocks,
                "wall_ms_per_step": wall_ms / args.steps, "snapshots": {"count": nsnap, "bytes_each_per_rank": int(snap_buf.nbytes)} if snap_buf is not None else None,
                "wall_glups_incl_snapshots": total_updates / (wall_ms * 1e-3) / 1e9, "exposed_halo_ms_per_update": halo_ms / (tsteps * args.steps) if n > 1 else None,
                "exposed_halo_ms_per_update_min_rank": halo_ms_min / (tsteps * args.steps) if n > 1 else None,
                "exchange": sim.exchange_name(),
                "roofline_glups_at_peak": peak / (16.0 / k)}
        print(json.dumps(line), flush=True)
    sim.close()
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
