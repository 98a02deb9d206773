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
Besides the headline the line carries: `roofline` (dominant kernel), `roofline_k1` (N=1: the depth-1 kernel measured in the
same run, so both temporal depths have a driver-run roofline fraction), `parity_preflight` (N>1: before anything is timed a
small awkward grid is advanced on the SAME decomposition, depth and exchange path, and every rank's slab is compared bit for
bit with a single-slab run of the same library on its own GPU - the single-GPU path is pinned to the oracle by the GPU
tests; a mismatch aborts the bench), `scaling_base_glups` / `parallel_efficiency` (N>1: rank 0's stand-alone rate on one
rank's share of the workload, measured in this run) and `strong_1024` (N>1: BASELINE configs[2], 1024^3 cut into N slabs,
with its own single-GPU base).
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
    ap.add_argument("--temporal-k", type=int, default=2, choices=[1, 2, 4],
                    help="ghost depth: 2 = two FTCS updates fused per HBM pass (temporal blocking); 4 = the same kernel with one "
                         "ghost exchange per FOUR updates (communication-avoiding); results are bit-identical")
    ap.add_argument("--grid", default="", help="nx,ny,nz: explicit global grid (experiments; overrides --workload)")
    ap.add_argument("--snapshot-every", type=int, default=0,
                    help="async snapshot of the field to pinned host memory every M bench steps inside the timed loop")
    ap.add_argument("--snapshot-stride", type=int, default=1,
                    help="with --snapshot-every: keep every s-th point per axis (decimated on the GPU before the copy)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-preflight", action="store_true", help="N>1: skip the bitwise parity pre-flight")
    ap.add_argument("--no-base", action="store_true", help="N>1: skip the stand-alone scaling base and the strong_1024 block")
    ap.add_argument("--no-k1", action="store_true", help="N=1: skip the depth-1 roofline block")
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



def bind_to_gpu_numa_node(index):
    """Run this process on the CPUs of the NUMA node its GPU hangs off (torchrun does not bind ranks): pinned buffers are
    then allocated next to the GPU's PCIe root, which matters when 8 ranks copy at once.  Best effort; returns what it did."""
    try:
        bus = subprocess.run(["nvidia-smi", "-i", str(index), "--query-gpu=pci.bus_id", "--format=csv,noheader"], capture_output=True,
                             text=True, timeout=20).stdout.strip().lower()
        if bus.startswith("0000"):
            bus = bus[4:]                                  # nvidia-smi prints an 8-digit domain, sysfs a 4-digit one
        node = int(open(f"/sys/bus/pci/devices/{bus}/numa_node").read())
        how = "from sysfs"
        if node < 0:
            # the platform does not say (virtualised PCI topology): spread the ranks evenly over the host's NUMA nodes in
            # device order, which is how 8-GPU boards are wired (GPUs 0-3 on one socket, 4-7 on the other)
            nodes = sorted(int(d[4:]) for d in os.listdir("/sys/devices/system/node") if d.startswith("node") and d[4:].isdigit())
            ngpu = int(os.environ.get("LOCAL_WORLD_SIZE", os.environ.get("WORLD_SIZE", "1")))
            if len(nodes) < 2 or ngpu < 2:
                return f"numa node unknown, {len(nodes)} host node(s): unbound"
            node, how = nodes[min(len(nodes) - 1, index * len(nodes) // ngpu)], f"guessed, {len(nodes)} host nodes"
        cpus = set()
        for part in open(f"/sys/devices/system/node/node{node}/cpulist").read().strip().split(","):
            lo, _, hi = part.partition("-")
            cpus.update(range(int(lo), int(hi or lo) + 1))
        cpus &= os.sched_getaffinity(0)
        if not cpus:
            return f"numa node {node}: no allowed CPUs"
        os.sched_setaffinity(0, cpus)
        return f"numa node {node} ({how}), {len(cpus)} CPUs"
    except Exception as e:      # no sysfs entry, no nvidia-smi, ...: run unbound
        return f"unbound ({type(e).__name__})"


def timed_steps(sim, tsteps, steps, warmup, barrier=None, snapshot=None):
    """`warmup` untimed + `steps` timed calls of heat3d_step(tsteps); device time from the library's CUDA events."""
    for _ in range(warmup):
        sim.step(tsteps)
    if barrier:
        barrier()
    else:
        sim.sync()
    dev_ms, launches, stencil, halo = 0.0, 0, 0, 0.0
    t0 = time.perf_counter()
    for it in range(steps):
        if snapshot:
            snapshot(it)
        sim.step(tsteps)
        tm = sim.timing()                     # waits for this step's events; device time of the call
        dev_ms += tm["last_call_ms"]
        launches += tm["last_call_launches"]
        stencil += tm["last_call_stencil_launches"]
        halo += tm["exposed_halo_ms"]
    return {"dev_ms": dev_ms, "launches": launches, "stencil_launches": stencil, "halo_ms": halo, "t0": t0}


def standalone_rate(h3, shape, arith, temporal_k, device, tsteps, steps=3, warmup=2):
    """GLUPS of ONE GPU on `shape` as a single slab (no exchange): the like-for-like base of a scaling figure."""
    nx, ny, nz = shape
    with h3.Heat3D(nx, ny, nz, arith=arith, temporal_k=temporal_k, devices=[device]) as sim:
        sim.init_reference_field()
        r = timed_steps(sim, tsteps, steps, warmup)
    return (nx - 2) * (ny - 2) * (nz - 2) * tsteps * steps / (r["dev_ms"] * 1e-3) / 1e9


def parity_preflight(h3, dist, torch, world, rank, local, arith, temporal_k, new_id):
    """N>1: a small awkward grid on the same decomposition / depth / exchange path as the timed run; each rank's slab must
    equal, bit for bit, the same planes of a single-slab run of this library on the rank's own GPU."""
    import numpy as np
    nx, ny, nz, nsteps = 150, 70, 23 * world, 15            # odd count: the depth-2 path ends with a depth-1 launch
    with h3.Heat3D(nx, ny, nz, arith=arith, temporal_k=temporal_k, nslabs=world, slab_first=rank, nlocal=1, devices=[local],
                   nccl_id=new_id(), flags=h3.FLAG_TIME_HALO) as sim:
        z0, z1 = sim.local_range()
        exchange = sim.exchange_name()
        sim.init_reference_field()
        sim.step(nsteps)
        mine = sim.download_planes(z0, z1 - z0)
        res = sim.step_residual(2)
        mine2 = sim.download_planes(z0, z1 - z0)
    with h3.Heat3D(nx, ny, nz, arith=arith, temporal_k=temporal_k, devices=[local]) as one:
        one.init_reference_field()
        one.step(nsteps)
        whole = one.download()
        res1 = one.step_residual(2)
        whole2 = one.download()
    ok = bool(np.array_equal(mine, whole[z0:z1]) and np.array_equal(mine2, whole2[z0:z1]) and res == res1)
    t = torch.tensor([1.0 if ok else 0.0], dtype=torch.float64, device="cuda")
    dist.all_reduce(t, op=dist.ReduceOp.MIN)
    return {"ranks": world, "exchange": exchange, "bitwise": bool(t.item() > 0.5), "grid": [nx, ny, nz], "updates": nsteps + 2,
            "temporal_k": temporal_k, "against": "single-slab run of the same library on each rank's GPU (oracle-pinned by the GPU tests)"}


def main():
    args = parse()
    if os.environ.get("HEAT3D_BENCH_WATCHDOG_S"):      # diagnostics: dump every thread's stack and exit if the run overstays
        import faulthandler
        faulthandler.dump_traceback_later(float(os.environ["HEAT3D_BENCH_WATCHDOG_S"]), exit=True)
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
    affinity = bind_to_gpu_numa_node(local)     # pinned host buffers then live on the GPU's own NUMA node
    n = args.gpus
    if world > 1:
        assert world == n, f"--gpus {n} but WORLD_SIZE={world}"
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    elif n > 1:
        raise SystemExit("for --gpus N>1 launch under torch.distributed.run with one rank per GPU")

    shape, wname, scaling, tsteps = workload(args, n)
    nx, ny, nz = shape
    cells = (nx - 2) * (ny - 2) * (nz - 2)

    def new_id():
        """A fresh 128-byte NCCL bootstrap id from rank 0 (one id serves one heat3d_create)."""
        if world == 1:
            return None
        box = [h3.nccl_unique_id() if rank == 0 else None]
        dist.broadcast_object_list(box, src=0)
        return box[0]

    arith = h3.ARITH_FMA if args.arith == "fma" else h3.ARITH_STRICT
    k = 1 if args.kernel == "baseline" else args.temporal_k
    peak, peak_src = 6650.0, "fallback"
    try:
        peak, peak_src = float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"]), "measured"
    except Exception:
        pass

    def max_over_ranks(x, op=None):
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=op or dist.ReduceOp.MAX)
        return float(t.item())

    # ---------------- N>1: parity of the decomposition before anything is timed ----------------
    preflight = None
    if world > 1 and not args.no_preflight:
        preflight = parity_preflight(h3, dist, torch, world, rank, local, arith, k, new_id)
        if not preflight["bitwise"]:
            if rank == 0:
                print(json.dumps({"error": "multi-GPU parity pre-flight failed", "parity_preflight": preflight}), flush=True)
            raise SystemExit(3)

    sim = h3.Heat3D(nx, ny, nz, arith=arith, kernel={"auto": 0, "baseline": 1, "tma": 2}[args.kernel], temporal_k=k,
                    nslabs=world, slab_first=rank, nlocal=1,
                    devices=[local], nccl_id=new_id(), flags=h3.FLAG_TIME_HALO if world > 1 else 0)
    z0, z1 = sim.local_range()

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        sim.sync()

    # ---------------- device-resident throughput ----------------
    sim.init_reference_field()
    ss = max(1, args.snapshot_stride)
    snap_shape = (z1 - z0, ny, nx) if ss == 1 else sim.strided_local_shape(ss, ss, ss)
    snap_buf, nsnap = (h3.pinned_empty(snap_shape), [0]) if args.snapshot_every else (None, [0])

    def snapshot(it):
        if snap_buf is not None and it % args.snapshot_every == 0:
            if ss == 1:
                sim.snapshot_async_local(snap_buf)      # D2H on the copy stream while the next updates run
            else:
                sim.snapshot_strided_async_local(snap_buf, ss, ss, ss)
            nsnap[0] += 1

    sampler = ClockSampler(local)

    def start_clock():
        barrier()
        if rank == 0:
            sampler.start()

    r = timed_steps(sim, tsteps, args.steps, args.warmup, barrier=start_clock, snapshot=snapshot)
    if snap_buf is not None:
        sim.snapshot_wait()
    barrier()
    wall_ms = (time.perf_counter() - r["t0"]) * 1e3
    clocks = sampler.stop() if rank == 0 else None
    launches, stencil_launches = r["launches"], r["stencil_launches"]      # kernels inside the headline's timed region
    extra_launches = 0                                                        # ... and inside the additional timed blocks of this line
    dev_ms = max_over_ranks(r["dev_ms"])
    wall_ms = max_over_ranks(wall_ms)
    halo_ms_min = max_over_ranks(r["halo_ms"], dist.ReduceOp.MIN) if world > 1 else r["halo_ms"]
    halo_ms = max_over_ranks(r["halo_ms"])
    nsnap = nsnap[0]
    total_updates = cells * tsteps * args.steps
    value = total_updates / (dev_ms * 1e-3) / 1e9
    kname = sim.kernel_name()

    def roofline_block(kernel_name, kk, local_cells, dev_ms_, n_launches, n_steps, grid_local):
        # 16 B per cell-update / k (one fp64 read + one write per k fused updates) x the cell-updates one launch performs
        bytes_per_launch = (16.0 / kk) * local_cells * tsteps * n_steps / max(1, n_launches)
        avg_launch_s = dev_ms_ * 1e-3 / max(1, n_launches)
        achieved = bytes_per_launch / avg_launch_s / 1e9
        traffic = None
        try:
            ent = json.load(open(os.path.join(ROOT, "profiles", "traffic.json"))).get(kernel_name)
            for e in (ent if isinstance(ent, list) else [ent] if ent else []):
                if e.get("grid") == grid_local:
                    traffic = e["dram_bytes_per_launch"]    # ncu dram__bytes_read.sum + dram__bytes_write.sum of one launch
        except Exception:
            pass
        return {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak, "traffic": traffic,
                "peak_source": f"{peak_src} (MEASURED_PEAKS.json hbm_gbs)" if peak_src == "measured" else "fallback 6.65 TB/s",
                "kernel": kernel_name, "temporal_k": kk, "algorithmic_bytes_per_cell_update": 16.0 / kk,
                "algorithmic_bytes_per_launch": bytes_per_launch, "avg_launch_ms": avg_launch_s * 1e3,
                "launches_timed": n_launches, "glups": local_cells * tsteps * n_steps / (dev_ms_ * 1e-3) / 1e9,
                "frac_of_nominal_8TBs": achieved / 8000.0}

    # roofline of the dominant kernel: algorithmic bytes per launch / average launch duration
    local_cells = (nx - 2) * (ny - 2) * (min(z1, nz - 1) - max(z0, 1))
    # (ghost depth 4 runs the depth-2 kernel: two updates per HBM pass either way)
    roofline = roofline_block(kname, min(k, 2), local_cells, dev_ms, stencil_launches, args.steps, [nx, ny, z1 - z0])

    # N=1: the depth-1 kernel in the same run (the other temporal depth's roofline fraction, driver-visible)
    roofline_k1 = None
    if n == 1 and k >= 2 and not args.no_k1:
        with h3.Heat3D(nx, ny, nz, arith=arith, kernel=2, temporal_k=1, devices=[local]) as s1:
            s1.init_reference_field()
            k1_steps = max(3, min(args.steps, 5))
            r1 = timed_steps(s1, tsteps, k1_steps, max(3, min(args.warmup, 3)))
            roofline_k1 = roofline_block(s1.kernel_name(), 1, local_cells, r1["dev_ms"], r1["stencil_launches"], k1_steps, [nx, ny, nz])
            extra_launches += r1["launches"]

    # N=1: the same kernel in nvcc's FMA contraction of the reference kernel (10 instead of 13 FP64 instructions per update,
    # bit-exact against the oracle's fma ordering, <= 1e-12 from the strict result) - reported beside the strict headline
    arith_fma = None
    if n == 1 and k >= 2 and args.arith == "strict" and not args.no_k1:
        with h3.Heat3D(nx, ny, nz, arith=h3.ARITH_FMA, temporal_k=k, devices=[local]) as sf:
            sf.init_reference_field()
            f_steps = max(3, min(args.steps, 5))
            rf = timed_steps(sf, tsteps, f_steps, max(3, min(args.warmup, 3)))
            arith_fma = {"glups": cells * tsteps * f_steps / (rf["dev_ms"] * 1e-3) / 1e9, "steps": f_steps, "kernel": sf.kernel_name(),
                         "note": "HEAT3D_ARITH_FMA: three fused multiply-adds per update, the contraction nvcc applies to the reference kernel"}
            extra_launches += rf["launches"]

    # N>1: like-for-like scaling base measured in this run (rank 0 alone on one rank's share), and the strong-scaling block
    scaling_base = parallel_eff = strong = None
    if world > 1 and not args.no_base:
        share = (nx, ny, z1 - z0) if scaling == "weak" else (nx, ny, nz)
        barrier()
        base = standalone_rate(h3, share, arith, k, local, tsteps) if rank == 0 else 0.0
        scaling_base = max_over_ranks(base)
        parallel_eff = value / (n * scaling_base)
        if args.workload == "auto" and not args.grid:
            sshape, ststeps = (1024, 1024, 1024), 125 * n
            scells = (sshape[0] - 2) * (sshape[1] - 2) * (sshape[2] - 2)
            barrier()
            with h3.Heat3D(*sshape, arith=arith, temporal_k=k, nslabs=world, slab_first=rank, nlocal=1, devices=[local],
                           nccl_id=new_id(), flags=h3.FLAG_TIME_HALO) as ss:
                ss.init_reference_field()

                def sbarrier():
                    torch.cuda.synchronize()
                    dist.barrier()
                    ss.sync()
                ssteps = max(3, min(args.steps, 5))
                rs = timed_steps(ss, ststeps, ssteps, 2, barrier=sbarrier)
                sbarrier()
                s_ms = max_over_ranks(rs["dev_ms"])
                s_halo = max_over_ranks(rs["halo_ms"])
                s_halo_min = max_over_ranks(rs["halo_ms"], dist.ReduceOp.MIN)
                extra_launches += rs["launches"]
            sval = scells * ststeps * ssteps / (s_ms * 1e-3) / 1e9
            barrier()
            sbase = max_over_ranks(standalone_rate(h3, sshape, arith, k, local, 125) if rank == 0 else 0.0)
            strong = {"workload": "1024^3 fp64 strong scaling, z-slab decomposition", "grid": list(sshape), "value": sval, "unit": "GLUPS",
                      "steps": ssteps, "updates_per_step": ststeps, "ms_per_step": s_ms / ssteps, "base_glups_1gpu": sbase,
                      "efficiency": sval / (n * sbase), "exposed_halo_ms_per_update": s_halo / (ststeps * ssteps),
                      "exposed_halo_ms_per_update_min_rank": s_halo_min / (ststeps * ssteps)}

    # ---------------- end to end through the C ABI with host buffers ----------------
    # Every step: its input field is copied from pinned host memory to the device, TSTEPS updates run, the result is
    # copied back to pinned host memory.  The streaming entry points keep PCIe busy in both directions while the GPU
    # steps: the upload of step i+1 (heat3d_upload_planes_async) and the download of step i-1 (heat3d_snapshot_async)
    # overlap the updates of step i.  `blocking` is the same loop with the blocking calls (upload -> step -> download).
    e2e = None
    if not args.no_e2e:
        planes = z1 - z0
        host_in = h3.pinned_empty((planes, ny, nx))
        host_out = [h3.pinned_empty((planes, ny, nx)) for _ in range(2)]
        sim.init_reference_field()
        sim.download_planes(z0, planes, out=host_in)
        e2e_steps = max(2, min(args.steps, 5))
        barrier()
        t0 = time.perf_counter()
        for _ in range(e2e_steps):
            sim.upload_planes(host_in, z0)              # H2D (pinned) + ghost refresh
            sim.step(tsteps)
            sim.download_planes(z0, planes, out=host_out[0])   # D2H, blocking
        barrier()
        blocking_s = max_over_ranks(time.perf_counter() - t0)
        barrier()
        t0 = time.perf_counter()
        sim.upload_planes_async(host_in, z0)
        for i in range(e2e_steps):
            sim.upload_commit()                         # input i is the current field
            if i + 1 < e2e_steps:
                sim.upload_planes_async(host_in, z0)    # input i+1 crosses PCIe while step i computes
            sim.step(tsteps)
            sim.snapshot_async_local(host_out[i & 1])   # result i leaves while step i+1 computes
        sim.snapshot_wait()
        barrier()
        e2e_s = max_over_ranks(time.perf_counter() - t0)
        streamed_ok = bool(np.array_equal(host_out[(e2e_steps - 1) & 1], host_out[e2e_steps & 1])) if e2e_steps >= 2 else True
        e2e = {"value": cells * tsteps * e2e_steps / e2e_s / 1e9, "unit": "GLUPS", "h2d_bytes_per_step": int(host_in.nbytes),
               "d2h_bytes_per_step": int(host_out[0].nbytes), "steps": e2e_steps, "ms_per_step": e2e_s / e2e_steps * 1e3,
               "path": "heat3d_upload_planes_async/_commit -> heat3d_step -> heat3d_snapshot_async, pinned host buffers, wall clock; "
                       "transfers of neighbouring steps overlap the updates",
               "blocking": {"value": cells * tsteps * e2e_steps / blocking_s / 1e9, "ms_per_step": blocking_s / e2e_steps * 1e3,
                            "path": "heat3d_upload_planes -> heat3d_step -> heat3d_download_planes"},
               "same_result_every_step": streamed_ok, "cpu_affinity": affinity}
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
                "roofline": roofline, "roofline_k1": roofline_k1, "arith_fma": arith_fma, "cpu_baseline": cpu, "e2e": e2e, "gpu_launches": int(launches), "gpu_launches_other_blocks": int(extra_launches), "clocks": clocks,
                "parity_preflight": preflight, "scaling_base_glups": scaling_base, "parallel_efficiency": parallel_eff, "strong_1024": strong,
                "wall_ms_per_step": wall_ms / args.steps, "snapshots": {"count": nsnap, "bytes_each_per_rank": int(snap_buf.nbytes), "stride": ss, "every_steps": args.snapshot_every,
                              "headline": "wall_glups_incl_snapshots"} if snap_buf is not None else None,
                "wall_glups_incl_snapshots": total_updates / (wall_ms * 1e-3) / 1e9, "exposed_halo_ms_per_update": halo_ms / (tsteps * args.steps) if n > 1 else None,
                "exposed_halo_ms_per_update_min_rank": halo_ms_min / (tsteps * args.steps) if n > 1 else None,
                "exchange": sim.exchange_name(),
                "roofline_glups_at_peak": peak / (16.0 / min(k, 2))}
        print(json.dumps(line), flush=True)
    sim.close()
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
