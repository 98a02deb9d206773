"""Worker for the multi-process tests (launched by torch.distributed.run, one rank per slab).

backend gloo (CPU, no GPU needed): the HOST-side logic of the N>1 path — slab bounds, shipping the 128-byte
bootstrap id, neighbour exchange order, max/sum reductions — with the oracle's plane stepper standing in for
the device (this file is under tests/: it may use the oracle).
backend nccl (GPU): the real thing — one heat3d context per rank, ghost planes over NCCL send/recv — compared
bit for bit with the single-domain oracle.
"""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as ge  # noqa: E402
from oracle import oracle  # noqa: E402


def main():
    backend = sys.argv[1]
    nx, ny, nz, nsteps = (int(v) for v in sys.argv[2:6])
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    h3 = ge.load_package()
    _, _, D = oracle.coefficients(nx, ny, nz)
    ref = oracle.steps(oracle.init_field(nx, ny, nz), D, nsteps)
    ref_prev = oracle.steps(oracle.init_field(nx, ny, nz), D, nsteps - 1)
    z0, z1 = h3.slab_bounds(nz, world)[rank]

    if backend == "gloo":
        dist.init_process_group("gloo")
        # bootstrap-id plumbing: 128 opaque bytes from rank 0 to everyone
        box = [bytes(range(128)) if rank == 0 else None]
        dist.broadcast_object_list(box, src=0)
        assert box[0] == bytes(range(128))
        # slab with one ghost plane per side, stepped by the oracle, ghosts exchanged like the library does
        lo, hi = max(z0 - 1, 0), min(z1 + 1, nz)
        cur = oracle.init_field(nx, ny, nz)
        nxt = cur.copy()
        for _ in range(nsteps):
            oracle.step_planes(cur, nxt, D, z0, z1)
            reqs = []
            if rank > 0:
                reqs.append(dist.isend(torch.from_numpy(nxt[z0]), rank - 1))
                reqs.append(dist.irecv(torch.from_numpy(nxt[z0 - 1]), rank - 1))
            if rank < world - 1:
                reqs.append(dist.isend(torch.from_numpy(nxt[z1 - 1]), rank + 1))
                reqs.append(dist.irecv(torch.from_numpy(nxt[z1]), rank + 1))
            for r in reqs:
                r.wait()
            cur, nxt = nxt, cur
        assert np.array_equal(cur[lo:hi], ref[lo:hi]), f"rank {rank}: slab differs from single domain"
        # residual = max over ranks of the local maxima (interior planes of the slab only)
        a, b = max(z0, 1), min(z1, nz - 1)
        local = float(np.max(np.abs(cur[a:b, 1:-1, 1:-1] - ref_prev[a:b, 1:-1, 1:-1]))) if b > a else 0.0
        t = torch.tensor([local], dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        assert max(t.item(), np.finfo(np.float64).tiny) == oracle.residual(ref, ref_prev)
        # ghost depth 4 (temporal_k = 4): blocks of four updates with ONE exchange of four planes per side.  Update j of a
        # block covers the owned planes plus 4-j ghost planes on every side that has a neighbour - the schedule of the
        # library's step_block4 (pass A = updates 1-2 ending two ghost planes deep, pass B = updates 3-4) - and the tail of
        # nsteps % 4 updates exchanges after every update.
        if all(b - a >= 4 for a, b in h3.slab_bounds(nz, world)):
            G = 4
            cur = oracle.init_field(nx, ny, nz)
            nxt = cur.copy()

            def exchange(buf, depth):
                reqs = []
                if rank > 0:
                    reqs.append(dist.isend(torch.from_numpy(buf[z0:z0 + depth]), rank - 1))
                    reqs.append(dist.irecv(torch.from_numpy(buf[z0 - depth:z0]), rank - 1))
                if rank < world - 1:
                    reqs.append(dist.isend(torch.from_numpy(buf[z1 - depth:z1]), rank + 1))
                    reqs.append(dist.irecv(torch.from_numpy(buf[z1:z1 + depth]), rank + 1))
                for r in reqs:
                    r.wait()

            done = 0
            while nsteps - done >= 4:
                for j in range(1, 5):
                    ext = G - j
                    oracle.step_planes(cur, nxt, D, z0 - ext if rank > 0 else z0, z1 + ext if rank < world - 1 else z1)
                    cur, nxt = nxt, cur
                exchange(cur, G)
                nxt[max(z0 - G, 0):min(z1 + G, nz)] = cur[max(z0 - G, 0):min(z1 + G, nz)]      # both buffers carry the Dirichlet shell
                done += 4
            while done < nsteps:
                oracle.step_planes(cur, nxt, D, z0, z1)
                exchange(nxt, G)
                cur, nxt = nxt, cur
                done += 1
            lo4, hi4 = max(z0 - G, 0), min(z1 + G, nz)
            assert np.array_equal(cur[lo4:hi4], ref[lo4:hi4]), f"rank {rank}: ghost-depth-4 slab differs from single domain"
    else:
        local = int(os.environ.get("LOCAL_RANK", rank))
        torch.cuda.set_device(local)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
        for kernel, tk, exchange in ((2, 1, "ipc-put"), (2, 2, "ipc-put"), (2, 4, "ipc-put"), (2, 1, "peer-store"), (2, 2, "peer-store"),
                                     (2, 2, "nccl"), (2, 4, "nccl"), (1, 1, "nccl")):
            os.environ["HEAT3D_EXCHANGE"] = {"nccl": "nccl", "ipc-put": "ipc", "peer-store": "store"}[exchange]
            box = [h3.nccl_unique_id() if rank == 0 else None]      # a bootstrap id serves ONE communicator
            dist.broadcast_object_list(box, src=0)
            with h3.Heat3D(nx, ny, nz, kernel=kernel, temporal_k=tk, nslabs=world, slab_first=rank, nlocal=1, devices=[local],
                           nccl_id=box[0], flags=h3.FLAG_TIME_HALO) as sim:
                assert sim.local_range() == (z0, z1)
                assert sim.exchange_name() == exchange, (sim.exchange_name(), exchange)
                sim.init_reference_field()
                sim.step(nsteps - 1)
                res = sim.step_residual(1)
                got = sim.download_planes(z0, z1 - z0)
                err = sim.mean_abs_error_vs_y()
                assert np.array_equal(got, ref[z0:z1]), f"rank {rank} kernel {kernel} k {tk} {exchange}: slab differs from single domain"
                assert res == oracle.residual(ref, ref_prev), (res, oracle.residual(ref, ref_prev))
                exp = oracle.mean_abs_error_vs_y(ref)
                assert abs(err - exp) < 1e-12, (err, exp, err - exp)   # serial sum (oracle) vs tree sum (device)
                # upload path: random field, ghosts refreshed by the exchange; then re-init and run again
                T = np.random.default_rng(5).random((nz, ny, nx))
                sim.upload_planes(np.ascontiguousarray(T[z0:z1]), z0)
                sim.step(5)
                assert np.array_equal(sim.download_planes(z0, z1 - z0), oracle.steps(T.copy(), D, 5)[z0:z1])
                sim.init_reference_field()
                sim.step(nsteps)
                assert np.array_equal(sim.download_planes(z0, z1 - z0), ref[z0:z1])
                # the reference's loop, residual examined every iteration (collective): iteration count and field must be the
                # single domain's whether the run converges on the first or the second update of a depth-2 pass, or not at all
                for eps, iter_max in ((0.3, 200), (0.28, 200), (0.18, 200), (1e-30, 13)):
                    To = oracle.init_field(nx, ny, nz)
                    want = oracle.run(To, D, iter_max, eps)
                    sim.init_reference_field()
                    assert sim.run(iter_max, eps, check_every=1) == want, (eps, want)
                    assert np.array_equal(sim.download_planes(z0, z1 - z0), To[z0:z1]), eps
            dist.barrier()
    dist.barrier()
    dist.destroy_process_group()
    print(f"rank {rank}/{world} {backend} ok", flush=True)


if __name__ == "__main__":
    main()
