"""Regenerates tests/golden/* from the reference's own host program (oracle/_ref/heat3D_ref).

Run in the dev container (where /root/reference exists):  python tests/golden/make_golden.py
The reference has no fixtures of its own for this path; these are outputs of
its src/heat3D.cpp compiled here with the single-rank MPI shim and the interior-update statement
it lacks (oracle/Makefile, oracle/ref_patch.sed).  Committed: manifest.json (stdout report lines,
out.dat md5/size, sha256 of the raw fp64 field, a few probe values) and the small raw fields
as compressed .npz.
"""
import hashlib
import json
import os
import sys
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import oracle  # noqa: E402

CASES = [
    # name, nx, ny, nz, iter_max, eps (as typed on the command line), store raw field?
    ("c27_converge", 27, 27, 27, 100000, "1e-6", True),
    ("aniso_33x17x65", 33, 17, 65, 300, "1e-30", True),
    ("tiny_5x4x3", 5, 4, 3, 7, "1e-30", True),
    ("c64_1000", 64, 64, 64, 1000, "1e-30", False),
]


def main():
    oracle.build(ref=True)
    here = os.path.dirname(os.path.abspath(__file__))
    manifest = {}
    for name, nx, ny, nz, it, eps, store in CASES:
        with tempfile.TemporaryDirectory() as wd:
            r = oracle.run_reference_binary(nx, ny, nz, it, eps, wd)
            dump = open(r["dump_path"], "rb").read()
            lines = [ln for ln in r["stdout"].splitlines()
                     if not ln.startswith("executable:") and not ln.startswith("Computational time")]
            raw = r["raw"]
            pts = [(nz // 2, ny // 2, nx // 2), (1, 1, 1), (nz - 2, ny - 2, nx - 2), (nz // 3, ny // 2, nx // 4)]
            probes = {f"{k},{j},{i}": float(raw[k, j, i]).hex() for (k, j, i) in pts}
            manifest[name] = dict(args=[nx, ny, nz, it, eps], stdout=lines,
                                  dump_md5=hashlib.md5(dump).hexdigest(), dump_bytes=len(dump),
                                  raw_sha256=hashlib.sha256(raw.tobytes()).hexdigest(),
                                  probes_hex=probes, raw_file=(name + ".npz") if store else None)
            if store:
                np.savez_compressed(os.path.join(here, name + ".npz"), T=raw)
            print(name, manifest[name]["dump_md5"], lines[-2:])
    with open(os.path.join(here, "manifest.json"), "w") as f:
        json.dump(manifest, f, indent=1, sort_keys=True)


if __name__ == "__main__":
    main()
