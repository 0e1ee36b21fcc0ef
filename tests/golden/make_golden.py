"""Generate the golden fixtures under tests/golden/ IN THE BUILD CONTAINER (needs /root/reference).

Two kinds of fixture, both produced from the reference itself (never from our code):

1. ``plink2twobit_lut.npy``  -- the 256 constants of d_conversion_table, parsed from
   /root/reference/src/cuda/snp_multiply_cuda.h:202 (the reference's own known-answer table).
2. ``dgemm_*.npz``           -- inputs (seeded, oracle.simulate) and the OUTPUTS of the unmodified
   reference CPU 5codes library (oracle/_ref/ref_miraculix.so, built by oracle/Makefile) for
   dgemm_compressed 'N' and 'T', centred and not, incl. missing codes, external f, and the
   deterministic B patterns of tests/dgemm_compressed/test_5codesapi{,_t}.f90:252-263.

Run:  python tests/golden/make_golden.py      (each centring mode runs in its own process
because the reference's options are process-global).
"""
from __future__ import annotations

import os
import re
import subprocess
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)

CASES = {
    # name: (snps, indiv, seed, missing, ncol, B kind)
    "dgemm_odd": (1003, 203, 0, 0.01, 5, "randn"),
    "dgemm_fortran_n": (640, 321, 1, 0.0, 10, "fortran10"),
    "dgemm_fortran_t": (517, 260, 2, 0.0, 15, "fortran1000"),
    "dgemm_tiny": (7, 5, 3, 0.0, 1, "randn"),
}


def make_lut() -> None:
    src = open("/root/reference/src/cuda/snp_multiply_cuda.h").read()
    m = re.search(r"d_conversion_table\[256\]\s*=\s*\{([^}]*)\}", src)
    vals = [int(v, 16) for v in m.group(1).split(",")]
    assert len(vals) == 256
    np.save(os.path.join(HERE, "plink2twobit_lut.npy"), np.array(vals, dtype=np.uint8))


def run_mode(centered: int) -> None:
    from oracle import ref, simulate
    ref.set_options(cores=4, centered=bool(centered), variant=256)
    for name, (snps, indiv, seed, missing, ncol, kind) in CASES.items():
        plink, Z, f = simulate.simulate(snps, indiv, seed=seed, missing=missing)
        rng = np.random.default_rng(seed + 100)
        fext = np.clip(f * 0.9 + 0.013, 0.0, 1.0)  # external f != data frequency
        h = ref.RefCompressed(plink, snps, indiv, fext, ncol)
        out = {"plink": plink, "f": fext, "snps": snps, "indiv": indiv}
        for tr, k in (("N", snps), ("T", indiv)):
            if kind == "randn":
                B = np.asfortranarray(rng.standard_normal((k, ncol)))
            else:
                B = simulate.fortran_pattern_B(k, ncol, 10 if kind == "fortran10" else 1000)
            out[f"B_{tr}"] = B
            out[f"C_{tr}"] = h.dgemm(tr, B)
        np.savez_compressed(os.path.join(HERE, f"{name}_c{centered}.npz"), **out)


if __name__ == "__main__":
    if len(sys.argv) > 1:
        run_mode(int(sys.argv[1]))
    else:
        subprocess.run(["make", "-s", "-j8", "-C", os.path.join(ROOT, "oracle"), "all"], check=True)
        make_lut()
        for c in (0, 1):
            subprocess.run([sys.executable, os.path.abspath(__file__), str(c)], check=True)
        print(sorted(os.listdir(HERE)))
