"""Generates tests/golden/golden.json (+ tiny_*.npz) by running the COMPILED REFERENCE
(oracle/_ref/liblbref_*.so = unmodified LB.c, -O2 -ffp-contract=off) on the recipes of
tests/cases.py.  Run here, where /root/reference exists:

    sh oracle/build_ref.sh && python tests/golden/make_golden.py

Outputs per case: sha256 of fA and fB over (9,nx,ny) without padding, the Reynolds number as a hex
float, min/max of fA; for the `full` cases also the arrays themselves."""
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)

from oracle.harness import Ref  # noqa: E402
from tests import cases  # noqa: E402


def main():
    out = {}
    for name, c in cases.CASES.items():
        fA, walls = cases.make_inputs(name)
        r = Ref(c["prec"], c["nx"], c["ny"], c["isa"])
        r.set_state(fA, walls)
        r.run(c["steps"])
        a, b = r.f_view("A"), r.f_view("B")
        out[name] = dict(c, fA_sha256=cases.digest(a), fB_sha256=cases.digest(b),
                         reynolds_hex=float(r.reynolds()).hex(), reynolds=repr(r.reynolds()),
                         fA_min=float(a.min()), fA_max=float(a.max()),
                         walls_sha256=cases.digest(walls), input_sha256=cases.digest(fA))
        if c.get("full"):
            np.savez_compressed(os.path.join(HERE, name + ".npz"), fA=a, fB=b)
        print(name, out[name]["fA_sha256"][:16], out[name]["reynolds"], flush=True)
    with open(os.path.join(HERE, "golden.json"), "w") as fh:
        json.dump(out, fh, indent=1, sort_keys=True)


if __name__ == "__main__":
    main()
