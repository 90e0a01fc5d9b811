#!/usr/bin/env python3
"""Generate tests/golden/*.npz from oracle/_ref (the reference's own DSP headers compiled
in place from /root/reference against oracle/shim).  Run in the container that has the
reference:  python tests/golden/make_golden.py
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))

import golden_cases  # noqa: E402
import oracle_lib as ol  # noqa: E402


def main():
    ref = ol.Oracle("ref")
    assert ref.kind == "reference"
    total = 0
    for name in golden_cases.CASES:
        res = golden_cases.run_case(ref, name)
        path = os.path.join(HERE, name + ".npz")
        np.savez_compressed(path, **res)
        sz = os.path.getsize(path)
        total += sz
        print("%-28s %8d bytes  %s" % (name, sz, {k: (v.shape, str(v.dtype)) for k, v in res.items() if k in ("out", "counts", "mono")}))
    print("total", total)


if __name__ == "__main__":
    main()
