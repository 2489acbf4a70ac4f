#!/usr/bin/env python
"""Generates tests/golden/ingest_golden.json from the REFERENCE ITSELF (oracle/_ref/libkiss_ref.so: ref_read_sequence =
the unmodified read_sequence of include/utils/io.hpp on an istringstream, then `% 4`).  Run in the build container:

    python tests/golden/make_ingest_golden.py
"""
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
import ingest_cases as ic  # noqa: E402
import oracle_lib as ol  # noqa: E402


def main():
    assert ol.have_ref(), "needs oracle/_ref/libkiss_ref.so (python __graft_entry__.py build where /root/reference exists)"
    out = {}
    cases = dict(ic.CASES)
    cases.update(ic.big_cases())
    for name, data in cases.items():
        codes = ol.ref_read_sequence(data)
        out[name] = {"bytes": len(data), "n": int(codes.size), "sha256": ol.sha_bytes(codes),
                     "head": codes[:32].tolist()}
    with open(os.path.join(HERE, "ingest_golden.json"), "w") as f:
        json.dump(out, f, indent=1)
    print(f"{len(out)} vectors")


if __name__ == "__main__":
    main()
