#!/usr/bin/env python
"""tests/golden/fmindex_golden.json from the REFERENCE ITSELF: the .fmi file its fmindex_build writes (oracle/_ref:
ref_fmindex_build_file), padding bits of the last bwt_ byte / b_ word cleared (the reference leaves them uninitialised).

    python tests/golden/make_fmindex_golden.py"""
import hashlib
import json
import os
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(HERE)), "oracle"))
import fmindex_oracle as fo  # noqa: E402
import oracle_lib as ol  # noqa: E402
import test_fmindex as tf  # noqa: E402

out = {}
for name in tf.NAMES:
    S = tf.CASES[name]()
    img = fo.mask_padding(ol.ref_fmindex_image(S), S.size + 1)
    out[name] = {"n": int(S.size), "bytes": len(img), "sha256": hashlib.sha256(img).hexdigest()}
json.dump(out, open(os.path.join(HERE, "fmindex_golden.json"), "w"), indent=1)
print(len(out), "vectors")
