"""Load tests/golden/lz4e_golden.json and rebuild the inputs of its cases."""
import hashlib
import json
import os

import numpy as np

import common
from common import sg

HERE = os.path.dirname(os.path.abspath(__file__))


def load():
    with open(os.path.join(HERE, "golden", "lz4e_golden.json")) as f:
        return json.load(f)


def seg_spec(kind, n, k=0):
    if kind == "single":
        return sg.segs_single(n)
    if kind == "pages":
        return sg.segs_pages(n)
    if kind == "odd":
        return sg.segs_odd(n, 16, k)
    if kind == "huge_bvec":
        return {"size": n, "segs": [(0, (1 << 24) + 1)], "skip": 0}
    raise ValueError(kind)


def sha(b):
    return hashlib.sha256(np.asarray(b, dtype=np.uint8).tobytes()).hexdigest()


def compress_inputs(cases):
    datas = [common.synth(c["n"], c["cls"], c["seed"]) for c in cases]
    specs = [seg_spec(c["seg"], c["n"], c.get("k", 0)) for c in cases]
    return datas, specs
