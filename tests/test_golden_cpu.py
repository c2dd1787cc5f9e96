"""Committed golden vectors (tests/golden/vectors.json) vs the oracle and vs the kernels' walk source (host build)."""
import json
import os

import numpy as np

from oracle import pyoracle as po
from tests import shim

V = json.load(open(os.path.join(os.path.dirname(__file__), "golden", "vectors.json")))
NAMES = shim.zone_names()


def _m(case):
    return np.array([[int(x)] for x in case["masks"]], dtype=np.uint64)


def test_next_vectors():
    for c in V["next"]:
        z = 0 if c["zone"] == "UTC" else NAMES.index(c["zone"])
        for f, want in zip(c["from"], c["want"]):
            got = shim.next_n(_m(c), np.array([f]), k=4, zone_id=np.array([z], np.int32))[0].tolist()
            ora = po.batch_next_n(_m(c), np.array([f]), k=4, zone_id=np.array([z], np.int32), zone_names=NAMES)[0].tolist()
            assert got == want == ora, (c["spec"], c["zone"], f)


def test_unmet_vectors():
    for c in V["unmet"]:
        z = 0 if c["zone"] == "UTC" else NAMES.index(c["zone"])
        dl = np.array([po.ZERO if c["deadline"] is None else c["deadline"]])
        got = shim.unmet(_m(c), np.array([c["earliest"]]), c["now"], 0, [c["policy"]], [c["active"]], [c["flags"]], dl,
                         zone_id=np.array([z], np.int32))
        for k, v in c["want"].items():
            assert int(got[k][0]) == v, (c["spec"], k)


def test_history_vectors():
    for c in V["history"]:
        perm, keep = po.history_prune(c["keys"], c["limit"])
        assert perm.tolist() == c["perm"] and keep.tolist() == c["keep"]
