"""robfig/cron v3.0.1's own known-answer rows (tests/golden/upstream_robfig.json, [UPSTREAM-RECALLED] -- see
tests/golden/make_upstream_robfig.py for provenance) against (1) the CPU oracle, (2) the g++ host build of the
kernels' walk source and (3) the product's host parser.  The GPU leg is tests/test_gpu_parity.py::test_upstream_robfig_vectors.
These are the only vectors for zones / DST / the DOM-DOW rule whose expected values come from outside this repo."""
import json
import os

import numpy as np
import pytest

from cron_operator_b200 import api
from oracle import pyoracle as po
from tests import shim

U = json.load(open(os.path.join(os.path.dirname(__file__), "golden", "upstream_robfig.json")))
NAMES = shim.zone_names()
PARSERS = {"standard": False, "seconds": True}


def _masks(spec, with_seconds):
    """parsed by the oracle's parser and by the product's host parser; both must agree"""
    s, kind, _, tz = po.parse_standard(spec, with_seconds)
    p = api._parse(spec, with_seconds)
    assert s.tuple() == p.masks() and tz == p.Location and kind == 0 and p.Every == 0, spec
    return np.array([[int(x)] for x in s.tuple()], dtype=np.uint64), tz


def _zone(case_zone, tz):
    name = tz or case_zone
    return np.array([0 if name == "UTC" else NAMES.index(name)], np.int32)


def test_next_rows_oracle_and_host_walk():
    assert len(U["next"]) >= 39
    for c in U["next"]:
        m, tz = _masks(c["spec"], True)
        assert tz in ("", c["zone"])
        z = _zone(c["zone"], tz)
        want = po.ZERO if c["want"] is None else c["want"]
        ora = int(po.batch_next_n(m, np.array([c["from"]]), 1, z, NAMES)[0, 0])
        got = int(shim.next_n(m, np.array([c["from"]]), k=1, zone_id=z)[0, 0])
        assert ora == want, ("oracle", c["time"], c["spec"], ora, want)
        assert got == want, ("host walk", c["time"], c["spec"], got, want)


def test_activation_rows_oracle_and_host_walk():
    for c in U["activation"]:
        m, tz = _masks(c["spec"], False)
        z = np.array([0], np.int32)
        at = c["at"]
        ora = int(po.batch_next_n(m, np.array([at - 1]), 1, z, NAMES)[0, 0])
        got = int(shim.next_n(m, np.array([at - 1]), k=1, zone_id=z)[0, 0])
        assert (ora == at) == c["expected"], ("oracle", c["time"], c["spec"])
        assert (got == at) == c["expected"] and got == ora, ("host walk", c["time"], c["spec"])


def test_parser_rows():
    for c in U["parse_ok"]:
        ws = PARSERS[c["parser"]]
        want = tuple(int(x) for x in c["masks"])
        assert po.parse_standard(c["spec"], ws)[0].tuple() == want, c["spec"]
        assert api._parse(c["spec"], ws).masks() == want, c["spec"]
    for c in U["parse_every"]:
        ws = PARSERS[c["parser"]]
        _, kind, every, _ = po.parse_standard(c["spec"], ws)
        assert (kind, every) == (1, c["every_sec"])
        assert api._parse(c["spec"], ws).Every == c["every_sec"]
    for c in U["parse_err"]:
        ws = PARSERS[c["parser"]]
        with pytest.raises(po.ParseError, match=c["err"]):
            po.parse_standard(c["spec"], ws)
        with pytest.raises(api.ParseError, match=c["err"]):
            api._parse(c["spec"], ws)
