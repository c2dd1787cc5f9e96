"""No-GPU checks of the product library: it loads, exports every symbol include/b200cron.h declares, its host-only
parser agrees with the oracle's restatement of robfig's parser, and compute entry points fail loudly without a GPU."""
import ctypes as C
import os
import random
import re

import pytest

from cron_operator_b200 import api
from oracle import pyoracle as po

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_library_exports_every_declared_symbol():
    hdr = open(os.path.join(ROOT, "include", "b200cron.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    names = sorted(set(re.findall(r"\b(b200cron_[a-z_0-9]+)\s*\(", hdr)))
    assert len(names) >= 18
    L = C.CDLL(api.lib_path()) if os.path.exists(api.lib_path()) else api.lib()
    for n in names:
        assert hasattr(L, n), f"{n} declared in include/b200cron.h but not exported"


def test_bindings_and_docs_name_only_declared_symbols():
    """the cgo shim (go/b200cron, never compiled here), the C++ mirror and the documents may only name entry points
    that include/b200cron.h declares -- keeps the uncompiled Go source and INTEGRATION.md from drifting off the ABI"""
    import glob
    hdr = open(os.path.join(ROOT, "include", "b200cron.h")).read()
    declared = set(re.findall(r"\b(b200cron_[a-z_0-9]+)\s*\(", hdr)) | {"b200cron_ctx", "b200cron_spec", "b200cron_specset"}
    files = glob.glob(os.path.join(ROOT, "go", "b200cron", "*.go")) + [
        os.path.join(ROOT, f) for f in ("INTEGRATION.md", "README.md", "DESIGN.md", os.path.join("include", "b200cron.hpp"))]
    assert len(files) >= 6
    for f in files:
        used = {u for u in re.findall(r"\b(b200cron_[a-z0-9_]+)\b", open(f).read()) if not u.endswith("_")}
        assert not (used - declared), (f, sorted(used - declared))


def test_no_cpu_fallback_without_gpu():
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is visible")
    with pytest.raises(api.B200CronError) as e:
        api.Context()
    assert "no usable CUDA device" in str(e.value) or "CUDA" in str(e.value)


CORPUS = [
    "* * * * *", "? ? ? ? ?", "*/5 * * * *", "*/1 * * * *", "0 0 1 JAN mon", "0 0 1 jan-MAR sun-sat", "@hourly",
    "@daily", "@midnight", "@weekly", "@monthly", "@yearly", "@annually", "@every 1h30m", "@every 1.5h", "@every 10ms",
    "@every 90s", "@every -5s", "@every 1h1m1s1ms1us1ns", "@every 2562047h", "@every 2562048h", "@every 1.h", "@every .5m",
    "@every 5", "@every h", "@every 1d", "@every 0", "@every +3m", "@every 1\xc2\xb5s",
    "TZ=Asia/Tokyo 0 9 * * *", "CRON_TZ=America/New_York 0 9 * * *", "TZ=UTC * * * * *", "TZ= * * * * *",
    "TZ=Europe/Berlin   @daily", "TZ=UTC", "CRON_TZ=UTC", "TZ=Asia/Tokyo\t* * * * *",
    "  0   0  *  *  *  ", "1,,2 * * * *", ", * * * *", "+5 * * * *", "*-5 * * * *", "5/15 * * * *", "1-5/2 * * * *",
    "60 31 30 2 *", "0 0 30 2 *", "", " ", "* * * *", "* * * * * *", "* * * * 7", "* 24 * * *", "* * 0 * *",
    "* * 32 * *", "* * * 13 *", "* * * 0 *", "5-1 * * * *", "*/0 * * * *", "1-2-3 * * * *", "1/2/3 * * * *",
    "a * * * *", "*/-5 * * * *", "-5 * * * *", "@nope", "@every bogus", "@", "@every ", "99999999999999999999 * * * *",
    "92233720368547758080 * * * *", "9223372036854775807 * * * *", "* * * foo *", "* * * jan-feb *", "* * * * mon-fri",
    "* * * * fri-mon", "* * * JAN,feb,MaR *", "* * mon * *", "* * * * sunday", "0-59 0-23 1-31 1-12 0-6", "*/60 * * * *",
    "*/59 * * * *", "0/0 * * * *", "?/2 * * * *", "*/2/3 * * * *", "1-* * * * *", "*,5 * * * *", "5,* * * * *",
    "1- * * * *", "/5 * * * *", "5/ * * * *", "0x10 * * * *", "1_0 * * * *", "１ * * * *", "* * * * *\n", "\t*\t*\t*\t*\t*\t",
]


def _fuzz(n, seed):
    """grammar-based: mostly valid terms per field, with a sprinkle of junk"""
    rng = random.Random(seed)
    bounds = [(0, 59), (0, 59), (0, 23), (1, 31), (1, 12), (0, 6)]
    names = {4: ["jan", "FEB", "Mar", "dec"], 5: ["sun", "MON", "Sat"]}
    junk = ["x", "-", "/", "*/", "60", "99", "1-", "-1", "1/2/3", "1-2-3", "@", "+", "**", "?*", "0x1", ""]

    def term(f):
        lo, hi = bounds[f]
        r = rng.random()
        if r < 0.06:
            return rng.choice(junk)
        v = lambda: (rng.choice(names[f]) if f in names and rng.random() < 0.2 else str(rng.randint(lo - (rng.random() < 0.03), hi + (rng.random() < 0.03))))  # noqa: E731
        if r < 0.30:
            base = rng.choice(["*", "?"])
        elif r < 0.60:
            base = v()
        else:
            a, b = rng.randint(lo, hi), rng.randint(lo, hi)
            if rng.random() < 0.9 and a > b:
                a, b = b, a
            base = f"{a}-{b}"
        if rng.random() < 0.3:
            base += "/" + str(rng.choice([0, 1, 1, 2, 3, 5, 7, 15, 30, 61]))
        return base

    out = []
    for _ in range(n):
        nf = rng.choice([5, 5, 5, 5, 5, 5, 6, 6, 4, 7])
        first = 0 if nf >= 6 else 1
        fields = []
        for i in range(nf):
            f = min(first + i, 5)
            fields.append(",".join(term(f) for _ in range(rng.choice([1, 1, 1, 2, 3]))) or "*")
        sep = rng.choice([" ", " ", "  ", "\t"])
        prefix = rng.choice(["", "", "", "TZ=Europe/Paris ", "CRON_TZ=Asia/Kolkata ", "TZ=Bad/Zone "])
        out.append(prefix + sep.join(fields))
    return out


@pytest.mark.parametrize("with_seconds", [False, True])
def test_parser_matches_oracle(with_seconds):
    texts = CORPUS + _fuzz(4000, 3 + with_seconds)
    n_ok = 0
    for t in texts:
        b = t.encode("utf-8", errors="surrogateescape") if isinstance(t, str) else t
        try:
            want = po.parse_standard(b, with_seconds)
        except po.ParseError:
            want = None
        try:
            got = api._parse(b, with_seconds)
        except api.ParseError:
            got = None
        if want is None or got is None:
            assert want is None and got is None, (t, want, got)
            continue
        n_ok += 1
        spec, kind, every, tz = want
        assert got.masks() == spec.tuple(), t
        assert got.Every == (every if kind == 1 else 0) and got.Location == tz, t
    assert n_ok > 200


def test_reference_known_answers_parse_side():
    # K1 (cron_controller_test.go:170-186): "60 31 30 2 *" is unparsable
    with pytest.raises(api.ParseError):
        api.ParseStandard("60 31 30 2 *")
    s = api.ParseStandard("*/5 * * * *")  # SURVEY.md 3.4 observed masks
    assert s.masks() == (0x1, 0x0084210842108421, 0x8000000000FFFFFF, 0x80000000FFFFFFFE, 0x8000000000001FFE,
                         0x800000000000007F)


def test_K5_default_job_name_and_actions():
    # cron_util_test.go:137-149: name "cron-test", time Unix(1234567890) -> "cron-test-1234567890"
    assert api.getDefaultJobName("cron-test", 1234567890) == "cron-test-1234567890"
    import numpy as np
    now = 1_767_268_800
    res = dict(decision=np.array([api.DEC_CREATE, api.DEC_SKIP_FORBID, api.DEC_REPLACE_CREATE, api.DEC_GATE_SUSPENDED]),
               next=np.array([now + 60, now + 300, now + 5, api.ZERO]), missed=np.array([5, 1, 200, 0]))
    acts = api.actions_from(res, ["a", "b", "c", "d"], now)
    assert acts[0]["create"] == f"a-{now + 60}" and acts[0]["requeue_after"] == 60 and acts[0]["set_last_schedule_time"] == now
    assert acts[1]["create"] is None and acts[1]["requeue_after"] == 300
    assert acts[2]["delete_active"] and acts[2]["too_many_missed"] and acts[2]["create"] == f"c-{now + 5}"
    assert acts[3]["requeue_after"] is None and acts[3]["create"] is None
