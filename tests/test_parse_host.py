"""The device parser's source on the CPU: parse_kernels.cuh parse_one (what every thread of parse_batch_kernel runs),
built with g++ into the test shim, against the oracle's restatement of robfig/cron v3.0.1 parser.go
(cronv3.ParseStandard, called at internal/controller/cron_controller.go:392).  The GPU suite runs the kernel itself on
6,000 fuzzed specs per field count; this runs the same source over far more text on a box without a GPU."""
import random

import pytest

from oracle import pyoracle as po
from tests import shim
from tests.test_abi_cpu import CORPUS, _fuzz

EVERY_TAG = 1 << 62


def _every_fuzz(n, seed):
    """'@every <duration>' strings: plain h/m/s terms (parsed on the device) and everything time.ParseDuration also takes
    or refuses (left to the host: status 2), plus descriptors and near-descriptors"""
    rng = random.Random(seed)
    units = ["ns", "us", "µs", "ms", "s", "m", "h", "d", "", "x", "S", "hh"]
    nums = ["1", "5", "90", "1.5", ".5", "0", "007", "1e3", "999999999", "1000000000", "9223372036", "99999999999999999999",
            "-1", "+2", "1.", " 3", ""]
    desc = ["@yearly", "@annually", "@monthly", "@weekly", "@daily", "@midnight", "@hourly", "@reboot", "@Daily", "@ daily",
            "@daily ", " @daily", "@everyday", "@every", "@every ", "@every  1s", "@every\t1s", "@", "@@daily"]
    out = []
    for _ in range(n):
        r = rng.random()
        if r < 0.75:
            body = "".join(rng.choice(nums) + rng.choice(units) for _ in range(rng.choice([1, 1, 2, 3, 4])))
            out.append("@every " + rng.choice(["", "", "", "-", "+"]) + body)
        else:
            out.append(rng.choice(["", "", "TZ=Asia/Tokyo ", "CRON_TZ=Local ", "TZ= "]) + rng.choice(desc))
    return out


def _check(texts, with_seconds):
    blobs = [t.encode("utf-8", errors="surrogateescape") if isinstance(t, str) else t for t in texts]
    masks, status, tz_off, tz_len = shim.parse_batch(blobs, with_seconds)
    seen = {0: 0, 1: 0, 2: 0}
    for i, b in enumerate(blobs):
        try:
            spec, kind, every, tz = po.parse_standard(b, with_seconds)
        except po.ParseError:
            spec = None
        got = tuple(int(x) for x in masks[:, i])
        st = int(status[i])
        seen[st] += 1
        if st == 2:  # a duration the device leaves to the host parser: whatever the verdict, it is an "@every"
            assert b"@every " in b and (spec is None or kind == 1), b
            assert got == (0,) * 6, b
        elif st == 1:
            assert spec is None, (b, spec and spec.tuple())
            assert got == (0,) * 6, b
        else:
            assert spec is not None, (b, got)
            assert got == spec.tuple(), (b, got, spec.tuple())
            if kind == 1:  # ConstantDelaySchedule in the planes
                assert got[4] == EVERY_TAG and got[3] == every, b
            if tz_off[i] == 0 and tz_len[i] == 0:
                name = ""
            elif tz_off[i] == 0xFFFFFFFF:
                name = "UTC"  # "TZ= ...": LoadLocation("") is UTC
            else:
                name = b[int(tz_off[i]):int(tz_off[i]) + int(tz_len[i])].decode(errors="replace")
            assert name == tz, (b, name, tz)
    return seen


@pytest.mark.parametrize("with_seconds", [False, True])
def test_device_parser_source_matches_oracle_on_corpus_and_fuzz(with_seconds):
    seen = _check(CORPUS + _fuzz(40000, 1234 + with_seconds), with_seconds)
    assert seen[0] > 2000 and seen[1] > 2000


def test_device_parser_source_on_every_durations_and_descriptors():
    texts = _every_fuzz(30000, 77)
    seen = _check(texts, False)
    _check(texts[:5000], True)
    assert seen[0] > 1000 and seen[2] > 1000  # both the device-parsed and the handed-over kind occur


def test_device_parser_simple_every_is_never_handed_over():
    """whole h/m/s terms of at most nine digits are the device's job (status 0, EVERY_TAG encoding), not status 2"""
    texts = ["@every 1s", "@every 90m", "@every 1h30m", "@every 2h0m5s", "@every 0s", "@every 999999999s", "@every 1m1m"]
    masks, status, _, _ = shim.parse_batch(texts)
    assert status.tolist() == [0] * len(texts)
    assert [int(x) for x in masks[3]] == [1, 5400, 5400, 7205, 1, 999999999, 120]
    assert all(int(x) == EVERY_TAG for x in masks[4])
