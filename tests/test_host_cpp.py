"""include/b200cron.hpp -- the C++ host mirror of the reference's schedule interface (ParseStandard, Schedule.Next,
getNextSchedule, the batched NextN / UnmetSince / HistoryPrune).  tests/host_cpp/mirror_test.cpp is built with g++
against libb200cron.so; the CPU test runs its host-only half, the GPU test compares what it prints with the oracle."""
import os
import subprocess

import numpy as np
import pytest

from cron_operator_b200 import api, synth
from oracle import pyoracle as po

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
SRC = os.path.join(HERE, "host_cpp", "mirror_test.cpp")
EXE = os.path.join(HERE, "host_cpp", "mirror_test")
NOW = synth.ts(2026, 1, 1, 12)


def _build():
    lib = api.lib_path()
    if not os.path.exists(lib):
        api.lib()
    deps = [SRC, os.path.join(ROOT, "include", "b200cron.hpp"), os.path.join(ROOT, "include", "b200cron.h"), lib]
    if not os.path.exists(EXE) or any(os.path.getmtime(d) > os.path.getmtime(EXE) for d in deps):
        libdir = os.path.dirname(lib)
        subprocess.check_call(["g++", "-std=c++17", "-O1", "-Wall", "-Werror", "-I" + os.path.join(ROOT, "include"), SRC,
                               "-L" + libdir, "-lb200cron", "-Wl,-rpath," + libdir, "-o", EXE])
    return EXE


def _run(mode):
    r = subprocess.run([_build(), mode], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout + r.stderr
    out = {}
    for line in r.stdout.splitlines():
        key, _, rest = line.partition(" ")
        out[key] = rest
    return out


def test_header_compiles_and_host_half_works():
    out = _run("cpu")
    for key, text in (("spec5", "*/5 * * * *"), ("spectz", "CRON_TZ=Asia/Tokyo 0 9 * * 1-5"), ("every", "@every 1h30m")):
        s = api.ParseStandard(text)
        f = out[key].split()
        assert [int(x) for x in f[:6]] == list(s.masks()) and f[6] == (s.Location or "-") and int(f[7]) == s.Every
    assert out["k1"].startswith("code -4 text ") and "5 fields" in out["k1"]
    import torch
    if not torch.cuda.is_available():
        assert "no usable CUDA device" in out["engine-error"]  # no CPU fallback behind the C++ mirror either


@pytest.mark.gpu
def test_reference_cases_and_batches_through_the_cpp_mirror():
    out = _run("gpu")
    assert out["k1"].startswith('unparsable cron "not a schedule"')       # K1, cron_controller.go:394
    assert out["k2"] == 'unschedulable cron "0 0 30 2 *"'                  # K2, :411
    assert out["k3"].split() == [str(NOW), str(NOW + 60)]                  # K3
    assert out["every"].split() == [str(NOW - 20), str(NOW + 90)]
    assert out["specset"] == "same"                                        # Engine::SpecSet == the host-pointer calls
    texts = ["*/5 * * * *", "TZ=America/New_York 30 2 * * *", "0 0 29 2 *", "15 3 * * 0"]
    specs = [api.ParseStandard(t) for t in texts]
    ctx = api.Context()
    try:
        names = ctx.zone_names
    finally:
        ctx.close()
    zid = np.array([names.index(s.Location) if s.Location else 0 for s in specs], dtype=np.int32)
    masks = api.pack_specs(specs)
    assert int(out["next"]) == int(po.batch_next_n(masks[:, :1].copy(), np.array([NOW]), 1)[0, 0])
    frm = np.array([NOW, synth.ts(2026, 3, 8, 2), NOW, NOW], dtype=np.int64)
    assert [int(x) for x in out["nextn"].split()] == po.batch_next_n(masks, frm, 3, zid, names).ravel().tolist()
    earliest = np.array([NOW - 3600, NOW - 5 * 86400, NOW - 86400, NOW + 10], dtype=np.int64)
    w = po.batch_unmet(masks, earliest, NOW, 0, np.array([0, 1, 2, 0], np.uint8), np.array([0, 2, 1, 0], np.uint32),
                       np.zeros(4, np.uint8), np.full(4, po.ZERO, np.int64), zid, names)
    want = []
    for i in range(4):
        want += [int(w["last"][i]), int(w["next"][i]), int(w["missed"][i]), int(w["status"][i]), int(w["decision"][i])]
    assert [int(x) for x in out["unmet"].split()] == want
    perm, keep = po.batch_history(np.array([0, 5, 5, 8], np.uint64), np.array([5, 3, 9, 3, 7, 2, 2, 1], np.int64),
                                  np.array([3, 1, -1], np.int64))
    assert out["history"].split() == [f"{int(p)}:{int(k)}" for p, k in zip(perm, keep)]
