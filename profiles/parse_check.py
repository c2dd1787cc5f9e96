"""torch-free GPU check of the device parser (same comparison as tests/test_gpu_parity.py
test_device_parse_matches_host_parser) followed by smoke(); used for a last short gpurun call."""
import sys
import time

sys.path.insert(0, ".")
t0 = time.time()
from cron_operator_b200 import api  # noqa: E402
from tests.test_abi_cpu import CORPUS, _fuzz  # noqa: E402

ctx = api.Context(devices=[0])
tot = 0
for with_seconds in (False, True):
    texts = CORPUS + _fuzz(6000, 11 + with_seconds)
    texts = [t.encode("utf-8", errors="surrogateescape") if isinstance(t, str) else t for t in texts]
    masks, status, tz = ctx.parse_batch(texts, with_seconds)
    for i, t in enumerate(texts):
        try:
            want = api._parse(t, with_seconds)
        except api.ParseError:
            want = None
        if status[i] == 2:
            assert b"@every " in t and (want is None or want.Every), t
        elif want is None:
            assert status[i] == 1 and not masks[:, i].any(), (t, status[i])
        else:
            assert status[i] == 0 and tuple(int(x) for x in masks[:, i]) == want.masks(), t
            if not want.Every:
                assert (tz[i] or "") == want.Location, (t, tz[i], want.Location)
        tot += 1
ctx.close()
print(f"device parser ok on {tot} strings ({time.time() - t0:.1f}s)", flush=True)
import __graft_entry__ as g  # noqa: E402
g.smoke()
print(f"total {time.time() - t0:.1f}s")
