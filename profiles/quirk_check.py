"""torch-free run of tests/test_gpu_parity.py::test_odd_transition_known_answers_and_ordinary_dst_scan on cuda:0
(a short gpurun call: pytest's collection hook would import torch, which takes a minute on a fresh box)"""
import sys
import time

sys.path.insert(0, ".")
t0 = time.time()
from cron_operator_b200 import api  # noqa: E402
from tests import test_gpu_parity as g  # noqa: E402

ctx = api.Context(devices=[0])
try:
    g.test_odd_transition_known_answers_and_ordinary_dst_scan(ctx, ctx.zone_names)
    print(f"odd-transition known answers + ordinary-DST scan ok on the GPU ({ctx.launch_count} launches, {time.time() - t0:.1f}s)")
finally:
    ctx.close()
