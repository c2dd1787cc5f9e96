#!/bin/bash
# last short call of the round: device parser (kernel rebuilt after parse_one was split out) + smoke, no torch import
timeout 40 python profiles/parse_check.py 2>&1 | tee gpurun_out/r2u_parse_smoke.log
