#!/bin/bash
# final check at HEAD: smoke + the whole GPU suite
python __graft_entry__.py smoke 2>&1 | tail -1
( time timeout 1200 python -m pytest tests -m gpu -q ) > gpurun_out/pytest_r2r.log 2>&1; tail -4 gpurun_out/pytest_r2r.log
