#!/bin/bash
# hand-traced odd-transition answers + ordinary-DST scan through b200cron_next_n on the GPU (torch-free, seconds)
timeout 22 python profiles/quirk_check.py 2>&1 | tee gpurun_out/r2v_quirks.log
