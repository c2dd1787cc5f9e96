"""b200cron: batched cron-schedule evaluation on B200 behind the reference's Schedule/ParseStandard shapes.

The compute path is the CUDA library csrc/libb200cron.so (C ABI in include/b200cron.h); this package is the thin
host-side mirror of the reference interface.  There is no CPU execution path: without the built library and a GPU
every compute call raises.
"""
