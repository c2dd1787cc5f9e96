"""Turns the .ncu-rep files brought back in gpurun_out/ into the markdown summaries committed under profiles/.
    python profiles/summarize.py <report.ncu-rep> <kernel substring> <title> <out.md> [command line that was profiled]"""
import csv
import subprocess
import sys

rep, kern, title, out = sys.argv[1:5]
cmd = sys.argv[5] if len(sys.argv) > 5 else ""
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout.splitlines()
rows = list(csv.reader(raw))
hdr, units = rows[0], rows[1]
data = [r for r in rows[2:] if kern in r[hdr.index("Kernel Name")]]
KEYS = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "launch__grid_size", "launch__block_size",
        "launch__registers_per_thread", "launch__waves_per_multiprocessor",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum",
        "smsp__thread_inst_executed_per_inst_executed.ratio", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "smsp__warps_eligible.avg.per_cycle_active", "sm__cycles_elapsed.avg", "smsp__cycles_active.avg",
        "smsp__average_warps_issue_stalled_no_instruction_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_wait_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio",
        "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active"]
with open(out, "w") as f:
    f.write(f"# {title}\n\n")
    if cmd:
        f.write(f"command (exited 0 without ncu first, same call): `{cmd}`\n\n")
    f.write("| metric | unit | " + " | ".join(f"launch {i + 1}" for i in range(len(data))) + " |\n|---|---|" + "---|" * len(data) + "\n")
    f.write("| kernel | | " + " | ".join(d[hdr.index("Kernel Name")][:60] for d in data) + " |\n")
    for k in KEYS:
        if k in hdr:
            i = hdr.index(k)
            f.write(f"| {k} | {units[i]} | " + " | ".join(d[i] for d in data) + " |\n")
print(open(out).read())
