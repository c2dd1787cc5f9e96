"""gpurun_out/ncu_pern_n{N}.csv -> profiles/r2_ncu_n{N}.md (achieved HBM GB/s of rank 0's kernels while all N ranks run)"""
import csv
import json
import sys

N = int(sys.argv[1])
peak = json.load(open("MEASURED_PEAKS.json"))["hbm_gbs"]
rows = list(csv.reader(open(f"gpurun_out/ncu_pern_n{N}.csv")))
hdr = None
launches = {}
for r in rows:
    if "Kernel Name" in r:
        hdr = r
        continue
    if hdr and len(r) == len(hdr):
        d = dict(zip(hdr, r))
        launches.setdefault((d["ID"], d["Kernel Name"]), {})[d["Metric Name"]] = (float(d["Metric Value"].replace(",", "")), d["Metric Unit"])
specs = 1_000_000_000 // N
out = [f"# r2: rank 0's kernels under ncu while all {N} rank(s) run their C5 shard ({specs} specs + 2e7 history segments per GPU)\n",
       "command: `bash profiles/ncu_rank0.sh N` (one plain process per GPU; rank 0 = `ncu --metrics "
       "dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --clock-control none`; per-launch times under ncu are "
       "serialised and cold: compare the GB/s, not the ms, with bench.py's)\n",
       f"HBM peak (MEASURED_PEAKS.json): {peak} GB/s\n",
       "| kernel | time | dram read | dram write | achieved DRAM GB/s | % of peak | algorithmic GB/s |", "|---|---|---|---|---|---|---|"]
scale = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
tscale = {"ns": 1e-9, "us": 1e-6, "usecond": 1e-6, "ms": 1e-3, "msecond": 1e-3, "nsecond": 1e-9, "second": 1}
for (lid, name), mm in launches.items():
    t = mm["gpu__time_duration.sum"][0] * tscale[mm["gpu__time_duration.sum"][1]]
    rd = mm["dram__bytes_read.sum"][0] * scale[mm["dram__bytes_read.sum"][1]]
    wr = mm["dram__bytes_write.sum"][0] * scale[mm["dram__bytes_write.sum"][1]]
    alg = 56 * specs if "next1" in name else None  # no seconds plane: 40 + 8 + 8 B per spec
    out.append(f"| {name[:48]} | {t * 1e3:.3f} ms | {rd / 1e9:.3f} GB | {wr / 1e9:.3f} GB | {(rd + wr) / t / 1e9:.0f} | "
               f"{(rd + wr) / t / 1e9 / peak * 100:.1f} | " + (f"{alg / t / 1e9:.0f} |" if alg else "- |"))
open(f"profiles/r2_ncu_n{N}.md", "w").write("\n".join(out) + "\n")
print("\n".join(out))
