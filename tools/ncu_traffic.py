"""DRAM traffic per launch of the decoder kernels, from a committed `ncu --set full` capture:
    python tools/ncu_traffic.py profiles/r02_prof_bench_k_tc.ncu-rep > profiles/r02_traffic.json
bench.py reads the JSON for `roofline.traffic` (dram__bytes_read.sum + dram__bytes_write.sum of the largest launch of
each kernel in the capture - the phase-0 forward launch, the Jacobian launch, the refinement launch)."""
import csv
import json
import subprocess
import sys

rep = sys.argv[1]
out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv", "--metrics",
                      "gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum"],
                     capture_output=True, text=True, check=True).stdout
rows = list(csv.reader(out.splitlines()))
h, u = rows[0], rows[1]
ix = {c: i for i, c in enumerate(h)}
scale = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
best = {}
for r in rows[2:]:
    name = r[ix["Kernel Name"]]
    rd = float(r[ix["dram__bytes_read.sum"]]) * scale[u[ix["dram__bytes_read.sum"]]]
    wr = float(r[ix["dram__bytes_write.sum"]]) * scale[u[ix["dram__bytes_write.sum"]]]
    ms = float(r[ix["gpu__time_duration.sum"]])
    if name not in best or ms > best[name]["ms_under_ncu"]:
        best[name] = {"dram_read_bytes": int(rd), "dram_write_bytes": int(wr), "traffic_bytes": int(rd + wr),
                      "ms_under_ncu": ms}
print(json.dumps({"source": rep, "kernels": best}, indent=1))
