"""DRAM traffic per launch of the decoder kernels, from a committed `ncu --set full` capture:
    python tools/ncu_traffic.py gpurun_out/r03_prof_bench_k_tc.ncu-rep > profiles/r03_traffic.json
The JSON records the hash of the kernel sources and the workload string of the tree it is made in (run it right after
the capture, on the tree that was captured): bench.py prints `roofline.traffic` from the newest such file only when
both still match, null otherwise.  The .ncu-rep itself stays out of git (20 MB); the text summaries are committed.
bench.py reads the JSON for `roofline.traffic` (dram__bytes_read.sum + dram__bytes_write.sum of the largest launch of
each kernel in the capture - the phase-0 forward launch, the Jacobian launch, the refinement launch)."""
import csv
import json
import os
import subprocess
import sys

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import bench  # noqa: E402  (kernel_src_sha16 / workload_desc: the identity of what was captured)

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
print(json.dumps({"source": rep, "kernel_src_sha16": bench.kernel_src_sha16(), "workload": bench.workload_desc(),
                  "kernels": best}, indent=1))
