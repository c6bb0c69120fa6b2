"""Per-kernel SASS instruction counts of the tcgen05 translation unit (evidence that the decoder kernels are
Blackwell-native: UTCHMMA = tcgen05.mma kind::f16, LDTM = tcgen05.ld, UBLKCP = cp.async.bulk / TMA 1-D,
UTCBAR = tcgen05.commit).   python tools/sass_counts.py > profiles/rNN_sass_counts.txt   (no GPU needed)"""
import collections
import os
import re
import subprocess

ROOT = os.path.abspath(os.path.join(os.path.dirname(__file__), ".."))
OBJ = os.path.join(ROOT, "dsp-slam-rgbd_b200", "csrc", "kernels_tc.o")
KEEP = ("UTCHMMA", "LDTM", "UBLKCP", "UTCBAR", "SYNCS", "UTCATOMSWS", "STS", "LDS", "FFMA", "BAR", "ATOMG", "F2FP")

sass = subprocess.run(["cuobjdump", "-sass", OBJ], capture_output=True, text=True, check=True).stdout
cur, cnt = None, collections.defaultdict(collections.Counter)
for line in sass.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        cur = m.group(1)
        continue
    m = re.search(r"^\s+/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", line)
    if cur and m:
        op = m.group(1)
        base = op.split(".")[0]
        if base in KEEP:
            cnt[cur][base] += 1
            if base == "UBLKCP" and "MULTICAST" in op:
                cnt[cur]["UBLKCP.MULTICAST"] += 1
        cnt[cur]["_total"] += 1
print("# cuobjdump -sass dsp-slam-rgbd_b200/csrc/kernels_tc.o (sm_100a): instruction counts per kernel")
for k, v in cnt.items():
    name = subprocess.run(["c++filt", k], capture_output=True, text=True).stdout.strip()
    print("%-62s %s" % (name[:62], dict(sorted(v.items()))))
