"""Recipe for `oracle/_ref/`: a verbatim, UNMODIFIED copy of the reference's own Python files for the hot path
(SURVEY.md 8(a)), taken from `/root/reference` where they lie, so that the reference itself can be timed and used
as the on-box checker on the GPU box (where `/root/reference` does not exist).

`oracle/_ref/` is git-ignored (reference sources never enter the history) but NOT gpurun-ignored: like the built
`.so` files it travels to the GPU box with the snapshot.  TEST INFRASTRUCTURE: only `bench.py --impl reference`,
`bench.py`'s `cpu_baseline` leg, `tools/ref_gpu.py` and tests import it, through `oracle/ref_harness.py`.

    python oracle/fetch_ref.py        # idempotent; called by __graft_entry__.build() when /root/reference exists
"""
import hashlib
import json
import os
import shutil

SRC = "/root/reference"
DST = os.path.join(os.path.dirname(os.path.abspath(__file__)), "_ref")
FILES = [
    "reconstruct/__init__.py", "reconstruct/optimizer.py", "reconstruct/loss.py", "reconstruct/loss_utils.py",
    "reconstruct/utils.py", "deep_sdf/__init__.py", "deep_sdf/deep_sdf_decoder.py", "deep_sdf/workspace.py",
    "configs/config_kitti.json", "configs/config_redwood_01053.json", "configs/config_freiburg_001.json",
]


def fetch():
    if not os.path.isdir(SRC):
        return False
    manifest = {}
    for rel in FILES:
        dst = os.path.join(DST, rel)
        os.makedirs(os.path.dirname(dst), exist_ok=True)
        shutil.copyfile(os.path.join(SRC, rel), dst)
        with open(dst, "rb") as f:
            manifest[rel] = hashlib.sha256(f.read()).hexdigest()
    with open(os.path.join(DST, "MANIFEST.json"), "w") as f:
        json.dump({"source": SRC, "sha256": manifest}, f, indent=1)
    return True


if __name__ == "__main__":
    print("oracle/_ref:", "fetched %d files" % len(FILES) if fetch() else "no /root/reference here, nothing done")
