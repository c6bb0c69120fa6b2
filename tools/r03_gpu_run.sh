#!/bin/bash
# Round-3 measurement pass on one B200 (run through gpurun from the repository root):
#   GPU tests, parity statistics of every mode (teacher-forced, free-running, held-out), the bench line, the reference
#   arm, the ncu launch list and one `ncu --set full` capture of the decoder kernels inside bench.py.
# Everything lands in gpurun_out/; the text summaries are copied to profiles/ by hand afterwards.
set -u
O=gpurun_out
mkdir -p $O
python -m pytest tests -x -q -m gpu > $O/r03_gputests.log 2>&1; echo "gpu tests rc=$?"; tail -2 $O/r03_gputests.log
python tools/parity_stats.py fp32 tc tc_exact tc_fast > $O/r03_parity_all_modes.jsonl 2> $O/r03_parity.err; echo "parity rc=$?"
python bench.py --steps 20 --warmup 3 > $O/r03_bench_1gpu.json 2> $O/r03_bench_1gpu.err; echo "bench rc=$?"
python bench.py --impl reference --steps 3 --warmup 1 > $O/r03_bench_reference_arm.json 2> $O/r03_bench_reference_arm.err; echo "reference arm rc=$?"
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file $O/r03_launches_bench.csv \
    python bench.py --steps 2 --warmup 1 --skip-configs --no-cpu > $O/r03_ncu_launches.log 2>&1; echo "launch list rc=$?"
ncu --set full --clock-control none --import-source on -k regex:k_tc --launch-skip 160 --launch-count 8 -f \
    -o $O/r03_prof_bench_k_tc python bench.py --steps 2 --warmup 1 --skip-configs --no-cpu > $O/r03_ncu_full.log 2>&1; echo "ncu full rc=$?"
tail -c 1500 $O/r03_bench_1gpu.json
