#!/bin/bash
# Round-2 evidence on one B200 (run under gpurun): GPU tests, bench line, stage times, then -- each only after its plain
# run exited 0 -- the ncu launch list of the bench command and one `--set full` capture of the kernels of one step.
# Outputs land in gpurun_out/r2_*; tools/ncu_summary.py turns the two ncu files into profiles/r2_bench_ncu.md and
# profiles/roofline_traffic.json.
set -u
O=gpurun_out
python -m pytest tests -m gpu -x -q > $O/r2_tests.log 2>&1; tail -2 $O/r2_tests.log
python bench.py > $O/r2_bench.json 2> $O/r2_bench.err || { tail -5 $O/r2_bench.err; exit 1; }
python tools/stages.py headline c4 c3 c2 c1 > $O/r2_stages.txt 2>&1 || exit 1
python tools/c5_step_breakdown.py > $O/r2_c5_step_breakdown.json 2> $O/r2_c5_step_breakdown.err
python tools/h2d_bw.py > $O/r2_h2d_bw_n1.json 2> $O/r2_h2d_bw_n1.err
python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-extras > /dev/null 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/r2_launches.csv \
    python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-extras > $O/r2_ncu_launch.log 2>&1
ncu --set full --import-source on --clock-control none -k regex:k_ --launch-skip 39 -c 13 -o $O/r2_step \
    python tools/stages.py headline > $O/r2_ncu_full.log 2>&1
ls -la $O | grep r2_
