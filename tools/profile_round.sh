#!/bin/bash
# Round-end evidence on one B200 (run under gpurun): GPU tests, bench line, stage times, then -- each
# only after its plain run exited 0 -- the ncu launch list of the bench command and one `--set full`
# capture of the kernels of one step.  Outputs land in gpurun_out/; tools/ncu_summary.py turns the
# two ncu files into profiles/r1_bench_ncu.md and profiles/roofline_traffic.json.
set -u
O=gpurun_out
python -m pytest tests -m gpu -x -q > $O/final_tests.log 2>&1; tail -2 $O/final_tests.log
python bench.py > $O/final_bench.json 2> $O/final_bench.err || exit 1
python tools/stages.py headline c4 c3 c2 c1 > $O/final_stages.txt 2>&1 || exit 1
python tools/scene_bench.py > $O/final_scene_bench.json 2> $O/final_scene_bench.err
python bench.py --workload c5 --no-cpu-baseline > $O/final_bench_c5.json 2> $O/final_bench_c5.err
python bench.py --steps 2 --warmup 3 --no-cpu-baseline > /dev/null 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/final_launches.csv \
    python bench.py --steps 2 --warmup 3 --no-cpu-baseline > $O/final_ncu_launch.log 2>&1
ncu --set full --import-source on --clock-control none -k regex:k_ --launch-skip 36 -c 12 -o $O/final_step \
    python tools/stages.py headline > $O/final_ncu_full.log 2>&1
ls -la $O | tail -12
