#!/bin/bash
# Copies the outputs of tools/profile_round2.sh (gpurun_out/r2_*) into profiles/ in the committed form:
# JSON pretty-printed, the ncu report summarised by tools/ncu_summary.py.
set -eu
cd "$(dirname "$0")/.."
pretty() { python - "$1" "$2" <<'PY'
import json, sys
lines = [l for l in open(sys.argv[1]) if l.lstrip().startswith("{")]
try:
    d = json.loads(lines[-1])
except Exception:
    d = json.load(open(sys.argv[1]))
open(sys.argv[2], "w").write(json.dumps(d, indent=1) + "\n")
PY
}
pretty gpurun_out/r2_bench.json profiles/r2_bench_headline.json
pretty gpurun_out/r2_c5_step_breakdown.json profiles/r2_c5_step_breakdown.json
pretty gpurun_out/r2_h2d_bw_n1.json profiles/r2_h2d_bw_n1.json
cp gpurun_out/r2_stages.txt profiles/r2_stages.txt
cp gpurun_out/r2_launches.csv profiles/r2_bench_launches.csv
python tools/ncu_summary.py --rep gpurun_out/r2_step.ncu-rep --launches gpurun_out/r2_launches.csv \
    --out profiles/r2_bench_ncu.md --traffic profiles/roofline_traffic.json > /dev/null
