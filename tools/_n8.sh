O=gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
timeout 120 $TR --nproc-per-node 8 --master-port 29513 tools/peer_allreduce_check.py > $O/r2_peer_allreduce_n8.json 2> $O/r2_peer_n8.err; echo peer8 rc=$?
timeout 120 $TR --nproc-per-node 4 --master-port 29514 tools/peer_allreduce_check.py > $O/r2_peer_allreduce_n4.json 2> $O/r2_peer_n4.err; echo peer4 rc=$?
timeout 120 $TR --nproc-per-node 4 --master-port 29515 tools/h2d_bw.py > $O/r2_h2d_bw_n4.json 2> $O/r2_h2d_n4.err; echo h2d4 rc=$?
timeout 120 $TR --nproc-per-node 8 --master-port 29516 tools/h2d_bw.py > $O/r2_h2d_bw_n8.json 2> $O/r2_h2d_n8.err; echo h2d8 rc=$?
timeout 300 $TR --nproc-per-node 8 --master-port 29517 bench.py --gpus 8 --steps 20 --warmup 5 > $O/r2_bench_n8.json 2> $O/r2_bench_n8.err; echo bench8 rc=$?
timeout 300 $TR --nproc-per-node 4 --master-port 29518 bench.py --gpus 4 --steps 20 --warmup 5 > $O/r2_bench_n4.json 2> $O/r2_bench_n4.err; echo bench4 rc=$?
grep -h "^{" $O/r2_peer_allreduce_n8.json $O/r2_peer_allreduce_n4.json $O/r2_h2d_bw_n4.json $O/r2_h2d_bw_n8.json | cut -c1-900
for f in $O/r2_bench_n8.json $O/r2_bench_n4.json; do python - "$f" <<'PY'
import json, sys
line=[l for l in open(sys.argv[1]).read().split("\n") if l.startswith("{")][-1]
d=json.loads(line)
print(sys.argv[1], d["n_gpus"], d["value"], d["ms_per_step"], "e2e", d["e2e"]["ms_per_step"])
print(json.dumps(d["extra"]["c5"]))
print({k: v.get("ms_per_step") for k, v in d["extra"].items() if isinstance(v, dict)})
PY
done
grep -v "^\*\*\*\|OMP_NUM" $O/r2_bench_n8.err | tail -5 | cut -c1-300
