set -x
mkdir -p gpurun_out
python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29521 bench.py --gpus 8 --steps 10 --warmup 3 > gpurun_out/r2_bench_n8.json 2> gpurun_out/r2_bench_n8.err; echo rc=$?; tail -3 gpurun_out/r2_bench_n8.err; wc -l gpurun_out/r2_bench_n8.json
python - <<'PY'
import json
d=json.load(open('gpurun_out/r2_bench_n8.json'))
print(d['n_gpus'], d['value'], d['e2e']['value'], d['e2e']['streaming']['value'], d['e2e_plugin']['value'], d['stream']['value'], d['stream']['order_check'], d['bicubic']['value'])
PY
python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29522 bench.py --gpus 8 --workload rrdb > gpurun_out/r2_bench_rrdb_n8.json 2>> gpurun_out/r2_bench_n8.err; echo rc=$?; head -c 400 gpurun_out/r2_bench_rrdb_n8.json
