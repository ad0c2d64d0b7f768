set -x
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29521 bench.py --gpus 8 --steps 10 --warmup 3 2>/dev/null | tail -1 > gpurun_out/r2_bench_n8_v2.json
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29522 bench.py --gpus 8 --workload rrdb 2>/dev/null | tail -1 > gpurun_out/r2_bench_rrdb_n8_v2.json
python - <<'PY'
import json
for f in ('gpurun_out/r2_bench_n8_v2.json','gpurun_out/r2_bench_rrdb_n8_v2.json'):
    l=json.loads(open(f).read()); print(f, l['value'], l['e2e']['value'], l['n_gpus'], l['clocks'])
PY
