set -x
mkdir -p gpurun_out
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 10 --warmup 3 > gpurun_out/r2_bench_n2.json 2> gpurun_out/r2_bench_n2.err; echo rc=$?; tail -3 gpurun_out/r2_bench_n2.err; wc -l gpurun_out/r2_bench_n2.json
python - <<'PY'
import json
d=json.load(open('gpurun_out/r2_bench_n2.json'))
print(d['n_gpus'], d['value'], d['e2e']['value'], d['e2e_plugin']['value'], d['stream']['value'], d['stream']['order_check'], d['bicubic']['value'], list(d.keys()))
PY
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --impl reference --gpus 2 --steps 1 --warmup 0 > gpurun_out/r2_bench_n2_ref.json 2>> gpurun_out/r2_bench_n2.err; echo rc=$?; head -c 300 gpurun_out/r2_bench_n2_ref.json
python -m pytest tests/test_async_gpu.py -q -k current_device 2>&1 | tail -2
