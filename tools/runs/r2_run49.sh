set -x
timeout 2400 python -m pytest tests -m gpu -x -q 2>&1 | tail -5
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 10 --warmup 3 2>/dev/null | tail -1 | python -c "import sys,json; l=json.loads(sys.stdin.read()); print('N2 cnn', l['value'], l['e2e']['value'], l['n_gpus'])"
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 2 --workload rrdb 2>/dev/null | tail -1 | python -c "import sys,json; l=json.loads(sys.stdin.read()); print('N2 rrdb', l['value'], l['e2e']['value'], l['n_gpus'], l['clocks'])"
