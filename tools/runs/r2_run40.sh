set -x
timeout 900 python -m pytest tests/test_rrdb_gpu.py -x -q 2>&1 | tail -3
timeout 900 python bench.py --workload rrdb --no-cpu-baseline 2>/dev/null | python -c "import sys,json; l=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('RRDB', l['value'], l['e2e']['value'], l['clocks'])"
python tools/rrdb_bench.py 1 > /dev/null 2>&1 || exit 1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"rconv3x3" -s 182 -c 8 -f \
  -o gpurun_out/r2_full_rrdb_v2 python tools/rrdb_bench.py 1 > gpurun_out/r2_ncu_rrdb_v2.log 2>&1
ls -la gpurun_out/*.ncu-rep
