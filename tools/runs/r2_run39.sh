set -x
timeout 2400 python -m pytest tests -m gpu -x -q 2>&1 | tail -6
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -2
timeout 900 python bench.py --workload rrdb > gpurun_out/r2_bench_rrdb_v2.json 2> gpurun_out/r2_bench_rrdb_v2.err; tail -c 600 gpurun_out/r2_bench_rrdb_v2.json
timeout 900 python bench.py > gpurun_out/r2_bench_v2.json 2> gpurun_out/r2_bench_v2.err; head -c 400 gpurun_out/r2_bench_v2.json
