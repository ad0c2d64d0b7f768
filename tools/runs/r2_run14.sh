set -x
mkdir -p gpurun_out
timeout 900 python bench.py --workload rrdb > gpurun_out/r2_bench_rrdb.json 2> gpurun_out/r2_bench_rrdb.err; tail -2 gpurun_out/r2_bench_rrdb.err; head -c 2500 gpurun_out/r2_bench_rrdb.json
timeout 300 python -m pytest tests/test_cnn_gpu.py -q -k "nvtx" 2>&1 | tail -2
python bench.py --workload wide --steps 1 --warmup 3 > /dev/null 2>&1 || exit 1
timeout 900 ncu --set full --clock-control none -k regex:"conv3x3_tc" -s 15 -c 5 -f -o gpurun_out/r2_full_wide python bench.py --workload wide --steps 1 --warmup 3 > gpurun_out/r2_ncu5.log 2>&1; ls -la gpurun_out/r2_full_wide.ncu-rep
