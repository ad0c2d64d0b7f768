set -x
mkdir -p gpurun_out
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"filter3x3_rows" -s 2 -c 1 -f -o gpurun_out/r2_full_f3 python tools/filters_bench.py > gpurun_out/r2_ncu7.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"lanczos4_2x" -s 2 -c 1 -f -o gpurun_out/r2_full_lz python tools/filters_bench.py > gpurun_out/r2_ncu8.log 2>&1
ls -la gpurun_out/r2_full_f3.ncu-rep gpurun_out/r2_full_lz.ncu-rep
