set -x
mkdir -p gpurun_out
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"filter3x3_rows|bilateral5|lanczos4_2x" -s 3 -c 3 -f -o gpurun_out/r2_full_filters python tools/filters_bench.py > gpurun_out/r2_ncu6.log 2>&1; ls -la gpurun_out/r2_full_filters.ncu-rep
