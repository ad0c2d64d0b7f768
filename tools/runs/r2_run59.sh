set -x
python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-secondary > /dev/null 2>&1 || exit 1
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r2_launches_v2.csv \
  python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-secondary > gpurun_out/r2_ncu1_v2.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"conv3x3_tc|uptail|gemm27|im2col27" -s 18 -c 6 -f \
  -o gpurun_out/r2_full_cnn_v2 python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-secondary > gpurun_out/r2_ncu2_v2.log 2>&1
ls -la gpurun_out/*.ncu-rep gpurun_out/r2_launches_v2.csv
