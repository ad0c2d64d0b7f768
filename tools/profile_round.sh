# Round-end evidence on one GPU (run through gpurun): tests, bench line, launch list, ncu --set full of one CNN step and of
# the bicubic kernel.  Each ncu pass runs only after the same command has exited 0 without ncu.  $1 = tag (e.g. r1_v6)
set -x
T=$1
python -m pytest tests -m gpu -x -q > gpurun_out/${T}_pytest.log 2>&1; tail -2 gpurun_out/${T}_pytest.log
python __graft_entry__.py --smoke > gpurun_out/${T}_smoke.log 2>&1; tail -2 gpurun_out/${T}_smoke.log
python bench.py > gpurun_out/${T}_bench.json 2> gpurun_out/${T}_bench.err || exit 1
python bench.py --steps 2 --warmup 3 --no-cpu-baseline > /dev/null 2>&1 || exit 1
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${T}_launches.csv \
  python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/${T}_ncu1.log 2>&1
python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-secondary > /dev/null 2>&1 || exit 1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"conv3x3_tc|uptail|gemm27|im2col27" -s 18 -c 6 -f \
  -o gpurun_out/${T}_full_cnn python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-secondary > gpurun_out/${T}_ncu2.log 2>&1
timeout 500 ncu --set full --clock-control none --import-source on -k regex:bicubic2x_strip -s 3 -c 1 -f \
  -o gpurun_out/${T}_full_bic python bench.py --workload bicubic --steps 1 --warmup 3 --no-cpu-baseline > gpurun_out/${T}_ncu3.log 2>&1
ls -la gpurun_out | tail -12
