# Round-end evidence on one GPU (run through gpurun): tests, smoke, bench line, launch list, ncu --set full of one CNN step,
# of the bicubic kernel and of one Real-ESRGAN block.  Each ncu pass runs only after the same command has exited 0 without
# ncu.  $1 = tag (e.g. r2).  The .ncu-rep files come back in gpurun_out/; tools/ncu_summary.sh turns them into the csv
# summaries under profiles/ that bench.py reads `roofline.traffic` from.
set -x
T=$1
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/${T}_pytest.log 2>&1; tail -2 gpurun_out/${T}_pytest.log
python __graft_entry__.py --smoke > gpurun_out/${T}_smoke.log 2>&1; tail -2 gpurun_out/${T}_smoke.log
python bench.py --steps 20 > gpurun_out/${T}_bench.json 2> gpurun_out/${T}_bench.err || exit 1
python bench.py --impl reference --steps 1 > gpurun_out/${T}_bench_reference.json 2>> gpurun_out/${T}_bench.err
python bench.py --workload wide --steps 3 > gpurun_out/${T}_bench_wide.json 2>> gpurun_out/${T}_bench.err
python tools/rrdb_bench.py > gpurun_out/${T}_rrdb_bench.json 2>> gpurun_out/${T}_bench.err
python tools/ref_cuda_path.py > gpurun_out/${T}_ref_cuda_path.json 2>> gpurun_out/${T}_bench.err
python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-secondary > /dev/null 2>&1 || exit 1
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${T}_launches.csv \
  python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-secondary > gpurun_out/${T}_ncu1.log 2>&1
python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-secondary > /dev/null 2>&1 || exit 1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"conv3x3_tc|uptail|gemm27|im2col27" -s 18 -c 6 -f \
  -o gpurun_out/${T}_full_cnn python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-secondary > gpurun_out/${T}_ncu2.log 2>&1
timeout 500 ncu --set full --clock-control none --import-source on -k regex:bicubic2x_strip -s 3 -c 1 -f \
  -o gpurun_out/${T}_full_bic python bench.py --workload bicubic --steps 1 --warmup 3 --no-cpu-baseline > gpurun_out/${T}_ncu3.log 2>&1
python tools/rrdb_bench.py 1 > /dev/null 2>&1 || exit 1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"rconv3x3" -s 60 -c 6 -f \
  -o gpurun_out/${T}_full_rrdb python tools/rrdb_bench.py 1 > gpurun_out/${T}_ncu4.log 2>&1
ls -la gpurun_out | tail -16
