# round 2, first GPU call: new boundary tests, TMEM ubench, smoke, bench, the reference's own CUDA path
set -x
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.max.sm,power.limit --format=csv
./tools/ubench/tmem_ld > gpurun_out/r2_tmem_ld.txt 2>&1; cat gpurun_out/r2_tmem_ld.txt
timeout 1500 python -m pytest tests -m gpu -x -q -k "not all_300" > gpurun_out/r2_run1_pytest.log 2>&1; tail -15 gpurun_out/r2_run1_pytest.log
timeout 300 python __graft_entry__.py --smoke > gpurun_out/r2_run1_smoke.log 2>&1; tail -3 gpurun_out/r2_run1_smoke.log
timeout 900 python bench.py --steps 20 > gpurun_out/r2_run1_bench.json 2> gpurun_out/r2_run1_bench.err; tail -5 gpurun_out/r2_run1_bench.err; head -c 3000 gpurun_out/r2_run1_bench.json
timeout 600 python tools/ref_cuda_path.py > gpurun_out/r2_ref_cuda_path.json 2> gpurun_out/r2_ref_cuda_path.err; cat gpurun_out/r2_ref_cuda_path.json; tail -3 gpurun_out/r2_ref_cuda_path.err
