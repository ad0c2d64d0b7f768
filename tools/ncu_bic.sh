# ncu --set full capture of the bicubic strip kernel (one launch) -> gpurun_out/$1.ncu-rep
timeout 500 ncu --set full --clock-control none --import-source on -k regex:bicubic2x_strip -s 3 -c 1 -f -o gpurun_out/$1 python bench.py --workload bicubic --steps 1 --warmup 1 --no-cpu-baseline > gpurun_out/ncu_bic.log 2>&1; echo rc=$?
