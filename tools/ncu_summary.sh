#!/bin/bash
# .ncu-rep (from tools/profile_round.sh, in gpurun_out/) -> csv summary with the columns bench.py and the judge read.
# usage: tools/ncu_summary.sh gpurun_out/r2_full_cnn.ncu-rep profiles/r2_ncu_full_cnn.csv
M="dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum,l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum,l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed,launch__registers_per_thread,lts__t_sector_hit_rate.pct,sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active,sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active,sm__pipe_tensor_subpipe_hmma_cycles_active.avg.pct_of_peak_sustained_active,sm__throughput.avg.pct_of_peak_sustained_elapsed,smsp__issue_active.avg.pct_of_peak_sustained_active,sm__warps_active.avg.pct_of_peak_sustained_active"
ncu -i "$1" --page raw --csv --metrics "$M" > "$2"
head -c 600 "$2"; echo; wc -l "$2"
