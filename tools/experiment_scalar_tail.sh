#!/bin/bash
# tools/experiment_scalar_tail.sh -- the one bounded kernel experiment the round-1 verdict asked for: does a SCALAR FP32 op
# next to each MUFU.RSQ hide the 0.17 FMA-pipe cycles a MUFU issue costs?  Variant 140 of the sweep build (`make sweep`)
# forms the two m * r^-3 products as scalar FMULs; variant 102 is the same kernel with the packed FMUL2.
#   gpurun --timeout 600 -- 'bash tools/experiment_scalar_tail.sh'   -> gpurun_out/scalar_tail.txt
export NBODY_LIB=libnbody_b200_sweep.so
out=gpurun_out/scalar_tail.txt
mkdir -p gpurun_out
python tools/sweep.py 262144 6 102 140 > $out 2>&1
M=sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active,sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active,sm__issue_active.avg.pct_of_peak_sustained_elapsed,gpu__time_duration.sum
for v in 102 140; do
    python tools/one.py 262144 3 $v >> $out 2>&1 &&
        ncu --metrics $M --clock-control none -k regex:force_kernel -s 1 -c 1 python tools/one.py 262144 3 $v 2>&1 | grep -E "force_kernel|pct|duration" >> $out
done
cat $out
