#!/bin/bash
# tools/experiment_rowsplit.sh -- the row-split family against what NBODY_VARIANT_AUTO picks without it
# (NBODY_NO_ROWSPLIT=1: small-N family below 8 192 bodies, stream-K kernels above), at N = 148 x (targets per CTA)
# where each shape fills the SMs exactly, at the BASELINE size 16 384 and at sizes in between (where the rule
# must keep the other kernels).   gpurun --timeout 600 -- 'bash tools/experiment_rowsplit.sh'
mkdir -p gpurun_out/row
out=gpurun_out/row/shapes.jsonl
: > $out
run() { NBODY_NO_ROWSPLIT=1 timeout 120 python tools/sweep.py "$@" 2>&1 | sed 's/, "frac.*/}/' >> $out; }
run 16384 640 0 152
run 3552 640 0 165
run 4144 640 0 166
run 4096 640 0 166
run 4736 640 0 167
run 5328 640 0 168
run 5920 640 0 160
run 7104 640 0 161
run 8288 640 0 162
run 8192 640 0 162
run 9472 640 0 163
run 10656 640 0 164
run 11840 640 0 150
run 14208 640 0 151
run 18944 640 0 153
run 21312 320 0 154
run 23680 320 0 170
run 28416 320 0 171
run 33152 320 0 172
run 32768 320 0 172
run 37888 320 0 173
echo "--- what AUTO picks with the rule on" >> $out
for n in 3000 5000 6500 10240 12288 13000 15000 17500 20480 26000 30000 36000; do
    timeout 60 python tools/sweep.py $n 320 0 2>&1 | sed 's/, "frac.*/}/' >> $out
    NBODY_NO_ROWSPLIT=1 timeout 60 python tools/sweep.py $n 320 0 2>&1 | sed 's/, "frac.*/}/' >> $out
done
cat $out
