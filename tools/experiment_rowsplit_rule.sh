#!/bin/bash
# tools/experiment_rowsplit_rule.sh -- what NBODY_VARIANT_AUTO picks with the row-split rule on, against the same run
# with NBODY_NO_ROWSPLIT=1, over unsharded sizes between the shapes' sweet spots: the rule must never lose.
mkdir -p gpurun_out/row
out=gpurun_out/row/rule.txt
: > $out
for n in ${SIZES:-2500 3000 3552 4096 5000 5920 6500 7104 8192 9000 10240 11000 12288 13000 14208 15000 16384 17500 18944 20480 22000 23680 26000 28416 30000 32768 36000 37888 40000}; do
    a=$(timeout 60 python tools/sweep.py $n 640 0 2>&1 | python -c "import json,sys; d=json.loads(sys.stdin.readline()); print(d['name'].split(':')[0], d['ms_per_step']*1e3)")
    b=$(NBODY_NO_ROWSPLIT=1 timeout 60 python tools/sweep.py $n 640 0 2>&1 | python -c "import json,sys; d=json.loads(sys.stdin.readline()); print(d['name'].split(':')[0], d['ms_per_step']*1e3)")
    echo "$n | $a | $b" >> $out
done
cat $out
