#!/bin/bash
# tools/experiment_small_below.sh -- where the stream-K kernels (with the spread fix-up) overtake the small-N family:
# us per step of the resident loop, AUTO kernel, small family forced (NBODY_SMALL_BELOW=65536) vs mid-size rule forced
# (NBODY_SMALL_BELOW=4096); best of three 3 200-step runs each (a short run right after an idle period can see the
# GPU's clocks still ramping up).   gpurun --timeout 900 -- 'bash tools/experiment_small_below.sh'
H=n-bodysimulation_b200/nbody_headless
mkdir -p gpurun_out
best() {   # best() ENVVALUE N -> "us checksum"
    for i in 1 2 3; do NBODY_SMALL_BELOW=$1 $H --n $2 --steps 3200 --ic uniform --resident --json | grep '^{'; done | python -c "
import json, sys
r = [json.loads(l) for l in sys.stdin]
print(min(x['device_ms'] / x['steps'] * 1e3 for x in r), r[0]['checksum_x'])"
}
for n in 4096 5120 5632 6144 6656 7168 7680 8192 8704 9216 9728 10240 10752 11264 11776 12288 13312; do
    a=($(best 65536 $n)); b=($(best 4096 $n))
    python -c "
a, b = float('${a[0]}'), float('${b[0]}')
print($n, 'small family', round(a, 2), 'us | stream-K', round(b, 2), 'us', '<-' if b < a else '')"
done | tee gpurun_out/small_below.txt
