#!/bin/bash
# tools/experiment_small_below.sh -- where the stream-K kernels (with the spread fix-up) overtake the small-N family:
# us per step of the resident loop, AUTO kernel, small family forced (NBODY_SMALL_BELOW=65536) vs mid-size rule forced
# (NBODY_SMALL_BELOW=4096).   gpurun --timeout 600 -- 'bash tools/experiment_small_below.sh'
H=n-bodysimulation_b200/nbody_headless
mkdir -p gpurun_out
for n in 4096 4608 5120 5632 6144 6656 7168 7680 8192 8704 9216 9728 10240 10752 11264 11776 12288 13312; do
    a=$(NBODY_SMALL_BELOW=65536 $H --n $n --steps 640 --ic uniform --resident --json | grep '^{')
    b=$(NBODY_SMALL_BELOW=4096 $H --n $n --steps 640 --ic uniform --resident --json | grep '^{')
    python - "$n" "$a" "$b" <<'PY'
import json, sys
n, a, b = sys.argv[1], json.loads(sys.argv[2]), json.loads(sys.argv[3])
ua, ub = a["device_ms"] / a["steps"] * 1e3, b["device_ms"] / b["steps"] * 1e3
print(n, "small family", round(ua, 2), "us | stream-K", round(ub, 2), "us", "<-" if ub < ua else "", "same bits" if a["checksum_x"] == b["checksum_x"] else "")
PY
done | tee gpurun_out/small_below.txt
