#!/bin/bash
# tools/experiment_spread.sh -- the spread fix-up of split tiles (force_kernel's SPREAD, cooperative launch) against the
# last-arriver fix-up (NBODY_NO_SPREAD_FIXUP=1) at mid-size N: same bits, time per step.  Run on the GPU box:
#   gpurun --timeout 600 -- 'bash tools/experiment_spread.sh'
set -u
out=gpurun_out/spread
mkdir -p "$out"
H=n-bodysimulation_b200/nbody_headless
echo "== bits: spread vs last arriver (resident loop and the reference-shaped loop)"
for n in 12288 16384 20000 33333 65000; do
    for fl in "--resident" ""; do
        a=$($H --n $n --steps 3 --ic plummer --seed 4 --dt 0.01 $fl --json | grep '^{')
        b=$(NBODY_NO_SPREAD_FIXUP=1 $H --n $n --steps 3 --ic plummer --seed 4 --dt 0.01 $fl --json | grep '^{')
        python - "$n" "$fl" "$a" "$b" <<'PY'
import json, sys
n, fl, a, b = sys.argv[1], sys.argv[2], json.loads(sys.argv[3]), json.loads(sys.argv[4])
same = a["checksum_x"] == b["checksum_x"] and a["checksum_v"] == b["checksum_v"]
print(n, fl or "simulate()", "identical" if same else "DIFFERENT", a["checksum_x"], b["checksum_x"])
PY
    done
done 2>&1 | tee "$out/bits.txt"
echo "== us per step, resident loop (graph replay), AUTO kernel: spread | last arriver"
for n in 12288 16384 20480 24576 32768 49152 65536; do
    a=$($H --n $n --steps 600 --ic uniform --resident --json | grep '^{')
    b=$(NBODY_NO_SPREAD_FIXUP=1 $H --n $n --steps 600 --ic uniform --resident --json | grep '^{')
    python - "$n" "$a" "$b" <<'PY'
import json, sys
n, a, b = sys.argv[1], json.loads(sys.argv[2]), json.loads(sys.argv[3])
print(n, round(a["device_ms"] / a["steps"] * 1e3, 2), "|", round(b["device_ms"] / b["steps"] * 1e3, 2), "us   G/s", a["g_interactions_per_s"], "|", b["g_interactions_per_s"])
PY
done 2>&1 | tee "$out/time.txt"
echo "== N = 1M: the large-N family (its kernels are separate instantiations, untouched)"
python tools/sweep.py 1048576 3 90 97 98 101 102 105 2>&1 | tee "$out/sweep_1m.jsonl"
