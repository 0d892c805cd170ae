#!/bin/bash
# tools/gpu_soak.sh G [N] [STEPS] -- a long resident run through the headless driver on 1 and on G GPUs (fused P2P
# exchange with the in-kernel flag barrier, and the NCCL exchange): energy drift, barrier health (the driver exits 1 on
# a barrier time-out), state agreement between the three.   gpurun --gpus 8 --timeout 900 -- 'bash tools/gpu_soak.sh 8'
G=${1:-2}; N=${2:-262144}; S=${3:-1000}
H=n-bodysimulation_b200/nbody_headless
out=gpurun_out/soak$G.txt
: > $out
common="--n $N --steps $S --ic plummer --seed 11 --dt 0.01 --resident --energy --json"
for mode in "" "--gpus $G --exchange p2p" "--gpus $G --exchange nccl"; do
    echo "== nbody_headless $common $mode" | tee -a $out
    timeout 600 $H $common $mode > gpurun_out/soak.tmp 2> gpurun_out/soak.err; rc=$?
    grep '^{' gpurun_out/soak.tmp | tee -a $out | cut -c1-400
    [ $rc -ne 0 ] && { echo "   FAILED rc=$rc" | tee -a $out; tail -3 gpurun_out/soak.err | tee -a $out; }
done
python - $out <<'PY' | tee -a $out
import json, sys
runs = [json.loads(l) for l in open(sys.argv[1]) if l.startswith('{"n"')]
en = [json.loads(l) for l in open(sys.argv[1]) if l.startswith('{"energy"')]
n = runs[0]["n"]
for r, e in zip(runs, en):
    print("gpus", r["gpus"], r["exchange"], "G/s", r["g_interactions_per_s"], "rel_drift", e["rel_drift"],
          "checksum_x vs 1 GPU", abs(r["checksum_x"] - runs[0]["checksum_x"]) / (3 * n), "checksum_v", abs(r["checksum_v"] - runs[0]["checksum_v"]) / (3 * n))
PY
