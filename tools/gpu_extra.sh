#!/bin/bash
# tools/gpu_extra.sh G -- the bench lines the north star asks for beside the default one, on G GPUs of one box:
# Plummer initial conditions and the NCCL all-gather exchange (the default line is uniform + fused P2P exchange).
#   gpurun --gpus 4 --timeout 900 -- 'bash tools/gpu_extra.sh 4 default'   # also the default line (G = 4 has no other record)
#   gpurun --gpus 8 --timeout 900 -- 'bash tools/gpu_extra.sh 8'
# Results: gpurun_out/extra<G>/{default,plummer,nccl}.json
set -u
G=${1:-2}
out=gpurun_out/extra$G
mkdir -p "$out"
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $G --master-addr 127.0.0.1 --master-port 29517"
[ "$G" = 1 ] && TR=python
run() {
    local name=$1; shift
    echo "== $name: bench.py --gpus $G $*"
    timeout 600 $TR bench.py --gpus $G --steps ${STEPS:-10} --warmup 3 "$@" >"$out/$name.out" 2>"$out/$name.err" || { echo "   FAILED"; tail -5 "$out/$name.err"; }
    grep '^{' "$out/$name.out" >"$out/$name.json"
    python - "$out/$name.json" <<'PY'
import json, sys
d = json.load(open(sys.argv[1]))
p = d.get("parity") or {}
print("  value", d["value"], "ms/step", d["ms_per_step"], "e2e", (d.get("e2e") or {}).get("value"), "frac", d["roofline"]["frac"],
      "parity", p.get("ok"), "err_vs_fp64", p.get("err_a_vs_fp64"), "timeouts", p.get("barrier_timeouts"), "ic", d["config"].get("ic"), "exchange", d["config"].get("exchange"))
PY
}
[ "${2:-}" = default ] && run default
run plummer --ic plummer --no-extras --no-capi --no-cpu-baseline
[ "$G" -gt 1 ] && run nccl --exchange nccl --no-extras --no-capi --no-cpu-baseline --no-series
echo "== extra done"
