#!/bin/bash
# tools/gpu_multi.sh -- the multi-GPU regression in one gpurun call:
#
#   gpurun --gpus 2 --timeout 900 -- 'bash tools/gpu_multi.sh 2'          # tests that need two GPUs + bench --gpus 2
#   gpurun --gpus 8 --timeout 1200 -- 'bash tools/gpu_multi.sh 8 bench'   # bench only (both arms), as the driver launches it
#
# Results land in gpurun_out/multi<G>/ (copy what should be judged into profiles/).
set -u
G=${1:-2}
what=${2:-all}
out=gpurun_out/multi$G
mkdir -p "$out"
fail=0
step() {
    local name=$1; shift
    echo "== $name"
    "$@" >"$out/$name.log" 2>&1
    local rc=$?
    tail -4 "$out/$name.log"
    if [ $rc -ne 0 ]; then echo "   FAILED rc=$rc"; fail=1; fi
}
nvidia-smi --query-gpu=index,name --format=csv,noheader | head -$G
if [ "$what" = all ]; then
    step pytest_multi timeout 900 python -m pytest tests/test_round2_gpu.py tests/test_headless_gpu.py tests/test_next_rows_gpu.py -q -m gpu -rs
fi
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $G --master-addr 127.0.0.1 --master-port 29511"
echo "== bench --impl reference --gpus $G"
timeout 600 $TR bench.py --impl reference --gpus $G --steps ${STEPS:-10} --warmup 3 >"$out/bench_ref.out" 2>"$out/bench_ref.err" || fail=1
grep '^{' "$out/bench_ref.out" >"$out/bench_ref.json" || fail=1
echo "== bench --gpus $G"
timeout 850 $TR bench.py --gpus $G --steps ${STEPS:-10} --warmup 3 ${BENCH_ARGS:-} >"$out/bench.out" 2>"$out/bench.err" || { fail=1; tail -20 "$out/bench.err"; }
grep '^{' "$out/bench.out" >"$out/bench.json" || fail=1
python - "$out" <<'PY'
import json, sys
o = sys.argv[1]
try:
    d = json.load(open(f"{o}/bench.json"))
    print("value", d["value"], "ms/step", d["ms_per_step"], "e2e", (d.get("e2e") or {}).get("value"), "per-GPU frac", d["roofline"]["frac"])
    print("per_rank", json.dumps(d["engine"]["per_rank"]))
    print("parity", json.dumps(d.get("parity")))
    for s in d.get("series") or []:
        print("series", json.dumps({k: v for k, v in s.items() if k not in ("per_rank", "reference")}))
    print("capi", json.dumps(d.get("capi")))
    print("ref_gpu_baseline", json.dumps(d.get("ref_gpu_baseline")))
    r = json.load(open(f"{o}/bench_ref.json"))
    print("reference arm", r["value"], r["cpu_baseline"]["cores"], "threads")
except Exception as e:  # noqa: BLE001
    print("unreadable:", e)
PY
echo "== multi-GPU regression $([ $fail -eq 0 ] && echo ok || echo FAILED)"
exit $fail
