#!/bin/bash
# tools/gpu_regression.sh -- the whole 1-GPU regression in one gpurun call (about 2.5 GPU-minutes):
#
#   gpurun --timeout 600 -- 'bash tools/gpu_regression.sh'            # tests, smoke, both bench arms
#   gpurun --timeout 900 -- 'bash tools/gpu_regression.sh launches'   # ... plus the ncu launch list of the bench command
#
# Results land in gpurun_out/regression/ (copy what should be judged into profiles/).  Every step runs under its own
# timeout; ncu runs only after the same command has exited 0 without it, and no number printed under ncu is used.
set -u
out=gpurun_out/regression
mkdir -p "$out"
fail=0
# run "$@" with its output in $out/<name>.log; print the tail; remember a failure (no pipes: $? must be the command's)
step() {
    local name=$1; shift
    echo "== $name"
    "$@" >"$out/$name.log" 2>&1
    local rc=$?
    tail -3 "$out/$name.log"
    if [ $rc -ne 0 ]; then echo "   FAILED rc=$rc"; fail=1; fi
}

step pytest_gpu timeout 600 python -m pytest tests -x -q -m gpu
step smoke timeout 120 python -c "import __graft_entry__ as g; g.smoke()"
echo "== bench (ours)"
timeout 300 python bench.py >"$out/bench.out" 2>"$out/bench.err" || fail=1
grep '^{' "$out/bench.out" >"$out/bench.json" || fail=1
echo "== bench --impl reference"
timeout 300 python bench.py --impl reference >"$out/bench_ref.out" 2>"$out/bench_ref.err" || fail=1
grep '^{' "$out/bench_ref.out" >"$out/bench_ref.json" || fail=1
python - "$out" <<'PY'
import json, sys
o = sys.argv[1]
for f in ("bench.json", "bench_ref.json"):
    try:
        d = json.load(open(f"{o}/{f}"))
        print(f, d["value"], d["unit"], "e2e", (d.get("e2e") or {}).get("value"), "frac", (d.get("roofline") or {}).get("frac"),
              (d.get("engine") or {}).get("variant"), (d.get("clocks") or {}).get("reasons"))
    except Exception as e:  # noqa: BLE001
        print(f, "unreadable:", e)
PY
if [ "${1:-}" = launches ] && [ $fail -eq 0 ]; then
    echo "== ncu launch list of: bench.py --steps 2 --warmup 1 --no-cpu-baseline"
    timeout 200 python bench.py --steps 2 --warmup 1 --no-cpu-baseline >/dev/null 2>&1 &&
        timeout 400 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file "$out/launches.csv" \
            python bench.py --steps 2 --warmup 1 --no-cpu-baseline >"$out/ncu.log" 2>&1 || fail=1
    wc -l "$out/launches.csv"
fi
echo "== regression $([ $fail -eq 0 ] && echo ok || echo FAILED)"
exit $fail
