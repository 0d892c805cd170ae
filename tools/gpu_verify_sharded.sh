#!/bin/bash
# tools/gpu_verify_sharded.sh G -- tools/verify_sharded.py on G GPUs, both exchanges, one and four steps:
#   gpurun --gpus 8 --timeout 900 -- 'bash tools/gpu_verify_sharded.sh 8'      -> gpurun_out/verify<G>.jsonl
G=${1:-2}
out=gpurun_out/verify$G.jsonl
: > $out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $G --master-addr 127.0.0.1 --master-port 29519"
fail=0
for args in "--exchange p2p --barrier device" "--exchange nccl" "--exchange p2p --barrier device --steps 4" "--exchange nccl --overlap 0" "--exchange p2p --barrier device --bodies 1048576"; do
    echo "== verify_sharded $args"
    timeout 300 $TR tools/verify_sharded.py $args 2>gpurun_out/verify$G.err | grep '^{' | tee -a $out | cut -c1-700
    [ ${PIPESTATUS[0]} -ne 0 ] && { fail=1; tail -3 gpurun_out/verify$G.err; }
done
echo "== verify done fail=$fail"
