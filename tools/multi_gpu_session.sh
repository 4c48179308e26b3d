#!/bin/bash
# One multi-GPU measurement session: correctness of the fused collective, the collective microbench,
# bench.py with both collectives, and the whole-train-step comparison. Usage: multi_gpu_session.sh NP [tag]
NP=${1:-2}; TAG=${2:-n$NP}; OUT=gpurun_out; mkdir -p $OUT
RUN="timeout 420 python -m torch.distributed.run --nnodes=1 --nproc-per-node $NP --master-addr 127.0.0.1"
echo "== fused_check NP=$NP"; $RUN --master-port 29601 tools/fused_check.py 2>&1 | grep -E "FUSED_CHECK_OK|Error|assert|ranks differ" | head -5
for c in 1 2; do
  echo "== arbench NP=$NP ctas=$c"; DDPB200_AR_CTAS=$c $RUN --master-port 29602 tools/arbench.py 2>&1 | grep "^{" | tee $OUT/arbench_${TAG}_c$c.json | cut -c1-600
done
echo "== bench nccl";  $RUN --master-port 29603 bench.py --gpus $NP --steps 10 --warmup 3 --allreduce nccl  2>&1 | tail -1 > $OUT/bench_${TAG}_nccl.json;  cut -c1-200 $OUT/bench_${TAG}_nccl.json
echo "== bench fused"; $RUN --master-port 29604 bench.py --gpus $NP --steps 10 --warmup 3 --allreduce fused 2>&1 | tail -1 > $OUT/bench_${TAG}_fused.json; cut -c1-200 $OUT/bench_${TAG}_fused.json
if [ "${TRAIN:-1}" = "1" ]; then
  echo "== train_bench"; $RUN --master-port 29605 tools/train_bench.py --model llama-1b --batch 4 --seq 2048 --steps 6 --warmup 3 2>&1 | grep "^{" | tee $OUT/train_${TAG}.json | cut -c1-400
fi
