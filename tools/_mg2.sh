mkdir -p gpurun_out/r02_multi_gpu_parity
O=gpurun_out/r02_multi_gpu_parity
nvidia-smi -L > $O/gpus_n2.txt
python -m pytest tests -m gpu -q -x 2>&1 | tail -15 > $O/pytest_gpu_2gpus.log
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511"
$TR tools/fused_check.py > $O/fused_check_n2.log 2>&1; echo "rc=$?" >> $O/fused_check_n2.log
$TR bench.py --gpus 2 --steps 20 --warmup 3 > $O/bench_n2.json 2> $O/bench_n2.err; echo "rc=$?" >> $O/bench_n2.err
$TR bench.py --gpus 2 --steps 20 --warmup 3 --separate-grads --no-parity-check > $O/bench_n2_separate.json 2> $O/bench_n2_separate.err
$TR bench.py --gpus 2 --steps 20 --warmup 3 --zero1 > $O/bench_n2_zero1.json 2> $O/bench_n2_zero1.err; echo "rc=$?" >> $O/bench_n2_zero1.err
$TR tools/arbench.py > $O/arbench_n2.json 2>&1
tail -3 $O/pytest_gpu_2gpus.log $O/fused_check_n2.log $O/bench_n2.err $O/bench_n2_zero1.err
