set -x
python tools/kbench.py --only bf16 --reps 3 --sustained 5 > gpurun_out/r02_ncu_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:adamw_kernel -s 3 -c 2 -o gpurun_out/r02_adamw_packed python tools/kbench.py --only bf16 --reps 3 --sustained 5 > gpurun_out/r02_ncu_packed.log 2>&1
DDPB200_ADAMW_PACKED=0 ncu --set full --clock-control none --import-source on -k regex:adamw_kernel -s 3 -c 2 -o gpurun_out/r02_adamw_scalar python tools/kbench.py --only bf16 --reps 3 --sustained 5 > gpurun_out/r02_ncu_scalar.log 2>&1
ls -la gpurun_out/
