for c in 2048 4096 8192 16384 32768 65536; do
  for mode in "" "DDPB200_ADAMW_PACKED=0" "DDPB200_ADAMW_WIDE=1"; do
    echo "== chunk $c $mode"
    env $mode python tools/kbench.py --only bf16 --sustained 200 --chunk $c 2>&1 | grep -E "adamw|clock"
  done
done
