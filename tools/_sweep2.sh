for rep in 1 2 3; do
for c in 4096 8192 16384; do
    echo "== bf16 chunk $c rep $rep"
    python tools/kbench.py --only bf16 --sustained 400 --chunk $c 2>&1 | grep -E "adamw|clock"
done
done
for c in 4096 8192 16384 65536; do
    echo "== f32 chunk $c"
    python tools/kbench.py --only f32 --sustained 200 --chunk $c 2>&1 | grep -E "adamw|clock"
done
