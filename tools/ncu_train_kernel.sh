# ncu --set full capture of training-step kernels whose DEMANGLED name matches a regex (template arguments included).
# Usage: bash tools/ncu_train_kernel.sh <out-name> '<regex>' [launch-count]
out=${1:-train_kernel}; re=${2:-conv_umma_kernel}; cnt=${3:-3}
mkdir -p gpurun_out
timeout 600 ncu --set full --clock-control none --import-source on --kernel-name-base demangled -k "regex:$re" \
    --launch-skip 0 --launch-count "$cnt" -o "gpurun_out/$out" python tools/train_perf.py --steps 1 --warmup 0 \
    > "gpurun_out/${out}.log" 2>&1
echo "ncu rc=$?"
