# Full verification pass on one B200 (what the driver runs at round end, plus the ncu captures the bench cites).
# Usage: bash tools/gpu_verify.sh <tag>     -> gpurun_out/<tag>_*
tag=${1:-r02x}
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/${tag}_pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/${tag}_pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/${tag}_smoke.log 2>&1; echo "smoke rc=$?"; tail -3 gpurun_out/${tag}_smoke.log
timeout 600 python bench.py --impl reference > gpurun_out/${tag}_bench_reference.json 2>/dev/null; echo "reference rc=$?"
timeout 600 python bench.py --layers > gpurun_out/${tag}_bench_n1.json 2> gpurun_out/${tag}_layers.txt; echo "bench rc=$?"; cat gpurun_out/${tag}_layers.txt | tail -14
# ncu: launch list of the bench command, then full captures (only after the plain runs above exited)
timeout 400 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/${tag}_launches.csv python bench.py --steps 2 --warmup 3 --only none --no-cpu-baseline > gpurun_out/${tag}_ncu_bench.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on --launch-skip 24 --launch-count 12 -o gpurun_out/${tag}_fwd_full python tools/prof_forward.py > gpurun_out/${tag}_ncu_fwd.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:'gaussian|noise|impulse|copy_u8|resample|normalize|resize|salt_pepper' --launch-skip 10 -o gpurun_out/${tag}_membound_full python tools/prof_membound.py > gpurun_out/${tag}_ncu_membound.log 2>&1
ls -la gpurun_out | tail -20
