# One GPU session of the development loop: tests, A/B runs, perf probes. Usage: bash tools/gpu_session.sh <tag>
tag=${1:-r02x}
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/${tag}_pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/${tag}_pytest_gpu.log
timeout 200 python tools/sr_perf.py esrgan 256 128 > gpurun_out/${tag}_sr_perf_esrgan.txt 2>&1; tail -3 gpurun_out/${tag}_sr_perf_esrgan.txt
timeout 300 python bench.py --only membound --no-cpu-baseline > gpurun_out/${tag}_bench_membound.json 2>/dev/null; python -c "
import json;d=json.loads(open('gpurun_out/${tag}_bench_membound.json').read().strip().splitlines()[-1])
for k in d['membound']['kernels']: print(k['kernel'][:60], round(k['ms'],4), round(k['gbs']), round(k['frac_of_hbm_peak'],3))"
