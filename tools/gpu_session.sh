# One GPU session of the development loop: tests, A/B runs, perf probes. Usage: bash tools/gpu_session.sh <tag>
tag=${1:-r02x}
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/${tag}_pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/${tag}_pytest_gpu.log
for o in 0 1; do timeout 120 python bench.py --only none --no-cpu-baseline --layers --opt up2_stream=$o > gpurun_out/${tag}_up2stream$o.json 2> gpurun_out/${tag}_up2stream$o.layers.txt; grep -E "up2 " gpurun_out/${tag}_up2stream$o.layers.txt; python -c "
import json;d=json.loads(open('gpurun_out/${tag}_up2stream$o.json').read().strip().splitlines()[-1]);print('up2_stream=$o',round(d['value']),d['ms_per_step'])"; done
timeout 200 python tools/sr_perf.py esrgan 256 128 > gpurun_out/${tag}_sr_perf_esrgan.txt 2>&1; cat gpurun_out/${tag}_sr_perf_esrgan.txt
timeout 300 python bench.py --only membound --no-cpu-baseline > gpurun_out/${tag}_bench_membound.json 2>/dev/null; python -c "
import json;d=json.loads(open('gpurun_out/${tag}_bench_membound.json').read().strip().splitlines()[-1])
for k in d['membound']['kernels']: print(k['kernel'][:60], round(k['ms'],4), round(k['gbs']), round(k['frac_of_hbm_peak'],3))"
