# One GPU session of the development loop: tests + the perf probes of every model. Usage: bash tools/gpu_session.sh <tag>
tag=${1:-r02x}
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/${tag}_pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/${tag}_pytest_gpu.log
timeout 200 python bench.py --only train --no-cpu-baseline --layers > gpurun_out/${tag}_bench_short.json 2> gpurun_out/${tag}_layers.txt; cat gpurun_out/${tag}_layers.txt | tail -13
python -c "
import json;d=json.loads(open('gpurun_out/${tag}_bench_short.json').read().strip().splitlines()[-1]);print('value',round(d['value']),d['ms_per_step'],'e2e',round(d['e2e']['value']),'train ms',d['train']['ms_per_step'])"
timeout 200 python tools/sr_perf.py esrgan 256 128 2>&1 | grep UNPROFILED
timeout 200 python tools/sr_perf.py srgan 64 128 2>&1 | tail -3
timeout 200 python tools/cgan_perf.py 256 2>&1 | tail -1
timeout 200 python tools/lat_debug.py 2>&1 | head -1
