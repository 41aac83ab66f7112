set -x
mkdir -p gpurun_out
# A/B of the direct-store epilogues (headline only)
for ds in 0 1 2 3; do timeout 120 python bench.py --only none --no-cpu-baseline --layers --opt direct_store=$ds > gpurun_out/r02b_ds$ds.json 2> gpurun_out/r02b_ds$ds.layers.txt; done
# latency probe
timeout 300 python tools/latency_probe.py > gpurun_out/r02b_latency_probe.json 2> gpurun_out/r02b_latency_probe.txt
# launch lists
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/r02b_launches.csv python bench.py --steps 2 --warmup 3 --only none --no-cpu-baseline > gpurun_out/r02b_ncu_bench.log 2>&1
timeout 200 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/r02b_b1_128_launches.csv python tools/prof_small.py 128 > /dev/null 2>&1
timeout 200 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/r02b_b1_512_launches.csv python tools/prof_small.py 512 > /dev/null 2>&1
# full captures
timeout 600 ncu --set full --clock-control none --import-source on --launch-skip 24 --launch-count 12 -o gpurun_out/r02b_fwd_full python tools/prof_forward.py > gpurun_out/r02b_ncu_fwd.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:'gaussian|noise|impulse|copy_u8|resample|normalize|resize' -o gpurun_out/r02b_membound_full python tools/prof_membound.py > gpurun_out/r02b_ncu_membound.log 2>&1
ls -la gpurun_out
