# weight-ring depth sweep, one layer at a time (others at their defaults): prints the layer's ms per setting
for spec in "3 down2.2" "4 bottleneck.0" "5 bottleneck.2" "7 upconv2.0" "8 upconv2.2"; do
  set -- $spec
  for w in 3 4 6 8 10 12; do
    python bench.py --only none --no-cpu-baseline --layers --steps 20 --opt wst$1=$w 2> gpurun_out/ws.txt > gpurun_out/ws.json || { echo "layer $2 wst=$w FAILED"; tail -2 gpurun_out/ws.txt; continue; }
    echo "layer $2 wst=$w $(grep "$2 " gpurun_out/ws.txt)"
  done
done
