# one whole frame (23 consecutive launches, warm) under ncu --set full; summaries under gpurun_out/
set -x
T=${TAG:-v6}
B="python bench.py --steps 1 --warmup 3 --no-cpu"
timeout 1500 ncu --set full --clock-control none --launch-skip 69 --launch-count 23 -o gpurun_out/r01_${T}_frame $B > gpurun_out/ncu_frame.log 2>&1
ncu -i gpurun_out/r01_${T}_frame.ncu-rep --page raw --csv > gpurun_out/r01_${T}_frame.raw.csv 2>/dev/null
ncu -i gpurun_out/r01_${T}_frame.ncu-rep --page details > gpurun_out/r01_${T}_frame.details.txt 2>/dev/null
ls -la gpurun_out | tail -5
