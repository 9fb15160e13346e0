# full ncu capture of one propagate launch (SKIP-th launch of the kernel), summaries under gpurun_out/
set -x
B="python bench.py --steps 1 --warmup 3 --no-cpu"
T=${TAG:-v5}
timeout 400 ncu --set full --clock-control none --import-source on -k regex:k_propagate -s ${SKIP:-3} -c 1 -o gpurun_out/r01_${T}_k_propagate $B > gpurun_out/ncu_k_propagate.log 2>&1
ncu -i gpurun_out/r01_${T}_k_propagate.ncu-rep --page raw --csv > gpurun_out/r01_${T}_k_propagate.raw.csv 2>/dev/null
ncu -i gpurun_out/r01_${T}_k_propagate.ncu-rep --page source --csv > gpurun_out/r01_${T}_k_propagate.source.csv 2>/dev/null
ncu -i gpurun_out/r01_${T}_k_propagate.ncu-rep --page details > gpurun_out/r01_${T}_k_propagate.details.txt 2>/dev/null
