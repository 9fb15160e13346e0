set -x
B="python bench.py --steps 1 --warmup 3 --no-cpu"
$B > gpurun_out/plain2.log 2>&1 || exit 1
for k in ${KERNELS:-k_propagate k_block_match k_update_centers k_assoc}; do
  timeout 400 ncu --set full --clock-control none --import-source on -k regex:$k -s 1 -c 1 -o gpurun_out/r01_${TAG:-v1}_$k $B > gpurun_out/ncu_$k.log 2>&1
  ncu -i gpurun_out/r01_${TAG:-v1}_$k.ncu-rep --page raw --csv > gpurun_out/r01_${TAG:-v1}_$k.raw.csv 2>/dev/null
  ncu -i gpurun_out/r01_${TAG:-v1}_$k.ncu-rep --page source --csv > gpurun_out/r01_${TAG:-v1}_$k.source.csv 2>/dev/null
  ncu -i gpurun_out/r01_${TAG:-v1}_$k.ncu-rep --page details > gpurun_out/r01_${TAG:-v1}_$k.details.txt 2>/dev/null
done
ls -la gpurun_out
