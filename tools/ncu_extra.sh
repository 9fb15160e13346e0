# ncu --set full of the kernels outside the default frame; summaries under gpurun_out/
set -x
T=${TAG:-v12}
python tools/extra_kernels.py > gpurun_out/extra_plain.log 2>&1 || { tail -5 gpurun_out/extra_plain.log; exit 1; }
timeout 900 ncu --set full --clock-control none -k regex:'k_suppress|k_project_inv|k_remove_inconsistency|k_ingest_bgr|k_quantise' --launch-skip 7 --launch-count 7 -o gpurun_out/r01_${T}_extra python tools/extra_kernels.py > gpurun_out/ncu_extra.log 2>&1
ncu -i gpurun_out/r01_${T}_extra.ncu-rep --page raw --csv > gpurun_out/r01_${T}_extra.raw.csv 2>/dev/null
ncu -i gpurun_out/r01_${T}_extra.ncu-rep --page details > gpurun_out/r01_${T}_extra.details.txt 2>/dev/null
rm -f gpurun_out/r01_${T}_extra.ncu-rep
tail -3 gpurun_out/extra_plain.log
