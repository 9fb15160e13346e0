# full ncu capture of one propagate launch (iteration IT of the third frame), summaries under gpurun_out/
# usage: TAG=r02_v1 IT=2 bash tools/ncu_prop.sh      (run only after the same command exited 0 without ncu)
T=${TAG:-r02}
IT=${IT:-2}
B="python tools/stage_time.py 15view_1080p 1"
$B > gpurun_out/${T}_plain.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/${T}_plain.log; exit 1; }
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_propagate -s $((10 + IT)) -c 1 -o gpurun_out/${T}_k_propagate $B > gpurun_out/${T}_ncu_k_propagate.log 2>&1
ncu -i gpurun_out/${T}_k_propagate.ncu-rep --page raw --csv > gpurun_out/${T}_k_propagate.raw.csv 2>/dev/null
ncu -i gpurun_out/${T}_k_propagate.ncu-rep --page source --csv > gpurun_out/${T}_k_propagate.source.csv 2>/dev/null
ncu -i gpurun_out/${T}_k_propagate.ncu-rep --page details > gpurun_out/${T}_k_propagate.details.txt 2>/dev/null
rm -f gpurun_out/${T}_k_propagate.ncu-rep
ls -la gpurun_out | grep ${T}
