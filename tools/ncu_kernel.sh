# full ncu capture of one launch of a kernel inside tools/stage_time.py (third frame), summaries under gpurun_out/
# usage: TAG=r02_v3 K=k_block_match_tma SKIP=2 [ENVV="MVS_BM_MODE=tma"] bash tools/ncu_kernel.sh
T=${TAG:-r02}; K=${K:-k_propagate}; S=${SKIP:-2}
B="python tools/stage_time.py 15view_1080p 1"
env $ENVV $B > gpurun_out/${T}_${K}_plain.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/${T}_${K}_plain.log; exit 1; }
env $ENVV timeout 600 ncu --set full --clock-control none --import-source on -k regex:$K -s $S -c 1 -o gpurun_out/${T}_$K $B > gpurun_out/${T}_ncu_$K.log 2>&1
ncu -i gpurun_out/${T}_$K.ncu-rep --page raw --csv > gpurun_out/${T}_$K.raw.csv 2>/dev/null
ncu -i gpurun_out/${T}_$K.ncu-rep --page source --csv > gpurun_out/${T}_$K.source.csv 2>/dev/null
ncu -i gpurun_out/${T}_$K.ncu-rep --page details > gpurun_out/${T}_$K.details.txt 2>/dev/null
rm -f gpurun_out/${T}_$K.ncu-rep
ls -la gpurun_out | grep ${T}_$K
