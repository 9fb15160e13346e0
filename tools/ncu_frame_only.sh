# launch list + whole-frame ncu table only (the bench lines come from tools/final_run.sh):  TAG=r02_v9 bash tools/ncu_frame_only.sh
T=${TAG:-r02}
mkdir -p gpurun_out
B="python tools/stage_time.py 15view_1080p 1"
$B > gpurun_out/${T}_plain.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/${T}_plain.log; exit 1; }
ncu --metrics gpu__time_duration.sum --clock-control none -s 60 -c 30 --csv --log-file gpurun_out/${T}_launches.csv $B > gpurun_out/${T}_ncu_list.log 2>&1
timeout 1500 ncu --set full --clock-control none --launch-skip 60 --launch-count 30 -o gpurun_out/${T}_frame $B > gpurun_out/${T}_ncu_frame.log 2>&1
ncu -i gpurun_out/${T}_frame.ncu-rep --page raw --csv > gpurun_out/${T}_frame.raw.csv 2>/dev/null
ncu -i gpurun_out/${T}_frame.ncu-rep --page details > gpurun_out/${T}_frame.details.txt 2>/dev/null
rm -f gpurun_out/${T}_frame.ncu-rep
python tools/ncu_table.py gpurun_out/${T}_frame.raw.csv > gpurun_out/${T}_frame_summary.md
head -8 gpurun_out/${T}_frame_summary.md | cut -c1-200
