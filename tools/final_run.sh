# round-end evidence on ONE B200: tests, smoke, bench lines (ours, reference arm, secondary workloads), launch list,
# whole-frame ncu table.   TAG=r02_v6 bash tools/final_run.sh   (NO_NCU=1 stops before the two ncu passes)
T=${TAG:-r02}
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/${T}_gpu_tests.log 2>&1; tail -3 gpurun_out/${T}_gpu_tests.log
python __graft_entry__.py smoke > gpurun_out/${T}_smoke.log 2>&1; tail -1 gpurun_out/${T}_smoke.log
python bench.py --steps 20 --warmup 5 > gpurun_out/${T}_bench.json 2> gpurun_out/${T}_bench.err
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/${T}_bench_ref.json 2> gpurun_out/${T}_bench_ref.err
for w in bundled_15view bundled_15view_legacy beer_garden_3x3 4x4_1080p 5x5_4k; do
  python bench.py --workload $w --steps 10 --warmup 3 --no-cpu --no-extra > gpurun_out/${T}_bench_$w.json 2> gpurun_out/${T}_bench_$w.err
done
python bench.py --workload slic_4k --steps 10 --warmup 3 > gpurun_out/${T}_bench_slic4k.json 2> gpurun_out/${T}_bench_slic4k.err
if [ -n "$NO_NCU" ]; then ls -la gpurun_out | grep ${T} | head -30; exit 0; fi
B="python tools/stage_time.py 15view_1080p 1"
$B > gpurun_out/${T}_plain.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -s 60 -c 30 --csv --log-file gpurun_out/${T}_launches.csv $B > gpurun_out/${T}_ncu_list.log 2>&1
timeout 1500 ncu --set full --clock-control none --launch-skip 60 --launch-count 30 -o gpurun_out/${T}_frame $B > gpurun_out/${T}_ncu_frame.log 2>&1
ncu -i gpurun_out/${T}_frame.ncu-rep --page raw --csv > gpurun_out/${T}_frame.raw.csv 2>/dev/null
ncu -i gpurun_out/${T}_frame.ncu-rep --page details > gpurun_out/${T}_frame.details.txt 2>/dev/null
rm -f gpurun_out/${T}_frame.ncu-rep
python tools/ncu_table.py gpurun_out/${T}_frame.raw.csv > gpurun_out/${T}_frame_summary.md
ls -la gpurun_out | grep ${T} | head -30
