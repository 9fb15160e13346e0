# round-end evidence run on one B200: tests, smoke, bench (ours + reference arm), launch list, ncu captures
set -x
T=${TAG:-v10}
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu_$T.log 2>&1; tail -3 gpurun_out/pytest_gpu_$T.log
python __graft_entry__.py smoke > gpurun_out/smoke_$T.log 2>&1; tail -2 gpurun_out/smoke_$T.log
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_ref_$T.json 2> gpurun_out/bench_ref_$T.err
python bench.py --steps 10 --warmup 3 > gpurun_out/bench_$T.json 2> gpurun_out/bench_$T.err
python bench.py --workload slic_4k --steps 10 --warmup 3 > gpurun_out/bench_slic4k_$T.json 2> gpurun_out/bench_slic4k_$T.err
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r01_launches_$T.csv python bench.py --steps 2 --warmup 3 --no-cpu > gpurun_out/ncu_list.log 2>&1
TAG=$T bash tools/ncu_frame.sh > gpurun_out/capture_frame_$T.log 2>&1
B="python bench.py --steps 1 --warmup 3 --no-cpu"
for k in k_propagate k_block_match; do
  timeout 400 ncu --set full --clock-control none --import-source on -k regex:$k -s 2 -c 1 -o gpurun_out/r01_${T}_$k $B > gpurun_out/ncu_$k.log 2>&1
  ncu -i gpurun_out/r01_${T}_$k.ncu-rep --page source --csv > gpurun_out/r01_${T}_$k.source.csv 2>/dev/null
  ncu -i gpurun_out/r01_${T}_$k.ncu-rep --page details > gpurun_out/r01_${T}_$k.details.txt 2>/dev/null
done
rm -f gpurun_out/r01_${T}_frame.ncu-rep
ls -la gpurun_out | tail -12
