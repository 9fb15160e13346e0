# multi-GPU bench lines on N GPUs of one box (gpurun --gpus N):  N=8 TAG=r02_v6 bash tools/scale_run.sh "15view_1080p 5x5_4k 4x4_1080p"
N=${N:-8}; T=${TAG:-r02}
mkdir -p gpurun_out
P=29620
for w in $1; do
  P=$((P+1))
  python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $P bench.py --gpus $N --workload $w --steps 10 --warmup 3 --no-extra > gpurun_out/${T}_scale_${w}_n${N}.json 2> gpurun_out/${T}_scale_${w}_n${N}.err
  python - <<PY
import json
try:
    d=json.load(open("gpurun_out/${T}_scale_${w}_n${N}.json"))
    vs=d.get("view_sharded") or {}
    print("$w N=$N value %.2f e2e %.2f | sharded %.3f ms (%s) exch %.3f identical %s others %s" % (d["value"], d["e2e"]["value"], vs.get("ms_per_step",0), vs.get("partition"), vs.get("exchange_ms_rank0",0), vs.get("bit_identical_to_1gpu"), {k: (round(o["ms_per_step"],3), o["bit_identical_to_1gpu"]) for k, o in (vs.get("other_partition") or {}).items()}))
except Exception as e:
    print("$w N=$N failed:", e); print(open("gpurun_out/${T}_scale_${w}_n${N}.err").read()[-800:])
PY
done
