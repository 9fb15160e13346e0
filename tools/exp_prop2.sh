#!/bin/bash
# propagate experiments on the GPU box: parity tests on the default build, then stage times of build variants
# usage: tools/exp_prop2.sh "<defs of variant 1>" "<defs of variant 2>" ...   (the default build is always measured)
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
O=gpurun_out/exp2
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv,noheader > ${O}_gpu.txt
python -c 'import __graft_entry__ as g; g.build()' > ${O}_build.log 2>&1
python -c 'import cl_multiview_stereo_b200 as m; m._native.lib()' >> ${O}_build.log 2>&1 || { echo "library does not load"; tail -5 ${O}_build.log; exit 1; }
if [ -z "$SKIP_TESTS" ]; then
  timeout 1500 python -m pytest tests -m gpu -x -q > ${O}_gpu_tests.log 2>&1; echo "pytest rc=$?" >> ${O}_gpu_tests.log
  tail -3 ${O}_gpu_tests.log
fi
: > ${O}.jsonl
for w in 15view_1080p bundled_15view noise_15view_1080p; do
  python tools/stage_time.py $w 5 >> ${O}.jsonl 2>> ${O}_err.log
done
for defs in "$@"; do
  MVS_NVCC_DEFS="$defs" python -c 'from cl_multiview_stereo_b200 import build as b; b.build_native()' >> ${O}_build.log 2>&1
  for w in 15view_1080p bundled_15view; do
    MVS_NVCC_DEFS="$defs" python tools/stage_time.py $w 5 >> ${O}.jsonl 2>> ${O}_err.log
  done
done
python -c 'from cl_multiview_stereo_b200 import build as b; b.build_native()' >> ${O}_build.log 2>&1
if [ -n "$WITH_BENCH" ]; then
python bench.py --steps 10 --warmup 3 --no-cpu --no-extra > ${O}_bench_free.json 2>> ${O}_err.log
MVS_E2E_SERIAL=1 python bench.py --steps 10 --warmup 3 --no-cpu --no-extra > ${O}_bench_serial.json 2>> ${O}_err.log
fi
python - <<'P'
import json
for l in open('gpurun_out/exp2.jsonl'):
    d = json.loads(l)
    print(d['workload'], d['defs'] or '-', d['mode'] or '-', d['ms_per_frame'], 'prop', d['stage_ms']['propagate'], 'bm', d['stage_ms']['block_match'], d['sha1']['disp'])
for n in ('serial', 'free'):
    try:
        d = json.loads(open(f'gpurun_out/exp2_bench_{n}.json').read().strip().splitlines()[-1])
        print(n, 'resident', d['ms_per_step'], 'e2e', d['e2e']['ms_per_step'])
    except Exception as e:
        print(n, 'failed', e)
P
