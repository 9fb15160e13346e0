#!/bin/bash
# centre-update experiments on the GPU box: SLIC parity tests on the default build, then stage times of variants
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
O=gpurun_out/expu
python -c 'import __graft_entry__ as g; g.build()' > ${O}_build.log 2>&1
python -c 'import cl_multiview_stereo_b200 as m; m._native.lib()' >> ${O}_build.log 2>&1 || { echo "library does not load"; tail -5 ${O}_build.log; exit 1; }
timeout 900 python -m pytest tests -m gpu -x -q -k "slic or fullframe or golden" > ${O}_tests.log 2>&1; echo "pytest rc=$?" >> ${O}_tests.log
tail -3 ${O}_tests.log
: > ${O}.jsonl
python tools/stage_time.py 15view_1080p 5 >> ${O}.jsonl 2>> ${O}_err.log
MVS_UPD_MODE=tma python tools/stage_time.py 15view_1080p 5 >> ${O}.jsonl 2>> ${O}_err.log
python bench.py --workload slic_4k --steps 5 --warmup 3 > ${O}_slic4k.json 2>> ${O}_err.log
MVS_UPD_MODE=tma python bench.py --workload slic_4k --steps 5 --warmup 3 > ${O}_slic4k_tma.json 2>> ${O}_err.log
for defs in "$@"; do
  MVS_NVCC_DEFS="$defs" python -c 'from cl_multiview_stereo_b200 import build as b; b.build_native()' >> ${O}_build.log 2>&1
  MVS_NVCC_DEFS="$defs" python tools/stage_time.py 15view_1080p 5 >> ${O}.jsonl 2>> ${O}_err.log
done
python -c 'from cl_multiview_stereo_b200 import build as b; b.build_native()' >> ${O}_build.log 2>&1
python - <<'P'
import json
for l in open('gpurun_out/expu.jsonl'):
    d = json.loads(l)
    print(d['workload'], d['defs'] or '-', d['mode'] or '-', d['ms_per_frame'], 'slic', d['stage_ms']['slic'], d['sha1']['idx'], d['sha1']['map'], d['sha1']['disp'])
for n in ('slic4k', 'slic4k_tma'):
    try:
        d = json.loads(open(f'gpurun_out/expu_{n}.json').read().strip().splitlines()[-1])
        print(n, d['value'], d['unit'], 'e2e', d['e2e']['value'], 'frac', d['roofline']['frac'])
    except Exception as e:
        print(n, 'failed', e)
P
tail -3 ${O}_err.log
