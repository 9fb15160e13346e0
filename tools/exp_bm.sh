#!/bin/bash
# block-matching experiments on the GPU box: parity tests, then stage times with the TMA and the gather kernel
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
python -c 'import __graft_entry__ as g; g.build()' > gpurun_out/exp_build.log 2>&1
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/exp_gpu_tests.log 2>&1; echo "pytest rc=$?" >> gpurun_out/exp_gpu_tests.log
tail -4 gpurun_out/exp_gpu_tests.log
: > gpurun_out/exp_bm.jsonl
python tools/stage_time.py 15view_1080p 5 >> gpurun_out/exp_bm.jsonl 2>> gpurun_out/exp_err.log
MVS_BM_MODE=gather python tools/stage_time.py 15view_1080p 5 >> gpurun_out/exp_bm.jsonl 2>> gpurun_out/exp_err.log
python tools/stage_time.py bundled_15view 5 >> gpurun_out/exp_bm.jsonl 2>> gpurun_out/exp_err.log
MVS_BM_MODE=gather python tools/stage_time.py bundled_15view 5 >> gpurun_out/exp_bm.jsonl 2>> gpurun_out/exp_err.log
for defs in "$@"; do
  MVS_NVCC_DEFS="$defs" python -c 'from cl_multiview_stereo_b200 import build as b; b.build_native()' >> gpurun_out/exp_build.log 2>&1
  MVS_NVCC_DEFS="$defs" python tools/stage_time.py 15view_1080p 5 >> gpurun_out/exp_bm.jsonl 2>> gpurun_out/exp_err.log
done
python -c 'from cl_multiview_stereo_b200 import build as b; b.build_native()' >> gpurun_out/exp_build.log 2>&1
cat gpurun_out/exp_bm.jsonl; tail -3 gpurun_out/exp_err.log
