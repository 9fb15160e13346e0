# A/B of the propagate kernel variants (MODE bit0 = 256-bit record gather, bit1 = nine labels first)
set -x
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; tail -3 gpurun_out/pytest_gpu.log
for m in 0 1 2 3; do MVS_PROP_MODE=$m python tools/prop_time.py 15view_1080p 3; done 2>&1 | grep MINB
MVS_PROP_MODE=1 MVS_PROP_MINB=4 python tools/prop_time.py 15view_1080p 3 2>&1 | grep MINB
MVS_PROP_MODE=3 MVS_PROP_MINB=4 python tools/prop_time.py 15view_1080p 3 2>&1 | grep MINB
MVS_PROP_MODE=1 MVS_PROP_MINB=6 python tools/prop_time.py 15view_1080p 3 2>&1 | grep MINB
