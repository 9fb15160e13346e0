// mathcheck.cu -- the device copies of the pinned math (mvs_math.cuh) evaluated on caller-supplied arguments, so that a
// test can hold them bit for bit against the oracle's copies (oracle/oracle_math.h) on the GPU itself.  A diagnostic
// entry point (mvs_math_eval), not part of the pipeline.
#include "common.cuh"
#include "mvs_math.cuh"

namespace mvs {

__global__ void k_math_eval(int which, const float *__restrict__ x, float *__restrict__ y, size_t n)
{
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float v = x[i];
    float r;
    switch (which) {
    case 0: r = expf_pinned<false>(v); break;
    case 1: r = expf_pinned<true>(v); break;       // valid for v <= 0 or NaN
    case 2: r = expf_pinned_small(v); break;       // valid for |v| < 0.3465
    case 3: r = cbrtf_pos(v); break;               // v > 0
    case 4: r = round_haz(v); break;
    default: r = __int_as_float(f2i(v)); break;    // 5: float -> int, the int's bits
    }
    y[i] = r;
}

int launch_math_eval(int which, const float *x_dev, float *y_dev, size_t n, cudaStream_t st)
{
    if (!n) return 0;
    k_math_eval<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(which, x_dev, y_dev, n);
    return 1;
}

} // namespace mvs
