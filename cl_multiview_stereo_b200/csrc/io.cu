// io.cu -- device-side ingest and export of the depth pipeline (sm_100a): byte work that the
// reference does in host loops around OpenCV.
//   k_ingest_bgr      loadImageIn (file_handler.cpp:6-15): BGR raster (cv::Mat, 3 B/px) -> vec3u
//                     {s0 = R, s1 = G, s2 = B, pad}; uploading 3 B/px instead of 4 and repacking here
//   k_quantise_disp   plot_full_image (depth_refinement.cpp:1476-1497): 8-bit image of a disparity map,
//                     (unsigned char)floor((d - min) / (max - min) * 255)
//   k_quantise_state  img_translate_state (depth_refinement.cpp:1500-1560): per-pixel image of one state
//                     element of the pixel's superpixel, elements 0 (d), 1 (smoothness), 2 (consistency)
//   k_quantise_spixl  img_translate (photo_consistency.cpp:414-436): the block-matching disparity s7 of the
//                     pixel's superpixel, floor-quantised like plot_full_image; img_translate_flatness
//                     (depth_refinement.cpp:299-321): (unsigned char)ceil(flatness.s0 * 255)
//   k_segmentation_lines  clSLIC::draw_segmentation_lines (clSLIC.cpp:447-478): interior pixels whose
//                     4-neighbourhood holds another label become {0,0,255}, the others take the input
//                     colour; the one-pixel frame and the pad byte of the output are never written
#include "common.cuh"
#include "mvs_math.cuh"

namespace mvs {

// 4 pixels per thread: three 32-bit loads, one 128-bit store
__global__ void __launch_bounds__(256) k_ingest_bgr(const uint32_t *__restrict__ bgr, uint4 *__restrict__ rgbx,
                                                    const uint8_t *__restrict__ bgr8, uchar4 *__restrict__ rgbx1,
                                                    size_t n, int vec)
{
    const size_t q = (size_t)blockIdx.x * blockDim.x + threadIdx.x;   // quad index
    const size_t p = 4 * q;
    if (vec && p + 4 <= n) {
        const uint32_t a = bgr[3 * q], b = bgr[3 * q + 1], c = bgr[3 * q + 2];   // B0 G0 R0 B1 | G1 R1 B2 G2 | R2 B3 G3 R3
        uint4 o;
        o.x = ((a >> 16) & 255u) | (a & 0xff00u) | ((a & 255u) << 16);
        o.y = ((b >> 8) & 255u) | ((b & 255u) << 8) | ((a >> 24) << 16);
        o.z = (c & 255u) | ((b >> 24) << 8) | (((b >> 16) & 255u) << 16);
        o.w = (c >> 24) | (((c >> 16) & 255u) << 8) | (((c >> 8) & 255u) << 16);
        rgbx[q] = o;
    } else {
        for (size_t k = p; k < n && k < p + 4; ++k) rgbx1[k] = make_uchar4(bgr8[3 * k + 2], bgr8[3 * k + 1], bgr8[3 * k], 0);
    }
}

// (unsigned char)floor(x): the reference's cast is only defined for 0 <= x < 256; outside it this pins
// what the reference's own platform does (x86 cvttss2si then the low byte; NaN and |x| >= 2^31 give 0)
__device__ __forceinline__ uint8_t to_u8(float x)
{
    const float fl = floorf(x);
    const int v = (fl >= -2147483648.0f && fl < 2147483648.0f) ? (int)fl : (int)0x80000000;
    return (uint8_t)(v & 255);
}

// four pixels per thread: one 128-bit load, one 32-bit store
__global__ void __launch_bounds__(256) k_quantise_disp(const float *__restrict__ disp, uint8_t *__restrict__ out, size_t n,
                                                       float mn, float range, int vec)
{
    const size_t p = 4 * ((size_t)blockIdx.x * blockDim.x + threadIdx.x);
    if (p >= n) return;
    if (vec && p + 4 <= n) {
        const float4 d = *reinterpret_cast<const float4 *>(disp + p);
        const uint32_t o = (uint32_t)to_u8(((d.x - mn) / range) * 255.0f) | ((uint32_t)to_u8(((d.y - mn) / range) * 255.0f) << 8) |
                           ((uint32_t)to_u8(((d.z - mn) / range) * 255.0f) << 16) | ((uint32_t)to_u8(((d.w - mn) / range) * 255.0f) << 24);
        *reinterpret_cast<uint32_t *>(out + p) = o;
    } else {
        for (size_t k = p; k < n && k < p + 4; ++k) out[k] = to_u8(((disp[k] - mn) / range) * 255.0f);
    }
}

__global__ void __launch_bounds__(256) k_quantise_state(const uint32_t *__restrict__ idx, const float4 *__restrict__ plane,
                                                        const float2 *__restrict__ cost, uint8_t *__restrict__ out, int N,
                                                        int M, int element, float mn, float range, int z0)
{
    const int p = blockIdx.x * blockDim.x + threadIdx.x;
    const int z = z0 + blockIdx.y;
    if (p >= N) return;
    const size_t q = (size_t)z * M + idx[(size_t)z * N + p];
    float v;
    if (element == 0) v = ((plane[q].x - mn) / range) * 255.0f;
    else if (element == 1) v = cost[q].x * 255.0f;
    else v = cost[q].y * 255.0f;
    out[(size_t)blockIdx.y * N + p] = to_u8(v);
}

// per-pixel image of one float of the pixel's superpixel record (`stride` floats per superpixel, element
// `off`).  mode 0: floor(((v - mn) / range) * 255); mode 1: ceil(v * 255)
__global__ void __launch_bounds__(256) k_quantise_spixl(const uint32_t *__restrict__ idx, const float *__restrict__ src,
                                                        uint8_t *__restrict__ out, int N, int M, int stride, int off,
                                                        int mode, float mn, float range, int z0)
{
    const int p = blockIdx.x * blockDim.x + threadIdx.x;
    const int z = z0 + blockIdx.y;
    if (p >= N) return;
    const float v = src[((size_t)z * M + idx[(size_t)z * N + p]) * stride + off];
    // to_u8 floors an already integral value in mode 1; it pins the out-of-range cast for both modes
    out[(size_t)blockIdx.y * N + p] = mode == 0 ? to_u8(((v - mn) / range) * 255.0f) : to_u8(ceilf(v * 255.0f));
}

// one thread per pixel of the interior; `out` is read-modify-written so that its pad byte survives
__global__ void __launch_bounds__(256) k_segmentation_lines(const uint32_t *__restrict__ idx, const uchar4 *__restrict__ in,
                                                            uchar4 *__restrict__ out, int W, int H, int z0)
{
    const int x = blockIdx.x * blockDim.x + threadIdx.x;
    const int y = blockIdx.y;
    if (x < 1 || x >= W - 1 || y < 1 || y >= H - 1) return;
    const size_t N = (size_t)W * H;
    const size_t p = (size_t)(z0 + blockIdx.z) * N + (size_t)y * W + x;
    const size_t q = (size_t)blockIdx.z * N + (size_t)y * W + x;
    const uint32_t l = idx[p];
    const bool edge = l != idx[p + 1] || l != idx[p - 1] || l != idx[p - W] || l != idx[p + W];
    uchar4 o = out[q];
    if (edge) { o.x = 0; o.y = 0; o.z = 255; }
    else { const uchar4 c = in[p]; o.x = c.x; o.y = c.y; o.z = c.z; }
    out[q] = o;
}

int launch_ingest_bgr(const FrameDev &f, const void *bgr_dev, int z0, int z1, cudaStream_t st)
{
    if (z1 <= z0) return 0;
    const size_t n = (size_t)(z1 - z0) * f.N;
    const size_t quads = (n + 3) / 4;
    uchar4 *dst = f.rgbx + (size_t)z0 * f.N;
    k_ingest_bgr<<<(unsigned)((quads + 255) / 256), 256, 0, st>>>((const uint32_t *)bgr_dev, (uint4 *)dst,
                                                                 (const uint8_t *)bgr_dev, dst, n,
                                                                 ((size_t)z0 * f.N) % 4 == 0 ? 1 : 0);
    return 1;
}

int launch_quantise_disp(const FrameDev &f, const float *src, uint8_t *out_dev, float mn, float range, int z0, int z1,
                         cudaStream_t st)
{
    if (z1 <= z0) return 0;
    const size_t n = (size_t)(z1 - z0) * f.N;
    const size_t quads = (n + 3) / 4;
    k_quantise_disp<<<(unsigned)((quads + 255) / 256), 256, 0, st>>>(src + (size_t)z0 * f.N, out_dev, n, mn, range,
                                                                    ((size_t)z0 * f.N) % 4 == 0 ? 1 : 0);
    return 1;
}

int launch_quantise_state(const FrameDev &f, int which, int element, uint8_t *out_dev, float mn, float range, int z0,
                          int z1, cudaStream_t st)
{
    if (z1 <= z0) return 0;
    dim3 g((f.N + 255) / 256, z1 - z0);
    k_quantise_state<<<g, 256, 0, st>>>(f.idx, f.plane[which], f.cost[which], out_dev, f.N, f.M, element, mn, range, z0);
    return 1;
}

int launch_quantise_spixl(const FrameDev &f, const float *src, int stride, int off, int mode, uint8_t *out_dev, float mn,
                          float range, int z0, int z1, cudaStream_t st)
{
    if (z1 <= z0) return 0;
    dim3 g((f.N + 255) / 256, z1 - z0);
    k_quantise_spixl<<<g, 256, 0, st>>>(f.idx, src, out_dev, f.N, f.M, stride, off, mode, mn, range, z0);
    return 1;
}

int launch_segmentation_lines(const FrameDev &f, uchar4 *out_dev, int z0, int z1, cudaStream_t st)
{
    if (z1 <= z0 || f.W < 3 || f.H < 3) return 0;
    dim3 g((f.W + 255) / 256, f.H, z1 - z0);
    k_segmentation_lines<<<g, 256, 0, st>>>(f.idx, f.rgbx, out_dev, f.W, f.H, z0);
    return 1;
}

} // namespace mvs
