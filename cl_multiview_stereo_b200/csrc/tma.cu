// tma.cu -- host side of tma.cuh: tensor maps through the driver entry point (libcuda is not linked).
#include "tma.cuh"
#include <cudaTypedefs.h>
#include <stdio.h>

namespace mvs {

static PFN_cuTensorMapEncodeTiled encode_fn()
{
    static PFN_cuTensorMapEncodeTiled fn = [] {
        void *p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) != cudaSuccess ||
            q != cudaDriverEntryPointSuccess)
            p = nullptr;
        return (PFN_cuTensorMapEncodeTiled)p;
    }();
    return fn;
}

int tma_make_image_map(CUtensorMap *map, const void *base, int px_bytes, int width, int height, int views, int box_w,
                       int box_h, const char **why)
{
    static char msg[160];
    *why = msg;
    PFN_cuTensorMapEncodeTiled enc = encode_fn();
    if (!enc) { snprintf(msg, sizeof msg, "driver entry point cuTensorMapEncodeTiled not found"); return 1; }
    // Lab (16-byte pixels) is described in 8-byte elements, labels (4-byte pixels) in 4-byte elements: a box row of up
    // to 256 elements then spans 128 Lab pixels / 256 labels
    const int eb = px_bytes % 8 == 0 ? 8 : 4;
    const int per_px = px_bytes / eb;
    if (px_bytes % eb || ((size_t)width * px_bytes) % 16 || ((size_t)box_w * px_bytes) % 16 || box_w * per_px > 256 ||
        box_h > 256 || box_w < 1 || box_h < 1 || ((uintptr_t)base & 15)) {
        snprintf(msg, sizeof msg, "image %dx%d (%d-byte pixels), box %dx%d: outside what a tiled tensor map can describe", width,
                 height, px_bytes, box_w, box_h);
        return 1;
    }
    cuuint64_t gdim[2] = {(cuuint64_t)width * per_px, (cuuint64_t)height * views};
    cuuint64_t gstride[1] = {(cuuint64_t)width * px_bytes};
    cuuint32_t box[2] = {(cuuint32_t)(box_w * per_px), (cuuint32_t)box_h};
    cuuint32_t estr[2] = {1, 1};
    CUresult r = enc(map, eb == 8 ? CU_TENSOR_MAP_DATA_TYPE_UINT64 : CU_TENSOR_MAP_DATA_TYPE_UINT32, 2, const_cast<void *>(base),
                     gdim, gstride, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE,
                     CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) { snprintf(msg, sizeof msg, "cuTensorMapEncodeTiled failed with CUresult %d", (int)r); return 1; }
    return 0;
}

} // namespace mvs
