// tma.cuh -- Tensor Memory Accelerator helpers (sm_100a): tiled tensor maps built on the host through the driver
// entry point (no link against libcuda), cp.async.bulk.tensor loads into shared memory, mbarrier completion.
// Used where a CTA consumes a DENSE box of an image: the neighbour-view Lab window of the block matcher
// (photo.cu) and the label / Lab tile of the centre update (slic.cu).
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>

namespace mvs {

// ---- host: tensor maps ----------------------------------------------------------------------------------------
// 2-D map over a row-major image stack whose PIXELS are `px_bytes` wide (a multiple of 8): the tensor is described in
// 8-byte elements so that one box row may span up to 256 * 8 = 2048 bytes (a box dimension is limited to 256 elements).
//   dim0 = x in 8-byte units (width * px_bytes / 8), dim1 = rows of ALL views (height * views), row pitch = width * px_bytes
// box = box_w pixels x box_h rows.  Out-of-range parts of a box are filled with zeros; a box that runs past the last row
// of a view reads the next view's rows (callers mask by their own bounds tests).
// returns 0 on success; on failure `why` names the driver's error
int tma_make_image_map(CUtensorMap *map, const void *base, int px_bytes, int width, int height, int views, int box_w,
                       int box_h, const char **why);

// ---- device ---------------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t *bar, uint32_t count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}

// make the initialised barrier visible to the async proxy (the TMA unit) before anything is issued on it
__device__ __forceinline__ void mbar_fence_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }

// generic-proxy accesses of shared memory (ordinary loads / stores) before this point are ordered before async-proxy
// accesses (TMA writes) issued after it: needed before a buffer that threads have read is refilled by TMA
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

__device__ __forceinline__ void mbar_arrive_expect_tx(uint64_t *bar, uint32_t bytes)
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}

__device__ __forceinline__ void mbar_wait(uint64_t *bar, uint32_t parity)
{
    uint32_t done;
    do {
        asm volatile("{\n\t.reg .pred p;\n\t"
                     "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
                     "selp.u32 %0, 1, 0, p;\n\t}"
                     : "=r"(done)
                     : "r"(smem_u32(bar)), "r"(parity)
                     : "memory");
    } while (!done);
}

// one box of a 2-D tensor map into shared memory (dst 128-byte aligned); completion = box bytes on `bar`
__device__ __forceinline__ void tma_load_2d(void *dst, const CUtensorMap *map, int c0, int c1, uint64_t *bar)
{
    asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
                 ::"r"(smem_u32(dst)), "l"(map), "r"(smem_u32(bar)), "r"(c0), "r"(c1)
                 : "memory");
}

__device__ __forceinline__ void tma_prefetch_desc(const CUtensorMap *map)
{
    asm volatile("prefetch.tensormap [%0];" ::"l"(map) : "memory");
}

} // namespace mvs
