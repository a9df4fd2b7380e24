/*
 * trace.cu -- the per-pixel tracer kernel (replaces the reference's `raytracing`
 * kernel, reference kernel.cu:228-259, and the trace<depth> recursion it inlines,
 * kernel.cu:131-225).
 *
 * One thread owns one pixel.  A warp covers an 8x4 pixel patch (coherent rays), a
 * 256-thread block a 32x8 region.  The template recursion of the reference
 * (primary + up to 4 mirror bounces) is an explicit loop; the per-depth Phong
 * colour and mirror weight wait in shared memory and are folded back innermost
 * first, so the blend order is the reference's (kernel.cu:218).
 *
 * Exact work-saving shortcuts (bit-identical results):
 *   - a shadow ray is traced only if N.L > 0: the reference's shadow loop can
 *     only set `angle` to 0 (kernel.cu:185-193);
 *   - only sky maps with a non-zero weight are fetched (see rt_sky).
 * The ray COUNT reported by the benchmark is the reference's (every shadow ray
 * built at kernel.cu:182 counts), independent of these shortcuts.
 */
#include "trace_global.cuh"

#define RT_TRACE_THREADS 256

template <bool BRUTE, bool COUNT>
__global__ void __launch_bounds__(RT_TRACE_THREADS)
rt_trace_kernel(const RtSceneDev sc, const RtFrameDev f, unsigned int* __restrict__ image,
                int* __restrict__ hit_ids, RtCountersDev* __restrict__ counters) {
    __shared__ float4 s_pending[RTB_MAX_DEPTH * RT_TRACE_THREADS];   /* (phong.rgb, kR) per depth */

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    /* block = 32x8 pixels, warp = 8x4 patch */
    const int x = blockIdx.x * 32 + (warp & 3) * 8 + (lane & 7);
    const int y = blockIdx.y * 8 + (warp >> 2) * 4 + (lane >> 3);
    if (x >= f.imgw || y >= f.imgh) return;
    if (!rt_shard_needs(f, x, y)) return;

    RtPixelCounters pc = {{0, 0, 0, 0, 0, 0, 0, 0, 0}, 0, 0, 0, 0};
    RtGlobalScene<BRUTE, COUNT> acc(sc);
#ifdef RT_TIMEMAP
    unsigned long long tm_t0;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(tm_t0));
#endif
    image[y * f.imgw + x] = rt_trace_pixel<RtGlobalScene<BRUTE, COUNT>, COUNT>(acc, f, x, y, s_pending, RT_TRACE_THREADS,
                                                                              threadIdx.x, hit_ids, pc);
#ifdef RT_TIMEMAP
    if (hit_ids) {   /* debug variant: ns this pixel's thread took */
        unsigned long long tm_t1;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(tm_t1));
        hit_ids[y * f.imgw + x] = (int)(tm_t1 - tm_t0);
    }
#endif
#ifdef RT_NODEMAP
    if (COUNT && hit_ids) hit_ids[y * f.imgw + x] = (int)acc.dbg[RT_NODEMAP];   /* debug variant: per-pixel node tests of one ray kind */
#endif
    if (COUNT) rt_flush_counters(pc, 1u, counters);
}

extern "C" cudaError_t rt_launch_trace(const RtSceneDev* sc, const RtFrameDev* f, unsigned int* image,
                                       int* hit_ids, RtCountersDev* counters, int brute, int count,
                                       cudaStream_t stream) {
    dim3 block(RT_TRACE_THREADS, 1, 1);
    dim3 grid((f->imgw + 31) / 32, (f->imgh + 7) / 8, 1);
    RT_LAUNCHED(1);
    if (brute) {
        if (count) rt_trace_kernel<true, true><<<grid, block, 0, stream>>>(*sc, *f, image, hit_ids, counters);
        else       rt_trace_kernel<true, false><<<grid, block, 0, stream>>>(*sc, *f, image, hit_ids, counters);
    } else {
        if (count) rt_trace_kernel<false, true><<<grid, block, 0, stream>>>(*sc, *f, image, hit_ids, counters);
        else       rt_trace_kernel<false, false><<<grid, block, 0, stream>>>(*sc, *f, image, hit_ids, counters);
    }
    return cudaGetLastError();
}
