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
#include "trace_pixel.cuh"

#define RT_TRACE_THREADS 256

/* scene accessor over the HBM-resident LBVH (any primitive count) */
template <bool BRUTE, bool COUNT>
struct RtGlobalScene {
    const RtSceneDev& sc;
    __device__ __forceinline__ RtGlobalScene(const RtSceneDev& s) : sc(s) {}
#ifdef RT_NODEMAP
    unsigned int dbg[3] = {0, 0, 0};   /* debug variant: node tests of primary / bounce / shadow rays */
    __device__ __forceinline__ RtHit closest(const vec3& o, const vec3& d, int depth, bool, RtLocalCounters& cnt) {
        const unsigned int before = cnt.node_tests;
        const RtHit h = rt_closest<BRUTE, COUNT>(sc, o, d, cnt);
        dbg[depth ? 1 : 0] += cnt.node_tests - before;
        return h;
    }
    __device__ __forceinline__ bool occluded(const vec3& o, const vec3& d, float tmax, const vec3&, int, RtLocalCounters& cnt) {
        const unsigned int before = cnt.node_tests;
        const bool r = rt_occluded<BRUTE, COUNT>(sc, o, d, tmax, cnt);
        dbg[2] += cnt.node_tests - before;
        return r;
    }
#else
    __device__ __forceinline__ RtHit closest(const vec3& o, const vec3& d, int, bool, RtLocalCounters& cnt) {
        return rt_closest<BRUTE, COUNT>(sc, o, d, cnt);
    }
    __device__ __forceinline__ bool occluded(const vec3& o, const vec3& d, float tmax, const vec3&, int, RtLocalCounters& cnt) {
        return rt_occluded<BRUTE, COUNT>(sc, o, d, tmax, cnt);
    }
#endif
    __device__ __forceinline__ vec3 normal(const RtHit& h, const vec3& P) { return rt_hit_normal<BRUTE>(sc, h, P); }
    __device__ __forceinline__ void material(unsigned int idx, float4& m0, float4& m1) {
        m0 = __ldg(sc.mats + 2 * idx);
        m1 = __ldg(sc.mats + 2 * idx + 1);
    }
};

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

    RtPixelCounters pc = {{0, 0, 0, 0}, 0, 0, 0, 0};
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
    if (brute) {
        if (count) rt_trace_kernel<true, true><<<grid, block, 0, stream>>>(*sc, *f, image, hit_ids, counters);
        else       rt_trace_kernel<true, false><<<grid, block, 0, stream>>>(*sc, *f, image, hit_ids, counters);
    } else {
        if (count) rt_trace_kernel<false, true><<<grid, block, 0, stream>>>(*sc, *f, image, hit_ids, counters);
        else       rt_trace_kernel<false, false><<<grid, block, 0, stream>>>(*sc, *f, image, hit_ids, counters);
    }
    return cudaGetLastError();
}
