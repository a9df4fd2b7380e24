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

template <bool BRUTE, bool COUNT, bool WIDE>
__global__ void __launch_bounds__(RT_TRACE_THREADS, COUNT ? 1 : 4)
rt_trace_kernel(const RtSceneDev sc, const RtFrameDev f, unsigned int* __restrict__ image,
                int* __restrict__ hit_ids, RtCountersDev* __restrict__ counters, unsigned int* __restrict__ err, int packet) {
    __shared__ float4 s_pending[RTB_MAX_DEPTH * RT_TRACE_THREADS];   /* (phong.rgb, kR) per depth */
    __shared__ int s_wstack[RT_TRACE_THREADS / 32][RT_PACKET_STACK];  /* per-warp stack of the packet traversal */

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    /* block = 32x8 pixels, warp = 8x4 patch */
    const int x = blockIdx.x * 32 + (warp & 3) * 8 + (lane & 7);
    const int y = blockIdx.y * 8 + (warp >> 2) * 4 + (lane >> 3);
    const bool mine = x < f.imgw && y < f.imgh && rt_shard_needs(f, x, y);
    if (!packet && !mine) return;

    RtPixelCounters pc = {{0, 0, 0, 0, 0, 0, 0, 0, 0}, 0, 0, 0, 0};
    RtGlobalPacketScene<BRUTE, COUNT, WIDE> acc(sc);
    if (!BRUTE && packet) {
        /* the patch's 32 primary rays walk the hierarchy as one packet (rt_packet_closest, trace_global.cuh) */
        if (!__any_sync(0xffffffffu, mine)) return;
        const vec3 o = v3(f.pos[0], f.pos[1], f.pos[2]);
        const vec3 d = mine ? rt_primary_dir(f, x, y) : v3(0.0f, 0.0f, 1.0f);
        RtHit best;
        best.t = __int_as_float(0x7f800000);
        best.tag = RT_MISS;
        best.slot = 0;
        if (mine)
            for (int p = 0; p < sc.n_planes; p++) {
                const float4 q0 = __ldg(sc.planes + 2 * p), q1 = __ldg(sc.planes + 2 * p + 1);
                float t;
                if (COUNT) pc.t.plane_tests++;
                if (rt_hit_plane(o, d, f4xyz(q0), f4xyz(q1), t)) rt_consider(best, t, __float_as_uint(q0.w), ~p);
            }
        rt_packet_closest<COUNT>(sc, mine, o, d, best, pc.t, s_wstack[warp], lane, err);
        acc.primary = best;
        acc.have_primary = true;
        if (!mine) return;
    }
#ifdef RT_TIMEMAP
    unsigned long long tm_t0;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(tm_t0));
#endif
    image[y * f.imgw + x] = rt_trace_pixel<RtGlobalPacketScene<BRUTE, COUNT, WIDE>, COUNT>(acc, f, x, y, s_pending, RT_TRACE_THREADS,
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
                                       int* hit_ids, RtCountersDev* counters, unsigned int* err, int brute, int count,
                                       int packet, cudaStream_t stream) {
    dim3 block(RT_TRACE_THREADS, 1, 1);
    dim3 grid((f->imgw + 31) / 32, (f->imgh + 7) / 8, 1);
    RT_LAUNCHED(1);
    if (brute) {
        if (count) rt_trace_kernel<true, true, false><<<grid, block, 0, stream>>>(*sc, *f, image, hit_ids, counters, err, packet);
        else       rt_trace_kernel<true, false, false><<<grid, block, 0, stream>>>(*sc, *f, image, hit_ids, counters, err, packet);
    } else if (sc->wide) {           /* the 4-wide view of the hierarchy is there (RT_OPT_WIDE_BVH, api.cu): walk that */
        if (count) rt_trace_kernel<false, true, true><<<grid, block, 0, stream>>>(*sc, *f, image, hit_ids, counters, err, packet);
        else       rt_trace_kernel<false, false, true><<<grid, block, 0, stream>>>(*sc, *f, image, hit_ids, counters, err, packet);
    } else {
        if (count) rt_trace_kernel<false, true, false><<<grid, block, 0, stream>>>(*sc, *f, image, hit_ids, counters, err, packet);
        else       rt_trace_kernel<false, false, false><<<grid, block, 0, stream>>>(*sc, *f, image, hit_ids, counters, err, packet);
    }
    return cudaGetLastError();
}
