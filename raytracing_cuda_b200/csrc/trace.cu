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
    /* to_light = L - P (the vector the caller normalised into d), li = index of the light */
    __device__ __forceinline__ bool occluded(const vec3& o, const vec3& d, float tmax, const vec3& to_light, int li,
                                             RtLocalCounters& cnt) {
        /* static indices: a dynamic one would copy the kernel parameter block to local memory */
        static_assert(RT_LGB_LIGHTS == 2, "select below");
        const unsigned int* __restrict__ lg_off = li == 0 ? sc.lg_offsets[0] : (li == 1 ? sc.lg_offsets[1] : nullptr);
        if (!BRUTE && lg_off) {
            /* light grid (light_grid.cu): the ray's line passes through the light, its direction from there is
             * -to_light; the bin of that direction lists every primitive such a line can hit */
            const float ax = fabsf(to_light.x), ay = fabsf(to_light.y), az = fabsf(to_light.z);
            const bool X = ax >= ay && ax >= az, Y = !X && ay >= az;
            const float m = X ? ax : (Y ? ay : az);
            if (m > 0.0f) {
                for (int p = 0; p < sc.n_planes; p++) {          /* planes stay outside the grid (rt_occluded) */
                    const float4 q0 = __ldg(sc.planes + 2 * p), q1 = __ldg(sc.planes + 2 * p + 1);
                    float t;
                    if (COUNT) cnt.plane_tests++;
                    if (__float_as_uint(q0.w) & RT_TAG_EMISSIVE) continue;
                    if (rt_hit_plane(o, d, f4xyz(q0), f4xyz(q1), t) && t < tmax) return true;
                }
                const float qm = X ? to_light.x : (Y ? to_light.y : to_light.z);
                const float qa = X ? to_light.y : (Y ? to_light.z : to_light.x);
                const float qb = X ? to_light.z : (Y ? to_light.x : to_light.y);
                const int face = (X ? 0 : (Y ? 2 : 4)) + (qm > 0.0f ? 1 : 0);      /* direction = -to_light */
                const float scale = (float)RT_LGB_G / (2.0f + 2.0f * 0.00390625f);  /* RS_LG_SCALE_OF(RT_LGB_G) */
                const float k = -scale * rt_rcp_approx(m);
                const float gs = __fmaf_rn(qa, k, scale * (1.0f + 0.00390625f));
                const float gt = __fmaf_rn(qb, k, scale * (1.0f + 0.00390625f));
                const int ix = min(max((int)gs, 0), RT_LGB_G - 1), iy = min(max((int)gt, 0), RT_LGB_G - 1);
                const unsigned int* __restrict__ off = lg_off + ((size_t)face * RT_LGB_G + iy) * RT_LGB_G + ix;
                const unsigned int e0 = __ldg(off), e1 = __ldg(off + 1);
                const unsigned int* __restrict__ ent = li == 0 ? sc.lg_entries[0] : sc.lg_entries[1];
                for (unsigned int e = e0; e < e1; e++) {
                    float t;
                    unsigned int tag;
                    if (rt_test_record<COUNT>(sc.prims + 3 * (size_t)__ldg(ent + e), o, d, t, tag, cnt) &&
                        !(tag & RT_TAG_EMISSIVE) && t < tmax)
                        return true;
                }
                return false;
            }
        }
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
