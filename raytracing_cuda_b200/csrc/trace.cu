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
#include "trace_common.cuh"

#define RT_TRACE_THREADS 256

template <bool BRUTE, bool COUNT>
__global__ void __launch_bounds__(RT_TRACE_THREADS)
rt_trace_kernel(const RtSceneDev sc, const RtFrameDev f, unsigned int* __restrict__ image,
                int* __restrict__ hit_ids, RtCountersDev* __restrict__ counters) {
    __shared__ float4 s_pending[RTB_MAX_DEPTH][RT_TRACE_THREADS];   /* (phong.rgb, kR) per depth */

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    /* block = 32x8 pixels, warp = 8x4 patch */
    const int x = blockIdx.x * 32 + (warp & 3) * 8 + (lane & 7);
    const int y = blockIdx.y * 8 + (warp >> 2) * 4 + (lane >> 3);
    if (x >= f.imgw || y >= f.imgh) return;
    if (!rt_shard_needs(f, x, y)) return;

    RtLocalCounters cnt = {0, 0, 0, 0};
    unsigned int n_shadow = 0, n_mirror = 0, n_sky = 0, n_shadow_skipped = 0;

    vec3 o = v3(f.pos[0], f.pos[1], f.pos[2]);
    vec3 d = rt_primary_dir(f, x, y);
    vec3 c = v3(0, 0, 0);
    int depth = 0;

    for (;;) {
        const RtHit h = rt_closest<BRUTE, COUNT>(sc, o, d, cnt);
        if (depth == 0 && hit_ids) hit_ids[y * f.imgw + x] = (h.tag == RT_MISS) ? -1 : (int)(h.tag & RT_TAG_INDEX);

        if (h.tag == RT_MISS) {
            if (COUNT) n_sky++;
            c = rt_sky(f, d);
            break;
        }
        const unsigned int idx = h.tag & RT_TAG_INDEX;
        const float4 m0 = __ldg(sc.mats + 2 * idx);       /* color.rgb, mirror */
        const float4 m1 = __ldg(sc.mats + 2 * idx + 1);   /* shine, specular, emissive, type */
        const vec3 color = f4xyz(m0);
        if (h.tag & RT_TAG_EMISSIVE) {                    /* reference kernel.cu:169 */
            c = color;
            break;
        }
        const vec3 P = rt_along(o, d, h.t);
        const vec3 N = rt_hit_normal<BRUTE>(sc, h, P);

        /* Phong, reference kernel.cu:171-206 */
        vec3 phong = vhadamard(color, v3(f.ambient[0], f.ambient[1], f.ambient[2]));
        for (int li = 0; li < f.n_lights; li++) {
            const RtLightDev L = f.lights[li];
            const vec3 vec = rt_sub(v3(L.px, L.py, L.pz), P);
            const float shadowDist = norm3df(vec.x, vec.y, vec.z);
            const float inv = __frcp_rn(shadowDist);
            const vec3 sdir = v3(__fmul_rn(vec.x, inv), __fmul_rn(vec.y, inv), __fmul_rn(vec.z, inv));
            const vec3 so = rt_along(P, sdir, 0.001f);
            float angle = fmaxf(0.0f, rt_dot(N, sdir));
            if (COUNT) n_shadow++;
            if (angle > 0.0f) {
                if (rt_occluded<BRUTE, COUNT>(sc, so, sdir, shadowDist, cnt)) angle = 0.0f;
            } else if (COUNT) {
                n_shadow_skipped++;
            }
            const float ai = angle * L.intensity;
            phong = vadd(phong, vscale(vhadamard(color, v3(L.cr, L.cg, L.cb)), ai));
            if (m1.x > 0.0f) {                            /* specular, reference kernel.cu:199-203 */
                const vec3 ldir = vscale(sdir, -1.0f);
                const float k = 2 * vdot(N, ldir);
                const vec3 sd = rt_normalize(vsub(ldir, vscale(N, k)));
                const float base = fmaxf(0.0f, -vdot(sd, d));
                const float p = __powf(base, m1.y);
                const float s = p * m1.x * angle;
                phong = vadd(phong, v3(s, s, s));
            }
        }

        const float kR = m0.w;
        if (kR > 0.0f && depth < RTB_MAX_DEPTH) {          /* mirror bounce, reference kernel.cu:209-215 */
            s_pending[depth][threadIdx.x] = make_float4(phong.x, phong.y, phong.z, kR);
            const float k = 2 * vdot(N, d);
            const vec3 rd = rt_normalize(vsub(d, vscale(N, k)));
            o = rt_along(P, rd, 0.001f);
            d = rd;
            depth++;
            if (COUNT) n_mirror++;
            continue;
        }
        /* no bounce (kR <= 0), or trace<5> returned black (kernel.cu:222-225): refColor = 0 */
        c = vadd(vscale(v3(0, 0, 0), kR), vscale(phong, 1 - kR));
        break;
    }
    /* unwind the recursion: reference kernel.cu:218 */
    for (int dd = depth - 1; dd >= 0; dd--) {
        const float4 pk = s_pending[dd][threadIdx.x];
        c = vadd(vscale(c, pk.w), vscale(v3(pk.x, pk.y, pk.z), 1 - pk.w));
    }

    c = vscale(c, 255.0f);
    image[y * f.imgw + x] = rt_pack(c.x, c.y, c.z);

    if (COUNT) {
        /* one atomic per warp and counter, not per thread */
        const unsigned int mask = __activemask();
        const bool leader = (threadIdx.x & 31) == (unsigned int)(__ffs(mask) - 1);
        const unsigned int vals[9] = {1u, n_shadow, n_mirror, cnt.node_tests, cnt.tri_tests, cnt.sphere_tests,
                                      cnt.plane_tests, n_sky, n_shadow_skipped};
#pragma unroll
        for (int k = 0; k < 9; k++) {
            const unsigned int s = __reduce_add_sync(mask, vals[k]);
            if (leader) atomicAdd(&counters->v[k], (unsigned long long)s);
        }
    }
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
