/*
 * trace_pixel.cuh -- the per-pixel ray loop shared by the tracer kernels
 * (restates reference kernel.cu:131-225, trace<depth>, and 244-257).
 *
 * The template recursion of the reference (primary + up to 4 mirror bounces) is an
 * explicit loop; the per-depth Phong colour and mirror weight wait in shared memory
 * and are folded back innermost first, so the blend order is the reference's
 * (kernel.cu:218).
 *
 * ACC is the scene accessor of the calling kernel:
 *     RtHit  closest(o, d, depth, off_plane0, cnt)  closest hit of a ray of generation `depth`; off_plane0: the ray
 *                                           is the bounce of a PRIMARY ray off plane 0 (the sea: mirror bins)
 *     bool   occluded(o, d, tmax, cnt)      any non-emissive hit with t < tmax (kernel.cu:188-193)
 *     vec3   normal(hit, P)                 surface normal at the hit (kernel.cu:66, 87, 123)
 *     void   material(idx, m0, m1)          (color.rgb, mirror) (shine, specular, emissive, type)
 *     bool   defers(depth, off_plane0)      rt_trace_chain<DEFER>: hand this ray to a compacted batch instead of
 *                                           tracing it here (only accessors used with DEFER need it)
 *
 * Exact work-saving shortcuts (bit-identical results):
 *   - a shadow ray is traced only if N.L > 0: the reference's shadow loop can
 *     only set `angle` to 0 (kernel.cu:185-193);
 *   - only sky maps with a non-zero weight are fetched (see rt_sky).
 * The ray COUNT reported by the benchmark is the reference's (every shadow ray
 * built at kernel.cu:182 counts), independent of these shortcuts.
 */
#ifndef RTB_TRACE_PIXEL_CUH
#define RTB_TRACE_PIXEL_CUH

#include "trace_common.cuh"

struct RtPixelCounters {
    RtLocalCounters t;
    unsigned int n_shadow, n_mirror, n_sky, n_shadow_skipped;
};

/* A ray of a pixel's chain: generation `depth` (0 = primary), and whether it is the bounce of a PRIMARY ray off
 * plane 0 (the sea: mirror bins).  Together with the s_pending entries below `depth` this is everything a pixel's
 * continuation needs, which is what lets a ray move to another thread (ray compaction, trace_small.cu). */
struct RtRay {
    vec3 o, d;
    int depth;
    bool off_plane0;
};

/* Runs a pixel's chain of rays from `ray` on.  Returns true when the pixel is finished (`rgba` holds the packed
 * colour, the recursion unwound through s_pending, reference kernel.cu:218).  With DEFER, returns false as soon
 * as the NEXT ray of the chain is one the accessor wants to hand to a compacted batch (acc.defers: it would
 * traverse the hierarchy): `ray` then describes that ray and s_pending[0 .. ray.depth) the colours waiting for it.
 * s_pending: [RTB_MAX_DEPTH][blockDim.x] float4, this thread's column is s_pending[depth * stride + tid]. */
template <class ACC, bool COUNT, bool DEFER>
__device__ __forceinline__ bool rt_trace_chain(ACC& acc, const RtFrameDev& f, int x, int y, RtRay& ray,
                                               float4* __restrict__ s_pending, int stride, int tid,
                                               int* __restrict__ hit_ids, RtPixelCounters& pc, unsigned int& rgba) {
    vec3 o = ray.o, d = ray.d;
    vec3 c = v3(0, 0, 0);
    int depth = ray.depth;
    bool off_plane0 = ray.off_plane0;

    for (;;) {
        const RtHit h = acc.closest(o, d, depth, off_plane0, pc.t);
        if (depth == 0 && hit_ids) hit_ids[y * f.imgw + x] = (h.tag == RT_MISS) ? -1 : (int)(h.tag & RT_TAG_INDEX);

        if (h.tag == RT_MISS) {
            if (COUNT) pc.n_sky++;
            c = rt_sky(f, d);
            break;
        }
        const unsigned int idx = h.tag & RT_TAG_INDEX;
        float4 m0, m1;
        acc.material(idx, m0, m1);
        const vec3 color = f4xyz(m0);
        if (h.tag & RT_TAG_EMISSIVE) {                    /* reference kernel.cu:169 */
            c = color;
            break;
        }
        const vec3 P = rt_along(o, d, h.t);
        const vec3 N = acc.normal(h, P);

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
            if (COUNT) pc.n_shadow++;
            if (angle > 0.0f) {
                if (acc.occluded(so, sdir, shadowDist, vec, li, pc.t)) angle = 0.0f;
            } else if (COUNT) {
                pc.n_shadow_skipped++;
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
            s_pending[depth * stride + tid] = make_float4(phong.x, phong.y, phong.z, kR);
            const float k = 2 * vdot(N, d);
            const vec3 rd = rt_normalize(vsub(d, vscale(N, k)));
            o = rt_along(P, rd, 0.001f);
            d = rd;
            off_plane0 = depth == 0 && h.slot == ~0;
            depth++;
            if (COUNT) pc.n_mirror++;
            if constexpr (DEFER) {
                if (acc.defers(depth, off_plane0)) {
                    ray.o = o;
                    ray.d = d;
                    ray.depth = depth;
                    ray.off_plane0 = off_plane0;
                    return false;
                }
            }
            continue;
        }
        /* no bounce (kR <= 0), or trace<5> returned black (kernel.cu:222-225): refColor = 0 */
        c = vadd(vscale(v3(0, 0, 0), kR), vscale(phong, 1 - kR));
        break;
    }
    /* unwind the recursion: reference kernel.cu:218 */
    for (int dd = depth - 1; dd >= 0; dd--) {
        const float4 pk = s_pending[dd * stride + tid];
        c = vadd(vscale(c, pk.w), vscale(v3(pk.x, pk.y, pk.z), 1 - pk.w));
    }
    c = vscale(c, 255.0f);
    rgba = rt_pack(c.x, c.y, c.z);
    return true;
}

/* the whole pixel on one thread: primary ray of pixel (x, y), then its chain to the end */
template <class ACC, bool COUNT>
__device__ __forceinline__ unsigned int rt_trace_pixel(ACC& acc, const RtFrameDev& f, int x, int y,
                                                       float4* __restrict__ s_pending, int stride, int tid,
                                                       int* __restrict__ hit_ids, RtPixelCounters& pc) {
    RtRay ray;
    ray.o = v3(f.pos[0], f.pos[1], f.pos[2]);
    ray.d = rt_primary_dir(f, x, y);
    ray.depth = 0;
    ray.off_plane0 = false;
    unsigned int rgba = 0;
    rt_trace_chain<ACC, COUNT, false>(acc, f, x, y, ray, s_pending, stride, tid, hit_ids, pc, rgba);
    return rgba;
}

/* one atomic per warp and counter, not per thread */
__device__ __forceinline__ void rt_flush_counters(const RtPixelCounters& pc, unsigned int n_pixels,
                                                  RtCountersDev* __restrict__ counters) {
    const unsigned int mask = __activemask();
    const bool leader = (threadIdx.x & 31) == (unsigned int)(__ffs(mask) - 1);
    const unsigned int vals[14] = {n_pixels, pc.n_shadow, pc.n_mirror, pc.t.node_tests, pc.t.tri_tests,
                                   pc.t.sphere_tests, pc.t.plane_tests, pc.n_sky, pc.n_shadow_skipped,
                                   pc.t.tri_exit_u, pc.t.tri_exit_v, pc.t.tri_full, pc.t.tri_accept, pc.t.sphere_hits};
#pragma unroll
    for (int k = 0; k < 14; k++) {
        const unsigned int s = __reduce_add_sync(mask, vals[k]);
        if (leader) atomicAdd(&counters->v[k], (unsigned long long)s);
    }
}

#endif
