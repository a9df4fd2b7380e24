/*
 * trace_global.cuh -- scene accessor over the HBM / L2-resident LBVH (any primitive count), shared by the plain
 * one-thread-per-pixel tracer (trace.cu) and the persistent tracer of big scenes (trace_big.cu).
 */
#ifndef RTB_TRACE_GLOBAL_CUH
#define RTB_TRACE_GLOBAL_CUH

#include "trace_pixel.cuh"

/* Shadow ray of light `li` against the light grid of a big scene (light_grid.cu): the ray's line passes through the
 * light, its direction from there is -to_light (to_light = L - P, the vector the caller normalised into d); the bin
 * of that direction lists every primitive such a line can hit.  Returns 1 = occluded, 0 = not occluded, -1 = this
 * light has no grid (the caller traverses the hierarchy).  Planes stay outside the grid and are tested here. */
template <bool COUNT>
__device__ __forceinline__ int rt_light_grid_occluded(const RtSceneDev& sc, const vec3& o, const vec3& d, float tmax,
                                                      const vec3& to_light, int li, RtLocalCounters& cnt) {
    /* static indices: a dynamic one would copy the kernel parameter block to local memory */
    static_assert(RT_LGB_LIGHTS == 2, "select below");
    const unsigned int* __restrict__ lg_off = li == 0 ? sc.lg_offsets[0] : (li == 1 ? sc.lg_offsets[1] : nullptr);
    if (!lg_off || __ldg(sc.lg_state + 4 * li) == 0u) return -1;     /* no grid, or its build overflowed */
    const float ax = fabsf(to_light.x), ay = fabsf(to_light.y), az = fabsf(to_light.z);
    const bool X = ax >= ay && ax >= az, Y = !X && ay >= az;
    const float m = X ? ax : (Y ? ay : az);
    if (!(m > 0.0f)) return -1;
    for (int p = 0; p < sc.n_planes; p++) {
        const float4 q0 = __ldg(sc.planes + 2 * p), q1 = __ldg(sc.planes + 2 * p + 1);
        float t;
        if (COUNT) cnt.plane_tests++;
        if (__float_as_uint(q0.w) & RT_TAG_EMISSIVE) continue;
        if (rt_hit_plane(o, d, f4xyz(q0), f4xyz(q1), t) && t < tmax) return 1;
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
            return 1;
    }
    return 0;
}

/* scene accessor over the HBM-resident LBVH (any primitive count); WIDE: over its 4-wide view */
template <bool BRUTE, bool COUNT, bool WIDE = false>
struct RtGlobalScene {
    const RtSceneDev& sc;
    __device__ __forceinline__ RtGlobalScene(const RtSceneDev& s) : sc(s) {}
#ifdef RT_NODEMAP
    unsigned int dbg[3] = {0, 0, 0};   /* debug variant: node tests of primary / bounce / shadow rays */
    __device__ __forceinline__ RtHit closest(const vec3& o, const vec3& d, int depth, bool, RtLocalCounters& cnt) {
        const unsigned int before = cnt.node_tests;
        const RtHit h = rt_closest<BRUTE, COUNT, WIDE>(sc, o, d, cnt);
        dbg[depth ? 1 : 0] += cnt.node_tests - before;
        return h;
    }
    __device__ __forceinline__ bool occluded(const vec3& o, const vec3& d, float tmax, const vec3&, int, RtLocalCounters& cnt) {
        const unsigned int before = cnt.node_tests;
        const bool r = rt_occluded<BRUTE, COUNT, WIDE>(sc, o, d, tmax, cnt);
        dbg[2] += cnt.node_tests - before;
        return r;
    }
#else
    __device__ __forceinline__ RtHit closest(const vec3& o, const vec3& d, int, bool, RtLocalCounters& cnt) {
        return rt_closest<BRUTE, COUNT, WIDE>(sc, o, d, cnt);
    }
    /* to_light = L - P (the vector the caller normalised into d), li = index of the light */
    __device__ __forceinline__ bool occluded(const vec3& o, const vec3& d, float tmax, const vec3& to_light, int li,
                                             RtLocalCounters& cnt) {
        if (!BRUTE) {
            const int g = rt_light_grid_occluded<COUNT>(sc, o, d, tmax, to_light, li, cnt);
            if (g >= 0) return g != 0;
        }
        return rt_occluded<BRUTE, COUNT, WIDE>(sc, o, d, tmax, cnt);
    }
#endif
    __device__ __forceinline__ vec3 normal(const RtHit& h, const vec3& P) { return rt_hit_normal<BRUTE>(sc, h, P); }
    __device__ __forceinline__ void material(unsigned int idx, float4& m0, float4& m1) {
        m0 = __ldg(sc.mats + 2 * idx);
        m1 = __ldg(sc.mats + 2 * idx + 1);
    }
};

#define RT_PACKET_STACK (RT_STACK_SIZE + 2)   /* per-warp stack of the packet traversal: one push per level at most */

/* Packet traversal: closest hits of the warp's 32 primary rays in one walk.  All 32 lanes call; `live` = this lane
 * has a ray.  `wstk` = the warp's stack in shared memory.  Every slab test is the per-axis-slack one
 * (rt_slab_robust): it is at least as tight as the common-slack test for every ray and needs no per-ray choice. */
template <bool COUNT>
__device__ __forceinline__ void rt_packet_closest(const RtSceneDev& sc, bool live, const vec3& o, const vec3& d, RtHit& best,
                                                  RtLocalCounters& cnt, int* __restrict__ wstk, int lane,
                                                  unsigned int* __restrict__ err) {
    if (sc.root == RT_NODE_SENTINEL) return;
    const RtRayBox rb = rt_ray_box(o, d);
    const float4* __restrict__ nodes = sc.nodes;
    int node = sc.root;                           /* warp-uniform from here on */
    int sp = 0;
    bool overflow = false;                        /* reported by one atomic BEHIND the loop (trace_common.cuh) */
    for (;;) {
        if (node >= 0) {
            const float4 n0 = __ldg(nodes + 4 * node);
            const float4 n1 = __ldg(nodes + 4 * node + 1);
            const float4 n2 = __ldg(nodes + 4 * node + 2);
            const float4 n3 = __ldg(nodes + 4 * node + 3);
            const float lim = fmaxf(best.t, 0.0f);
            float t0, t1;
            const bool h0 = live && rt_slab_robust(rb, n0.x, n0.y, n0.z, n0.w, n1.x, n1.y, lim, t0);
            const bool h1 = live && rt_slab_robust(rb, n1.z, n1.w, n2.x, n2.y, n2.z, n2.w, lim, t1);
            if (COUNT && live) cnt.node_tests += 2;
            const unsigned int b0 = __ballot_sync(0xffffffffu, h0), b1 = __ballot_sync(0xffffffffu, h1);
            const int c0 = __float_as_int(n3.x), c1 = __float_as_int(n3.y);
            if (b0 && b1) {
                /* the child most rays reach first goes first (rays that meet only one child vote for it) */
                const unsigned int first1 = __ballot_sync(0xffffffffu, h1 && (!h0 || t1 < t0));
                const bool one_first = 2 * __popc(first1) > __popc(b0 | b1);
                if (sp < RT_PACKET_STACK) {
                    if (lane == 0) wstk[sp] = one_first ? c0 : c1;
                    sp++;
                } else {
                    overflow = true;
                }
                node = one_first ? c1 : c0;
                continue;
            }
            if (b0) { node = c0; continue; }
            if (b1) { node = c1; continue; }
        } else {
            const int slot = ~node;
            float t;
            unsigned int tag;
            if (live && rt_test_record<COUNT>(sc.prims + 3 * (size_t)slot, o, d, t, tag, cnt)) rt_consider(best, t, tag, slot);
        }
        if (sp == 0) break;
        __syncwarp();                             /* lane 0's pushes are visible to everybody */
        sp--;
        node = wstk[sp];
        __syncwarp();                             /* ... and read before the slot is pushed over */
    }
    if (overflow && lane == 0) atomicOr(err, 1u);
}

/* RtGlobalScene whose primary-ray hit is handed in (packet traversal) */
template <bool BRUTE, bool COUNT, bool WIDE = false>
struct RtGlobalPacketScene : RtGlobalScene<BRUTE, COUNT, WIDE> {
    RtHit primary;
    bool have_primary;
    __device__ __forceinline__ RtGlobalPacketScene(const RtSceneDev& s) : RtGlobalScene<BRUTE, COUNT, WIDE>(s), have_primary(false) {}
    __device__ __forceinline__ RtHit closest(const vec3& o, const vec3& d, int depth, bool off, RtLocalCounters& cnt) {
        if (depth == 0 && have_primary) return primary;
        return RtGlobalScene<BRUTE, COUNT, WIDE>::closest(o, d, depth, off, cnt);
    }
};

#endif
