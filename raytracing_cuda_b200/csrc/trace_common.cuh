/*
 * trace_common.cuh -- per-ray device routines shared by the tracer kernels:
 * closest hit / any hit over the LBVH (or brute force), sky lookup, Phong
 * shading with hard shadows, pixel packing.  Restates reference
 * kernel.cu:131-225 (trace<depth>) with the linear object loops
 * (kernel.cu:144-151, 188-193) replaced by LBVH traversal.
 *
 * Closest hit keeps the reference's tie rule: strict '<' over increasing object
 * index (kernel.cu:145) == lexicographic minimum of (t, index); traversal order
 * therefore cannot change the winner.
 */
#ifndef RTB_TRACE_COMMON_CUH
#define RTB_TRACE_COMMON_CUH

#include "intersect.cuh"

struct RtHit {
    float t;
    unsigned int tag;   /* original index | flags; 0xFFFFFFFF = miss */
    int slot;           /* >=0: BVH slot / record index; <0: ~plane list index */
};

struct RtLocalCounters {
    unsigned int node_tests, tri_tests, sphere_tests, plane_tests;
    /* triangle tests by how far they got (rt_hit_triangle_tier; the rest of tri_tests was culled by det < 0.001) and
     * sphere tests that hit: the flop model of SURVEY 8d charges a test by its exit */
    unsigned int tri_exit_u, tri_exit_v, tri_full, tri_accept, sphere_hits;
};

/* instrumented leaf tests: the same arithmetic, plus the counts */
__device__ __forceinline__ bool rt_count_triangle(const vec3& o, const vec3& d, const vec3& v0, const vec3& e1,
                                                  const vec3& e2, float& t, RtLocalCounters& cnt) {
    int tier;
    const bool hit = rt_hit_triangle_tier(o, d, v0, e1, e2, t, tier);
    cnt.tri_tests++;
    cnt.tri_exit_u += tier == 1;
    cnt.tri_exit_v += tier == 2;
    cnt.tri_full += tier >= 3;
    cnt.tri_accept += tier == 4;
    return hit;
}
__device__ __forceinline__ bool rt_count_sphere(const vec3& o, const vec3& d, const vec3& c, float r2, float& t,
                                                RtLocalCounters& cnt) {
    const bool hit = rt_hit_sphere(o, d, c, r2, t);
    cnt.sphere_tests++;
    cnt.sphere_hits += hit;
    return hit;
}

#define RT_MISS 0xFFFFFFFFu

__device__ __forceinline__ vec3 f4xyz(const float4& q) { return v3(q.x, q.y, q.z); }

__device__ __forceinline__ void rt_consider(RtHit& best, float t, unsigned int tag, int slot) {
    unsigned int idx = tag & RT_TAG_INDEX;
    if (t < best.t || (t == best.t && idx < (best.tag & RT_TAG_INDEX))) {
        best.t = t;
        best.tag = tag;
        best.slot = slot;
    }
}

/* one leaf record (3 float4 at `rec`) against the ray; returns hit and t */
template <bool COUNT>
__device__ __forceinline__ bool rt_test_record(const float4* __restrict__ rec, const vec3& o, const vec3& d,
                                               float& t, unsigned int& tag, RtLocalCounters& cnt) {
    const float4 q0 = __ldg(rec);
    const float4 q1 = __ldg(rec + 1);
    tag = __float_as_uint(q0.w);
    if (tag & RT_TAG_SPHERE) {
        if (COUNT) return rt_count_sphere(o, d, f4xyz(q0), q1.y, t, cnt);
        return rt_hit_sphere(o, d, f4xyz(q0), q1.y, t);
    }
    if (tag & RT_TAG_PLANE) {
        if (COUNT) cnt.plane_tests++;
        return rt_hit_plane(o, d, f4xyz(q0), f4xyz(q1), t);
    }
    const float4 q2 = __ldg(rec + 2);
    if (COUNT) return rt_count_triangle(o, d, f4xyz(q0), f4xyz(q1), f4xyz(q2), t, cnt);
    return rt_hit_triangle(o, d, f4xyz(q0), f4xyz(q1), f4xyz(q2), t);
}

/*
 * ANY = false: closest hit into `best` (planes already folded in by the caller).
 * ANY = true : returns true at the first non-emissive primitive with t < tmax
 *              (reference kernel.cu:188-193; negative t counts, as in the reference).
 */
template <bool ANY, bool COUNT, bool ROBUST>
__device__ __forceinline__ bool rt_traverse_loop(const RtSceneDev& sc, const RtRayBox& rb, const vec3& o, const vec3& d,
                                                 RtHit& best, float tmax, RtLocalCounters& cnt) {
    int node = sc.root;
    int stack[RT_STACK_SIZE];
    int sp = 0;
    /* a push beyond the stack is reported, never silent (rt_get_device_errors) -- by ONE atomic behind the loop: the
     * same atomic inside it, never executed, cost the terrain frame 41 % (1.15 -> 1.62 ms) */
    bool overflow = false;
    const float4* __restrict__ nodes = sc.nodes;
    while (true) {
        if (node >= 0) {
            const float4 n0 = __ldg(nodes + 4 * node);
            const float4 n1 = __ldg(nodes + 4 * node + 1);
            const float4 n2 = __ldg(nodes + 4 * node + 2);
            const float4 n3 = __ldg(nodes + 4 * node + 3);
            const float lim = ANY ? tmax : fmaxf(best.t, 0.0f);
            float t0, t1;
            const bool h0 = ROBUST ? rt_slab_robust(rb, n0.x, n0.y, n0.z, n0.w, n1.x, n1.y, lim, t0)
                                   : rt_slab(rb, n0.x, n0.y, n0.z, n0.w, n1.x, n1.y, lim, t0);
            const bool h1 = ROBUST ? rt_slab_robust(rb, n1.z, n1.w, n2.x, n2.y, n2.z, n2.w, lim, t1)
                                   : rt_slab(rb, n1.z, n1.w, n2.x, n2.y, n2.z, n2.w, lim, t1);
            if (COUNT) cnt.node_tests += 2;
            int c0 = __float_as_int(n3.x), c1 = __float_as_int(n3.y);
            if (h0 && h1) {
                if (t1 < t0) { int tmp = c0; c0 = c1; c1 = tmp; }
                if (sp < RT_STACK_SIZE) stack[sp++] = c1;     /* (first 4 / 8 / 16 entries in shared memory: 1.18 vs 1.16 ms) */
                else overflow = true;
                node = c0;
                continue;
            }
            if (h0) { node = c0; continue; }
            if (h1) { node = c1; continue; }
        } else {
            const int slot = ~node;
            float t;
            unsigned int tag;
            if (rt_test_record<COUNT>(sc.prims + 3 * slot, o, d, t, tag, cnt)) {
                if (ANY) {
                    if (!(tag & RT_TAG_EMISSIVE) && t < tmax) return true;
                } else {
                    rt_consider(best, t, tag, slot);
                }
            }
        }
        if (sp == 0) break;
        node = stack[--sp];
    }
    if (overflow && sc.err) atomicOr(sc.err, 1u);
    return false;
}

template <bool ANY, bool COUNT>
__device__ __forceinline__ bool rt_traverse(const RtSceneDev& sc, const vec3& o, const vec3& d, RtHit& best,
                                            float tmax, RtLocalCounters& cnt) {
    if (sc.root == RT_NODE_SENTINEL) return false;
    const RtRayBox rb = rt_ray_box(o, d);
    /* rays with a near-zero direction component: per-axis slack (intersect.cuh) */
    if (rb.slack > RT_ROBUST_SLACK) return rt_traverse_loop<ANY, COUNT, true>(sc, rb, o, d, best, tmax, cnt);
    return rt_traverse_loop<ANY, COUNT, false>(sc, rb, o, d, best, tmax, cnt);
}

/* The same walk over the 4-wide view (RtSceneDev.wide): one fetch of 7 float4 brings the boxes of up to four
 * grandchildren; the ones the ray meets are visited nearest first (closest hit) or in any order (ANY).  The boxes are
 * the binary hierarchy's own and the leaf tests decide, so the result is the binary walk's, bit for bit. */
template <bool ANY, bool COUNT, bool ROBUST>
__device__ __forceinline__ bool rt_traverse_wide_loop(const RtSceneDev& sc, const RtRayBox& rb, const vec3& o, const vec3& d,
                                                      RtHit& best, float tmax, RtLocalCounters& cnt) {
    int node = sc.root;
    int stack[RT_STACK_SIZE];
    int sp = 0;
    bool overflow = false;                       /* reported behind the loop, see rt_traverse_loop */
    const float4* __restrict__ wide = sc.wide;
    while (true) {
        if (node >= 0) {
            const float4* __restrict__ w = wide + (size_t)RT_WIDE_F4 * node;
            const float4 a0 = __ldg(w), a1 = __ldg(w + 1), a2 = __ldg(w + 2), a3 = __ldg(w + 3), a4 = __ldg(w + 4),
                         a5 = __ldg(w + 5);
            const int4 ln = __ldg((const int4*)(w + 6));
            const float lim = ANY ? tmax : fmaxf(best.t, 0.0f);
            float t0, t1, t2, t3;
            bool h0 = ROBUST ? rt_slab_robust(rb, a0.x, a0.y, a0.z, a0.w, a1.x, a1.y, lim, t0)
                             : rt_slab(rb, a0.x, a0.y, a0.z, a0.w, a1.x, a1.y, lim, t0);
            bool h1 = ROBUST ? rt_slab_robust(rb, a1.z, a1.w, a2.x, a2.y, a2.z, a2.w, lim, t1)
                             : rt_slab(rb, a1.z, a1.w, a2.x, a2.y, a2.z, a2.w, lim, t1);
            bool h2 = ROBUST ? rt_slab_robust(rb, a3.x, a3.y, a3.z, a3.w, a4.x, a4.y, lim, t2)
                             : rt_slab(rb, a3.x, a3.y, a3.z, a3.w, a4.x, a4.y, lim, t2);
            bool h3 = ROBUST ? rt_slab_robust(rb, a4.z, a4.w, a5.x, a5.y, a5.z, a5.w, lim, t3)
                             : rt_slab(rb, a4.z, a4.w, a5.x, a5.y, a5.z, a5.w, lim, t3);
            h2 = h2 && ln.z != RT_NODE_SENTINEL;          /* slots 0 and 1 always hold a child */
            h3 = h3 && ln.w != RT_NODE_SENTINEL;
            if (COUNT) cnt.node_tests += 2 + (ln.z != RT_NODE_SENTINEL) + (ln.w != RT_NODE_SENTINEL);
            int c0 = ln.x, c1 = ln.y, c2 = ln.z, c3 = ln.w;
            const float inf = __int_as_float(0x7f800000);
            if (!h0) t0 = inf;
            if (!h1) t1 = inf;
            if (!h2) t2 = inf;
            if (!h3) t3 = inf;
#ifdef RT_WIDE_NEAREST_ONLY
            if (!ANY) {
                /* the nearest child to the front, the others in slot order */
#define RT_CSWAP(ta, ca, tb, cb) do { if (tb < ta) { const float tt = ta; ta = tb; tb = tt; const int cc = ca; ca = cb; cb = cc; } } while (0)
                RT_CSWAP(t0, c0, t1, c1);
                RT_CSWAP(t0, c0, t2, c2);
                RT_CSWAP(t0, c0, t3, c3);
#undef RT_CSWAP
            } else {
#else
            if (!ANY) {
                /* sort the four (t, link) pairs by t, misses (t = inf) last: 5 compare-exchanges */
#define RT_CSWAP(ta, ca, tb, cb) do { if (tb < ta) { const float tt = ta; ta = tb; tb = tt; const int cc = ca; ca = cb; cb = cc; } } while (0)
                RT_CSWAP(t0, c0, t1, c1);
                RT_CSWAP(t2, c2, t3, c3);
                RT_CSWAP(t0, c0, t2, c2);
                RT_CSWAP(t1, c1, t3, c3);
                RT_CSWAP(t1, c1, t2, c2);
#undef RT_CSWAP
            } else {
#endif
                /* any order: hits to the front */
                if (!(t0 < inf)) { if (t1 < inf) { t0 = t1; c0 = c1; t1 = inf; } else if (t2 < inf) { t0 = t2; c0 = c2; t2 = inf; } else if (t3 < inf) { t0 = t3; c0 = c3; t3 = inf; } }
            }
            if (t0 < inf) {
                const int more = (t1 < inf) + (t2 < inf) + (t3 < inf);
                if (sp + more <= RT_STACK_SIZE) {
                    if (t3 < inf) stack[sp++] = c3;            /* farthest first: the nearest pops first */
                    if (t2 < inf) stack[sp++] = c2;
                    if (t1 < inf) stack[sp++] = c1;
                } else {
                    overflow = true;
                }
                node = c0;
                continue;
            }
        } else {
            const int slot = ~node;
            float t;
            unsigned int tag;
            if (rt_test_record<COUNT>(sc.prims + 3 * (size_t)slot, o, d, t, tag, cnt)) {
                if (ANY) {
                    if (!(tag & RT_TAG_EMISSIVE) && t < tmax) return true;
                } else {
                    rt_consider(best, t, tag, slot);
                }
            }
        }
        if (sp == 0) break;
        node = stack[--sp];
    }
    if (overflow && sc.err) atomicOr(sc.err, 1u);
    return false;
}

template <bool ANY, bool COUNT>
__device__ __forceinline__ bool rt_traverse_wide(const RtSceneDev& sc, const vec3& o, const vec3& d, RtHit& best,
                                                 float tmax, RtLocalCounters& cnt) {
    if (sc.root == RT_NODE_SENTINEL) return false;
    const RtRayBox rb = rt_ray_box(o, d);
    if (rb.slack > RT_ROBUST_SLACK) return rt_traverse_wide_loop<ANY, COUNT, true>(sc, rb, o, d, best, tmax, cnt);
    return rt_traverse_wide_loop<ANY, COUNT, false>(sc, rb, o, d, best, tmax, cnt);
}

/* closest hit over planes + BVH (BRUTE: over all records in original order) */
template <bool BRUTE, bool COUNT, bool WIDE = false>
__device__ __forceinline__ RtHit rt_closest(const RtSceneDev& sc, const vec3& o, const vec3& d,
                                            RtLocalCounters& cnt) {
    RtHit best;
    best.t = __int_as_float(0x7f800000);
    best.tag = RT_MISS;
    best.slot = 0;
    if (BRUTE) {
        for (int i = 0; i < sc.n_objects; i++) {
            float t;
            unsigned int tag;
            if (rt_test_record<COUNT>(sc.recs + 3 * i, o, d, t, tag, cnt) && t < best.t) {
                best.t = t;
                best.tag = tag;
                best.slot = i;
            }
        }
        return best;
    }
    for (int p = 0; p < sc.n_planes; p++) {
        const float4 q0 = __ldg(sc.planes + 2 * p), q1 = __ldg(sc.planes + 2 * p + 1);
        float t;
        if (COUNT) cnt.plane_tests++;
        if (rt_hit_plane(o, d, f4xyz(q0), f4xyz(q1), t)) rt_consider(best, t, __float_as_uint(q0.w), ~p);
    }
    if (WIDE) rt_traverse_wide<false, COUNT>(sc, o, d, best, 0.0f, cnt);
    else rt_traverse<false, COUNT>(sc, o, d, best, 0.0f, cnt);
    return best;
}

/* reference kernel.cu:188-193 */
template <bool BRUTE, bool COUNT, bool WIDE = false>
__device__ __forceinline__ bool rt_occluded(const RtSceneDev& sc, const vec3& o, const vec3& d, float tmax,
                                            RtLocalCounters& cnt) {
    if (BRUTE) {
        for (int i = 0; i < sc.n_objects; i++) {
            float t;
            unsigned int tag;
            if (__float_as_uint(__ldg(&sc.recs[3 * i].w)) & RT_TAG_EMISSIVE) continue;
            if (rt_test_record<COUNT>(sc.recs + 3 * i, o, d, t, tag, cnt) && t < tmax) return true;
        }
        return false;
    }
    for (int p = 0; p < sc.n_planes; p++) {
        const float4 q0 = __ldg(sc.planes + 2 * p), q1 = __ldg(sc.planes + 2 * p + 1);
        float t;
        if (COUNT) cnt.plane_tests++;
        if (__float_as_uint(q0.w) & RT_TAG_EMISSIVE) continue;
        if (rt_hit_plane(o, d, f4xyz(q0), f4xyz(q1), t) && t < tmax) return true;
    }
    RtHit dummy;
    if (WIDE) return rt_traverse_wide<true, COUNT>(sc, o, d, dummy, tmax, cnt);
    return rt_traverse<true, COUNT>(sc, o, d, dummy, tmax, cnt);
}

/* surface normal at the hit (reference kernel.cu:66, 87, 123) */
template <bool BRUTE>
__device__ __forceinline__ vec3 rt_hit_normal(const RtSceneDev& sc, const RtHit& h, const vec3& P) {
    if (!BRUTE && h.slot < 0) return f4xyz(__ldg(sc.planes + 2 * (~h.slot) + 1));
    const float4* rec = (BRUTE ? sc.recs : sc.prims) + 3 * h.slot;
    if (h.tag & RT_TAG_PLANE) return f4xyz(__ldg(rec + 1));
    if (h.tag & RT_TAG_SPHERE) return rt_normalize(rt_sub(P, f4xyz(__ldg(rec))));
    return rt_normalize(rt_cross(f4xyz(__ldg(rec + 1)), f4xyz(__ldg(rec + 2))));
}

/* a / b in FP64 for a constant divisor b with rc = RN(1/b): q0 = a*rc, r = a - q0*b (exact in one FMA),
 * q = q0 + r*rc is the correctly rounded quotient (Markstein), i.e. bit-identical to the IEEE division the
 * reference executes -- 3 DFMA-class instructions instead of a ~30-instruction DDIV sequence.
 * tests/test_numerics.py checks the identity with exact rational arithmetic. */
__device__ __forceinline__ double rt_div_const(double a, double b, double rc) {
    const double q0 = a * rc;
    return __fma_rn(__fma_rn(-q0, b, a), rc, q0);
}

/* reference kernel.cu:154-164 with the truncating / wrapping uchar4 operators of structs.h:86-91.
 * Maps with weight 0 contribute (uchar)(v*0.0f) = 0, so skipping them is bit-identical.
 * fmodf(v, 1) == v - truncf(v) exactly (the difference of a float and its integer part is representable). */
__device__ __forceinline__ vec3 rt_sky(const RtFrameDev& f, const vec3& d) {
    const double PI = RTB_PI;
    const double TWO_PI = 2.0f * PI;
    float y = (float)(1 - rt_div_const(asinf(d.y) + PI / 2.0f, PI, 1.0 / PI));
    const float xv = (float)(rt_div_const(atan2f(d.x, d.z) + PI, TWO_PI, 1.0 / TWO_PI) + f.dayTime);
    float x = xv - truncf(xv);
    unsigned char r = 0, g = 0, b = 0;
#pragma unroll
    for (int k = 0; k < 4; k++) {
        const float w = f.skyVars[k];
        if (w != 0.0f || f.sky_all4) {
            const uchar4 v = tex2D<uchar4>(f.sky[k], x, y);
            r = (unsigned char)(r + (unsigned char)(v.x * w));
            g = (unsigned char)(g + (unsigned char)(v.y * w));
            b = (unsigned char)(b + (unsigned char)(v.z * w));
        }
    }
    return vscale(v3((float)r, (float)g, (float)b), 1.0f / 255.0f);
}

/* reference kernel.cu:26-32 */
__device__ __forceinline__ unsigned int rt_pack(float r, float g, float b) {
    r = clampf(r, 0.0f, 255.0f);
    g = clampf(g, 0.0f, 255.0f);
    b = clampf(b, 0.0f, 255.0f);
    return (unsigned int)(((int)b << 16) | ((int)g << 8) | (int)r);
}

/* reference kernel.cu:244-253, in the fused shape of the rebuilt reference's SASS */
__device__ __forceinline__ vec3 rt_primary_dir(const RtFrameDev& f, int x, int y) {
    const float partX = __fmul_rn((float)x, rt_rcp_approx((float)(f.imgw - 1)));
    const float partY = __fmul_rn((float)y, rt_rcp_approx((float)(f.imgh - 1)));
    vec3 t;
    {
        float vd = __fmaf_rn(partX, __fadd_rn(f.RD[0], -f.LD[0]), f.LD[0]);
        float vu = __fmaf_rn(partX, __fadd_rn(f.RU[0], -f.LU[0]), f.LU[0]);
        t.x = __fmaf_rn(-partY, __fadd_rn(vu, -vd), vu);
    }
    {
        float vd = __fmaf_rn(partX, __fadd_rn(f.RD[1], -f.LD[1]), f.LD[1]);
        float vu = __fmaf_rn(partX, __fadd_rn(f.RU[1], -f.LU[1]), f.LU[1]);
        t.y = __fmaf_rn(-partY, __fadd_rn(vu, -vd), vu);
    }
    {
        float vd = __fmaf_rn(partX, __fadd_rn(f.RD[2], -f.LD[2]), f.LD[2]);
        float vu = __fmaf_rn(partX, __fadd_rn(f.RU[2], -f.LU[2]), f.LU[2]);
        t.z = __fmaf_rn(-partY, __fadd_rn(vu, -vd), vu);
    }
    return rt_normalize(t);
}

/* does this shard have to trace pixel (x,y)?  own tile, or 1-pixel apron of an own tile */
__device__ __forceinline__ bool rt_shard_owns(const RtFrameDev& f, int x, int y) {
    return rt_tile_owner(x, y, f.tile_w, f.tile_h, f.n_shards, f.stride) == f.shard_id;
}
__device__ __forceinline__ bool rt_shard_needs(const RtFrameDev& f, int x, int y) {
    if (f.n_shards <= 1) return true;
    const int x0 = max(x - 1, 0), x1 = min(x + 1, f.imgw - 1);
    const int y0 = max(y - 1, 0), y1 = min(y + 1, f.imgh - 1);
    /* the 3x3 neighbourhood spans at most 2x2 tiles: test its corners */
    return rt_shard_owns(f, x0, y0) || rt_shard_owns(f, x1, y0) || rt_shard_owns(f, x0, y1) ||
           rt_shard_owns(f, x1, y1) || rt_shard_owns(f, x, y);
}

/* All 32 lanes of a warp call (lane = its lane id, (x, y) = its pixel of the patch).  Which pixels of the patch [x0, x0+8) x [y0, y0+4) does this shard trace (own tiles + the 1-pixel FXAA apron)?
 * Whole frames: all of them.  Shards whose tiles are at least 10x6 pixels: the patch grown by one pixel spans at most
 * 2x2 tiles, so four owner look-ups (lanes 0..3, one corner each) settle the common cases -- no corner owned: nothing to
 * trace; all four owned: everything; only a patch on a tile border asks per pixel. */
__device__ __forceinline__ bool rt_patch_pixel_needed(const RtFrameDev& f, int x0, int y0, int x, int y, int lane) {
    const bool inside = x < f.imgw && y < f.imgh;
    if (f.n_shards <= 1) return inside;
    if (f.tile_w >= 10 && f.tile_h >= 6) {
        const int cx = (lane & 1) ? min(x0 + 8, f.imgw - 1) : max(x0 - 1, 0);
        const int cy = (lane & 2) ? min(y0 + 4, f.imgh - 1) : max(y0 - 1, 0);
        const unsigned int own = __ballot_sync(0xffffffffu, lane < 4 && rt_shard_owns(f, cx, cy)) & 15u;
        if (own == 0u) return false;
        if (own == 15u) return inside;
    }
    return inside && rt_shard_needs(f, x, y);
}

#endif
