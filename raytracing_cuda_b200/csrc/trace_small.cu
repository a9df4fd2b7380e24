/*
 * trace_small.cu -- tracer for scenes whose whole LBVH fits in shared memory (the
 * reference's built-in scene: 133 objects, reference scene.h:11; here up to
 * RT_SMALL_MAX_PRIMS).  Same per-pixel arithmetic as trace.cu (trace_pixel.cuh), so the
 * two kernels agree bit for bit; only WHICH primitives get tested differs.
 *
 *   - rt_cull_tiles_kernel (binning pre-pass, one warp per 64x16-pixel bin, one CTA per 4x2
 *     block of bins): primary rays do not traverse the hierarchy at all.  The CTA culls every
 *     primitive against the frustum of its block (4 planes through the eye, back-facing
 *     triangles dropped -- they fail `det < 0.001`, kernel.cu:104), each warp the survivors
 *     against its bin, into a short, index-ordered candidate list in HBM (ballot compaction,
 *     64 B per bin); the same against the frustum mirrored in the sea plane (mirror bins).  For
 *     a screen-tile shard it also marks which 16x8 work tiles hold pixels to trace (own tiles +
 *     1-pixel FXAA apron) and appends them to a compact work list.  CTAs behind the screen bins
 *     fill the light bins when scene or lights have changed;
 *   - rt_trace_small_kernel: persistent CTAs (4 per SM); the LBVH nodes, leaf records,
 *     materials and planes are staged into shared memory once per CTA (the reference keeps
 *     its 133 objects in __constant__, kernel.cu:14).  After that every WARP is autonomous:
 *     it fetches single 8x4 patches from an atomic counter, the previous frame's most
 *     expensive patches first (order history; without one: bottom-up, so that the mirror
 *     chains in the lower half start early and the frame ends on cheap sky patches); no CTA
 *     barrier in the loop.  A primary ray tests just its bin's list, a shadow ray the <= 7
 *     occluders of its light's direction bin, the ray a primary ray sends off the sea plane its
 *     bin's mirror list;
 *   - the remaining bounces traverse the shared-memory LBVH with a per-thread stack of
 *     16-bit entries in shared memory (no local memory, sentinel at the bottom), child links
 *     pre-scaled to byte offsets so that a node fetch is one add + four LDS.128;
 *   - ray compaction (RT_OPT_COMPACTION 1; measured slower, off by default): a ray of the second or a later generation that has to TRAVERSE the hierarchy (no
 *     candidate list covers it) is not traced by the lane that spawned it.  The lane parks the ray -- origin,
 *     direction, pixel, generation and the colours waiting for it -- in a queue in the CTA's shared memory
 *     (ballot + popc slot assignment) and its warp goes on to the next patch; any warp of the CTA that finds 32
 *     rays waiting pops them and runs one generation of all of them as a FULL warp (traversal, shading, shadow
 *     rays; what bounces again is parked again).  Per-ray arithmetic is untouched, so pixels do not change.
 *
 * Conservative culling: the tile is grown by one pixel on every side and each plane test
 * gets a world-space pad (2^-13 of the largest coordinate in play), both orders of
 * magnitude above the rounding of the ray set-up and of the acceptance region of the leaf
 * tests; spheres are grown by the rounding noise of d2 = L.L - tca^2 (kernel.cu:55-58).
 * A primitive is only dropped when it cannot be hit by any ray of the tile, so the
 * (t, index) minimum over the candidates equals the reference's loop over all objects.
 */
#include "light_bins.cuh"
#include "frustum.cuh"

#define RT_SMALL_THREADS 256
#define RT_SMALL_STACK 23            /* per-thread stack entries (16-bit) in shared memory, above the sentinel */
#ifndef RT_SMALL_CTAS_PER_SM
#define RT_SMALL_CTAS_PER_SM 4
#endif
#define RT_SMALL_STACK_STRIDE (2 * RT_SMALL_THREADS)   /* bytes between two levels of one thread's stack */
/* light bins (see `light bins` below) */
#define RS_LG_G RT_LGRID_G
#define RS_LG_CELL 16                        /* bins per edge of the coarse cell one CTA handles */
#define RS_LG_CELLS (RS_LG_G / RS_LG_CELL)
#define RS_LG_SCALE RS_LG_SCALE_OF(RS_LG_G)
#define RS_LG_BLOCKS_PER_LIGHT (6 * RS_LG_CELLS * RS_LG_CELLS)

/* Shared-memory addressing: the hot loops use 32-bit shared-window addresses and explicit ld/st.shared, so a
 * node or leaf fetch is one add + LDS.128s with immediate offsets.  Node and leaf links are byte offsets from
 * the start of the staged scene (`base`): node i -> 64 i, leaf slot j -> ~(64 n_nodes + RS_LEAF j).
 * A staged leaf is RS_LEAF = 64 bytes: the three float4 of the record in HBM (rt_device.h) and, as the fourth, the unit
 * normal of a triangle -- normalize(cross(e1, e2)) (kernel.cu:123) depends on the triangle only, so every CTA computes
 * it once while it stages the scene, with the same inline functions the hit code used, and a shaded hit fetches it with
 * one LDS.128 at rec + 48 instead of ~35 instructions (cross, norm3df, correctly rounded reciprocal, three products). */
#define RS_LEAF 64
struct RtSmallShared {
    bool defer;              /* rays that would traverse the hierarchy are handed to a compacted batch (rt_trace_chain) */
    unsigned int base;       /* shared address of the staged nodes (leaf records follow them) */
    unsigned int stack0;     /* shared address of this thread's stack column */
    const unsigned short* list;   /* primary candidates of the current tile: [0] count, then leaf byte offsets */
    const unsigned short* mlist;  /* mirror bins: candidates of the rays primary rays of the tile send off plane 0 (NULL: off) */
    const uint4* lgrid;           /* light bins (NULL: shadow rays traverse the hierarchy) */
    int root, n_prims, n_planes, prims_off;
    bool deep;               /* hierarchy deeper than RT_SMALL_STACK: secondary rays loop linearly instead */
};

__device__ __forceinline__ float4 rs_lds128(unsigned int a) {
    float4 v;
    asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(a));
    return v;
}
__device__ __forceinline__ void rs_push(unsigned int& sp, int v) {
    asm volatile("st.shared.u16 [%0], %1;" ::"r"(sp), "h"((short)v));
    sp += RT_SMALL_STACK_STRIDE;
}
__device__ __forceinline__ int rs_pop(unsigned int& sp) {
    short v;
    sp -= RT_SMALL_STACK_STRIDE;
    asm volatile("ld.shared.s16 %0, [%1];" : "=h"(v) : "r"(sp));
    return (int)v;
}

/* leaf record at shared address `rec` */
__device__ __forceinline__ bool rs_test_record(unsigned int rec, const vec3& o, const vec3& d, float& t,
                                               unsigned int& tag, RtLocalCounters& cnt, bool count) {
    const float4 q0 = rs_lds128(rec);
    const float4 q1 = rs_lds128(rec + 16);
    tag = __float_as_uint(q0.w);
    if (tag & RT_TAG_SPHERE) {
        if (count) return rt_count_sphere(o, d, f4xyz(q0), q1.y, t, cnt);
        return rt_hit_sphere(o, d, f4xyz(q0), q1.y, t);
    }
    const float4 q2 = rs_lds128(rec + 32);
    if (count) return rt_count_triangle(o, d, f4xyz(q0), f4xyz(q1), f4xyz(q2), t, cnt);
    return rt_hit_triangle(o, d, f4xyz(q0), f4xyz(q1), f4xyz(q2), t);
}

#define RS_PLANES_BYTES 512            /* 16 planes x 2 float4, in front of the staged nodes */
#define RS_STACK_END (-32768)          /* sentinel at the bottom of every per-thread stack (never a valid link) */

/* LBVH traversal in shared memory.  Returns true on an occluder (ANY).  The per-thread stack holds
 * RT_SMALL_STACK entries above the sentinel; the caller only comes here when the hierarchy is no deeper
 * than that (RtSceneDev.depth), so pushes are unchecked. */
template <bool ANY, bool COUNT>
__device__ __forceinline__ bool rs_traverse(const RtSmallShared& s, const vec3& o, const vec3& d, RtHit& best,
                                            float tmax, RtLocalCounters& cnt) {
    if (s.n_prims == 0) return false;
    int node = s.root;
    const RtRayBox rb = rt_ray_box(o, d);
    const unsigned int base = s.base;
    unsigned int sp = s.stack0;
    rs_push(sp, RS_STACK_END);
    for (;;) {
        if (node >= 0) {
            const unsigned int a = base + (unsigned int)node;
            const float4 n0 = rs_lds128(a), n1 = rs_lds128(a + 16), n2 = rs_lds128(a + 32), n3 = rs_lds128(a + 48);
            const float lim = ANY ? tmax : fmaxf(best.t, 0.0f);
            float t0, t1;
            const bool h0 = rt_slab(rb, n0.x, n0.y, n0.z, n0.w, n1.x, n1.y, lim, t0);
            const bool h1 = rt_slab(rb, n1.z, n1.w, n2.x, n2.y, n2.z, n2.w, lim, t1);
            if (COUNT) cnt.node_tests += 2;
            const int c0 = __float_as_int(n3.x), c1 = __float_as_int(n3.y);
            if (h0 | h1) {
                const bool second_first = h1 && (!h0 || t1 < t0);
                const int near = second_first ? c1 : c0;
                if (h0 && h1) rs_push(sp, c0 ^ c1 ^ near);     /* measured faster than an unconditional store + select */
                node = near;
                continue;
            }
        } else {
            const unsigned int rec = base + (unsigned int)(~node);
            float t;
            unsigned int tag;
            if (rs_test_record(rec, o, d, t, tag, cnt, COUNT)) {
                if (ANY) {
                    if (!(tag & RT_TAG_EMISSIVE) && t < tmax) return true;
                } else {
                    rt_consider(best, t, tag, ~node);
                }
            }
        }
        node = rs_pop(sp);
        if (node == RS_STACK_END) break;
    }
    return false;
}

template <bool COUNT>
struct RtSmallScene {
    const RtSmallShared& s;
    __device__ __forceinline__ RtSmallScene(const RtSmallShared& sh) : s(sh) {}

    __device__ __forceinline__ void planes_closest(const vec3& o, const vec3& d, RtHit& best, RtLocalCounters& cnt) {
        for (int p = 0; p < s.n_planes; p++) {
            const float4 q0 = rs_lds128(s.base - RS_PLANES_BYTES + 32 * p), q1 = rs_lds128(s.base - RS_PLANES_BYTES + 32 * p + 16);
            float t;
            if (COUNT) cnt.plane_tests++;
            if (rt_hit_plane(o, d, f4xyz(q0), f4xyz(q1), t)) rt_consider(best, t, __float_as_uint(q0.w), ~p);
        }
    }
    __device__ __forceinline__ void linear_closest(const vec3& o, const vec3& d, RtHit& best, RtLocalCounters& cnt) {
        for (int j = 0; j < s.n_prims; j++) {
            const int off = s.prims_off + RS_LEAF * j;
            float t;
            unsigned int tag;
            if (rs_test_record(s.base + off, o, d, t, tag, cnt, COUNT)) rt_consider(best, t, tag, off);
        }
    }
    /* RtHit.slot: >= 0 leaf byte offset from `base`, < 0 ~plane index */
    __device__ __forceinline__ RtHit closest(const vec3& o, const vec3& d, int depth, bool off_plane0, RtLocalCounters& cnt) {
        RtHit best;
        best.t = __int_as_float(0x7f800000);
        best.tag = RT_MISS;
        best.slot = 0;
        planes_closest(o, d, best, cnt);
        /* primary ray: the tile's candidate list instead of the hierarchy; the ray a primary ray sends off plane 0:
         * the tile's mirror list (see `mirror bins`) */
        const unsigned short* cand = depth == 0 ? s.list : (off_plane0 ? s.mlist : nullptr);
        if (cand) {
            const int n = cand[0];
            if (n != 0xFFFF) {
                for (int i = 1; i <= n; i++) {
                    const int off = cand[i];
                    float t;
                    unsigned int tag;
                    if (rs_test_record(s.base + off, o, d, t, tag, cnt, COUNT)) rt_consider(best, t, tag, off);
                }
                return best;
            }
        }
        if (s.deep) linear_closest(o, d, best, cnt);     /* hierarchy deeper than the shared-memory stacks */
        else rs_traverse<false, COUNT>(s, o, d, best, 0.0f, cnt);
        return best;
    }
    /* to_light = L - P (the vector the caller normalised into d), li = index of the light */
    __device__ __forceinline__ bool occluded(const vec3& o, const vec3& d, float tmax, const vec3& to_light, int li,
                                             RtLocalCounters& cnt) {
        for (int p = 0; p < s.n_planes; p++) {
            const float4 q0 = rs_lds128(s.base - RS_PLANES_BYTES + 32 * p), q1 = rs_lds128(s.base - RS_PLANES_BYTES + 32 * p + 16);
            float t;
            if (COUNT) cnt.plane_tests++;
            if (__float_as_uint(q0.w) & RT_TAG_EMISSIVE) continue;
            if (rt_hit_plane(o, d, f4xyz(q0), f4xyz(q1), t) && t < tmax) return true;
        }
        if (s.lgrid && li < RT_LGRID_LIGHTS) {
            /* light bins: the ray's line passes through the light, its direction from there is -to_light */
            const float ax = fabsf(to_light.x), ay = fabsf(to_light.y), az = fabsf(to_light.z);
            const bool X = ax >= ay && ax >= az, Y = !X && ay >= az;
            const float m = X ? ax : (Y ? ay : az);
            const float qm = X ? to_light.x : (Y ? to_light.y : to_light.z);
            const float qa = X ? to_light.y : (Y ? to_light.z : to_light.x);
            const float qb = X ? to_light.z : (Y ? to_light.x : to_light.y);
            const int face = (X ? 0 : (Y ? 2 : 4)) + (qm > 0.0f ? 1 : 0);      /* direction = -to_light */
            const float k = -RS_LG_SCALE * rt_rcp_approx(m);
            const float gs = __fmaf_rn(qa, k, RS_LG_SCALE * (1.0f + RS_LG_DELTA));
            const float gt = __fmaf_rn(qb, k, RS_LG_SCALE * (1.0f + RS_LG_DELTA));
            const int ix = min(max((int)gs, 0), RS_LG_G - 1), iy = min(max((int)gt, 0), RS_LG_G - 1);
            uint4 e = __ldg(s.lgrid + ((li * 6 + face) * RS_LG_G + iy) * RS_LG_G + ix);
            unsigned int n = e.x & 0xFFFFu;
            if (n != 0xFFFFu && m > 0.0f) {
                for (; n; n--) {
                    e.x = __funnelshift_r(e.x, e.y, 16);
                    e.y = __funnelshift_r(e.y, e.z, 16);
                    e.z = __funnelshift_r(e.z, e.w, 16);
                    e.w >>= 16;
                    float t;
                    unsigned int tag;
                    if (rs_test_record(s.base + (e.x & 0xFFFFu), o, d, t, tag, cnt, COUNT) && t < tmax) return true;
                }
                return false;
            }
        }
        RtHit dummy;
        if (!s.deep) return rs_traverse<true, COUNT>(s, o, d, dummy, tmax, cnt);
        {
            for (int j = 0; j < s.n_prims; j++) {
                float t;
                unsigned int tag;
                if (rs_test_record(s.base + s.prims_off + RS_LEAF * j, o, d, t, tag, cnt, COUNT) &&
                    !(tag & RT_TAG_EMISSIVE) && t < tmax)
                    return true;
            }
        }
        return false;
    }
    /* rt_trace_chain<DEFER>: a ray of generation >= 1 that no candidate list covers (it would traverse) */
    __device__ __forceinline__ bool defers(int depth, bool off_plane0) const {
        return s.defer && depth >= 1 && !(off_plane0 && s.mlist && s.mlist[0] != 0xFFFF);
    }
    __device__ __forceinline__ vec3 normal(const RtHit& h, const vec3& P) {
        if (h.slot < 0) return f4xyz(rs_lds128(s.base - RS_PLANES_BYTES + 32 * (~h.slot) + 16));
        const unsigned int rec = s.base + (unsigned int)h.slot;
        if (h.tag & RT_TAG_SPHERE) return rt_normalize(rt_sub(P, f4xyz(rs_lds128(rec))));
        return f4xyz(rs_lds128(rec + 48));            /* staged with the leaf: see RS_LEAF */
    }
    __device__ __forceinline__ void material(unsigned int idx, float4& m0, float4& m1) {
        const unsigned int a = s.base + (unsigned int)(s.prims_off + RS_LEAF * s.n_prims) + 32u * idx;
        m0 = rs_lds128(a);
        m1 = rs_lds128(a + 16);
    }
};

/* Programmatic dependent launch (griddepcontrol): a kernel launched with the programmatic-serialization attribute
 * may start while its predecessor on the stream still runs.  wait = the predecessor has completed and its
 * writes are visible (no-op without the attribute); launch = the successor may be launched. */
__device__ __forceinline__ void rs_griddep_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void rs_griddep_launch() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }

#define RT_SMALL_WARPS (RT_SMALL_THREADS / 32)
#define RT_SMALL_TILE_W 16
#define RT_SMALL_TILE_H 8
#define RT_SMALL_PATCHES 4            /* 8x4 patches per tile */
#define RS_CLASSES RT_SMALL_ORDER_CLASSES   /* class c = floor(log2(cycles / 1024)), clamped */
#define RS_ORDER_HEAD RT_SMALL_ORDER_HEAD              /* uints before the class lists of an order buffer: [0..7] class counts */
#define RS_CNT_TILES 32               /* tile_counter[]: [0] next fetch, [1] finished warps, [32] listed tiles */
#define RT_SMALL_LIST 32              /* ushorts per bin list: [0] = count (0xFFFF: too many, traverse), [1..31] */
/* candidate lists are kept per BIN of RT_SMALL_BIN_W x RT_SMALL_BIN_H pixels (whole tiles) */
#ifndef RT_SMALL_BIN_W
#define RT_SMALL_BIN_W 64
#endif
#ifndef RT_SMALL_BIN_H
#define RT_SMALL_BIN_H 16
#endif

/* ---------------------------------------------------------------- binning pre-pass */

#define RS_LG_PRIM_F4 (sizeof(RsLgPrim) / 16)

/* one CTA (256 threads) = one coarse cell of 16x16 bins of one face of one light; a warp = 8x4 bins of it.
 * Three levels: CTA (every primitive against the cell), warp (the cell's survivors against its 8x4 bins), lane
 * (the warp's survivors against its bin). */
__device__ __forceinline__ void rs_light_cell(const RtSceneDev& sc, const RtFrameDev& f, uint4* __restrict__ lgrid,
                                              int cell_id) {
    __shared__ float4 s_prim[RT_SMALL_MAX_PRIMS * RS_LG_PRIM_F4];
    __shared__ unsigned short s_slot[RT_SMALL_MAX_PRIMS];
    __shared__ unsigned short s_warp[8][RT_SMALL_MAX_PRIMS];
    __shared__ unsigned int s_mask[8][RT_SMALL_MAX_PRIMS];     /* per (warp, survivor): the bins of the warp's block it can hit */
    __shared__ int s_n, s_nw[8];
    const int li = cell_id / RS_LG_BLOCKS_PER_LIGHT, rr = cell_id - li * RS_LG_BLOCKS_PER_LIGHT;
    const int face = rr / (RS_LG_CELLS * RS_LG_CELLS), cell = rr - face * (RS_LG_CELLS * RS_LG_CELLS);
    const int cy = cell / RS_LG_CELLS, cx = cell - cy * RS_LG_CELLS;
    const int n = sc.n_bvh_prims;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
#ifdef RS_LG_TIMING
    const long long lg_t0 = clock64();
#endif
    const float lpos[3] = {f.lights[li].px, f.lights[li].py, f.lights[li].pz};
    const vec3 lp = v3(lpos[0], lpos[1], lpos[2]);
    const float pad = rs_cull_pad(sc, lpos);
    if (threadIdx.x == 0) s_n = 0;
    __syncthreads();
    {
        const RsLgRect rc = rs_lg_rect<RS_LG_G>((float)(cx * RS_LG_CELL), (float)(cy * RS_LG_CELL),
                                       (float)(cx * RS_LG_CELL + RS_LG_CELL), (float)(cy * RS_LG_CELL + RS_LG_CELL));
        for (int j = threadIdx.x; j < n; j += blockDim.x) {
            union { RsLgPrim p; float4 q[RS_LG_PRIM_F4]; } u;
            if (rs_lg_prim(face, lp, sc.prims + 3 * j, pad, u.p) && rs_lg_may_hit(u.p, rc, pad)) {
                const int at = atomicAdd(&s_n, 1);             /* any order: the lists serve an any-hit test */
#pragma unroll
                for (int k = 0; k < (int)RS_LG_PRIM_F4; k++) s_prim[at * RS_LG_PRIM_F4 + k] = u.q[k];
                s_slot[at] = (unsigned short)j;
            }
        }
    }
    __syncthreads();
    const int nc = s_n;
    const int wx = cx * RS_LG_CELL + 8 * (warp & 1), wy = cy * RS_LG_CELL + 4 * (warp >> 1);
    int nw = 0;
    if (nc) {
        const RsLgRect rw = rs_lg_rect<RS_LG_G>((float)wx, (float)wy, (float)(wx + 8), (float)(wy + 4));
        for (int i0 = 0; i0 < nc; i0 += 32) {
            const int i = i0 + lane;
            bool keep = false;
            if (i < nc) {
                union { RsLgPrim p; float4 q[RS_LG_PRIM_F4]; } u;
#pragma unroll
                for (int k = 0; k < (int)RS_LG_PRIM_F4; k++) u.q[k] = s_prim[i * RS_LG_PRIM_F4 + k];
                keep = rs_lg_may_hit(u.p, rw, pad);
            }
            const unsigned int mk = __ballot_sync(0xffffffffu, keep);
            if (keep) s_warp[warp][nw + __popc(mk & ((1u << lane) - 1u))] = (unsigned short)i;
            nw += __popc(mk);
        }
        __syncwarp();
    }
    /* Level 3, shared by the whole CTA: the (warp, survivor) pairs of all eight warps are dealt round-robin, lane =
     * bin of that warp's 8x4 block, one ballot per pair.  (Each warp walking its own survivors left the cell to its
     * densest warp: 24..52 survivors at ~800 cycles each for a lone warp, 12..20 us -- the whole pre-pass.) */
    if (lane == 0) s_nw[warp] = nw;
    __syncthreads();
    {
        int pw = 0, pbase = 0, pend = s_nw[0];          /* pairs [pbase, pend) belong to warp pw */
        int total = 0;
#pragma unroll
        for (int q = 0; q < 8; q++) total += s_nw[q];
        RsLgRect rb;
        int have_rect = -1;
        for (int pr = warp; pr < total; pr += 8) {
            while (pr >= pend) { pw++; pbase = pend; pend += s_nw[pw]; }
            if (have_rect != pw) {
                const int bx = cx * RS_LG_CELL + 8 * (pw & 1) + (lane & 7), by = cy * RS_LG_CELL + 4 * (pw >> 1) + (lane >> 3);
                rb = rs_lg_rect<RS_LG_G>((float)bx, (float)by, (float)(bx + 1), (float)(by + 1));
                have_rect = pw;
            }
            const int i = s_warp[pw][pr - pbase];
            union { RsLgPrim p; float4 q[RS_LG_PRIM_F4]; } u;
#pragma unroll
            for (int k = 0; k < (int)RS_LG_PRIM_F4; k++) u.q[k] = s_prim[i * RS_LG_PRIM_F4 + k];
            const unsigned int hit = __ballot_sync(0xffffffffu, rs_lg_may_hit(u.p, rb, pad));
            if (lane == 0) s_mask[pw][pr - pbase] = hit;
        }
    }
    __syncthreads();
    const int gx = wx + (lane & 7), gy = wy + (lane >> 3);
    unsigned int w[4] = {0u, 0u, 0u, 0u};
    int count = 0;
    {
        const int prims_off = 64 * (n > 1 ? n - 1 : 0);
        for (int k0 = 0; k0 < nw; k0++) {
            if (!((s_mask[warp][k0] >> lane) & 1u)) continue;
            count++;
            if (count < 8) {
                const unsigned int off = (unsigned int)(prims_off + RS_LEAF * (int)s_slot[s_warp[warp][k0]]);
#pragma unroll
                for (int k = 1; k < 8; k++)                  /* static register indices */
                    if (k == count) w[k >> 1] |= off << (16 * (k & 1));
            }
        }
    }
    w[0] |= count < 8 ? (unsigned int)count : 0xFFFFu;
    lgrid[((size_t)(li * 6 + face) * RS_LG_G + gy) * RS_LG_G + gx] = make_uint4(w[0], w[1], w[2], w[3]);
#ifdef RS_LG_TIMING
    /* debug variant: (cycles, cell survivors, this warp's survivors) of warp w of cell c into the last light's bins */
    if (lane == 0)
        lgrid[(size_t)(RT_LGRID_LIGHTS - 1) * 6 * RS_LG_G * RS_LG_G + cell_id * 8 + warp] =
            make_uint4((unsigned int)(clock64() - lg_t0), (unsigned int)nc, (unsigned int)nw, 0u);
#endif
}

/* ---- mirror bins.  A primary ray that hits plane 0 (the reference's sea, scene.cpp:183-190) at P continues as
 * P + 0.001 r, r = normalize(d - 2 (N.d) N) (kernel.cu:212-213).  With a unit normal that is the mirror image
 * of the primary ray: its line passes through the eye's image E' behind the plane with the primary direction's
 * image, so the rays a tile sends off the plane lie in the tile's frustum mirrored, apex E' -- the same cull as
 * the primary one with the mirrored corners, plus: nothing wholly on E's far side of the plane can be met (the ray
 * starts on the plane and leaves it towards the eye's side).  Rounding: P sits within |t| 2^-22 (< 1e-3 world
 * units) of the exact point and r within 2^-23 of the exact direction, both far inside the pad (2^-13 of the
 * largest coordinate) every plane test here carries.  The exact leaf tests decide; bit-identical to the traversal. */
struct RsMirror {
    bool on;            /* plane 0 is a unit-normal, non-emissive mirror */
    vec3 pn, p0, eye_m; /* its normal, a point on it, the eye's image */
    float pad_m, side;  /* side: +1 / -1 = the eye's side of the plane, 0 = eye (nearly) in the plane: no side cull */
};

__device__ __forceinline__ RsMirror rs_mirror_of(const RtSceneDev& sc, const vec3& eye, float pad, bool want) {
    RsMirror m;
    m.on = false;
    m.pn = m.p0 = v3(0, 0, 0);
    float hgt = 0.0f;
    if (want && sc.n_planes > 0) {
        const float4 q0 = __ldg(sc.planes), q1 = __ldg(sc.planes + 1);
        const unsigned int tag = __float_as_uint(q0.w);
        m.pn = f4xyz(q1);
        m.p0 = f4xyz(q0);
        const float mirror = __ldg(sc.mats + 2 * (tag & RT_TAG_INDEX)).w;
        hgt = vdot(vsub(eye, m.p0), m.pn);
        m.on = mirror > 0.0f && !(tag & RT_TAG_EMISSIVE) && fabsf(vdot(m.pn, m.pn) - 1.0f) <= 1e-6f;
    }
    m.eye_m = vsub(eye, vscale(m.pn, 2.0f * hgt));
    const float apex_m[3] = {m.eye_m.x, m.eye_m.y, m.eye_m.z};
    m.pad_m = fmaxf(pad, rs_cull_pad(sc, apex_m));
    m.side = hgt > m.pad_m ? 1.0f : (hgt < -m.pad_m ? -1.0f : 0.0f);
    return m;
}

__device__ __forceinline__ RsFrustum rs_mirror_frustum(const RsMirror& m, const vec3* corner) {
    vec3 corner_m[4];
#pragma unroll
    for (int k = 0; k < 4; k++) corner_m[k] = vsub(corner[k], vscale(m.pn, 2.0f * vdot(corner[k], m.pn)));
    return rs_frustum_from_corners(corner_m);
}

/* One CTA = a block of RS_CB_W x RS_CB_H bins (256 x 32 pixels), one warp per bin.  Two levels: the CTA tests every
 * primitive (one per thread) against the frusta of the whole BLOCK and keeps the survivors, in index order, in shared
 * memory; each warp then tests only those against its bin.  A warp walking all primitives for both of its lists ran
 * ~1,900 instructions in series (3.9 M warp instructions per 1080p frame, ncu) -- the pre-pass is that chain, and the
 * tracer behind it waits for its last CTA; most of a block's primitives fail the block test already. */
#define RS_CB_W 4
#define RS_CB_H 2

__device__ __forceinline__ void
rs_cull_block(const RtSceneDev& sc, const RtFrameDev& f, unsigned short* __restrict__ tile_lists, int bins_x, int bins_y,
              unsigned int* __restrict__ work_list, unsigned int* __restrict__ work_count, int work_tiles_x,
              int work_tiles_y, unsigned short* __restrict__ mirror_lists, int block_id) {
    /* work_list = NULL for whole frames; mirror_lists = NULL: no mirror bins */
    __shared__ unsigned short s_surv[2][RT_SMALL_MAX_PRIMS];   /* [0] primary, [1] mirror survivors of the block */
    __shared__ unsigned int s_ball[2][8];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int n = sc.n_bvh_prims, n_nodes = n > 1 ? n - 1 : 0;
    const int prims_off = 64 * n_nodes;
    const int cbx = (bins_x + RS_CB_W - 1) / RS_CB_W;
    const int cy = block_id / cbx, cx = block_id - cy * cbx;
    const vec3 eye = v3(f.pos[0], f.pos[1], f.pos[2]);
    const float pad = rs_cull_pad(sc, f.pos);
    const RsMirror mir = rs_mirror_of(sc, eye, pad, mirror_lists != nullptr);
    {
        const int x0 = cx * RS_CB_W * RT_SMALL_BIN_W, y0 = cy * RS_CB_H * RT_SMALL_BIN_H;
        const int x1 = min(x0 + RS_CB_W * RT_SMALL_BIN_W - 1, f.imgw - 1), y1 = min(y0 + RS_CB_H * RT_SMALL_BIN_H - 1, f.imgh - 1);
        vec3 corner[4];
        rs_tile_corners(f, x0, y0, x1, y1, corner);
        const RsFrustum fb = rs_frustum_from_corners(corner);
        const RsFrustum fmb = rs_mirror_frustum(mir, corner);
        static_assert(RT_SMALL_MAX_PRIMS <= 256, "one primitive per thread of the 256-thread CTA");
        const int j = threadIdx.x;                /* n <= RT_SMALL_MAX_PRIMS = the CTA's threads */
        bool kp = false, km = false;
        if (j < n) {
            const float4 r0 = __ldg(sc.prims + 3 * j), r1 = __ldg(sc.prims + 3 * j + 1), r2 = __ldg(sc.prims + 3 * j + 2);
            kp = !fb.ok || rs_tile_may_hit(fb, eye, r0, r1, r2, pad);
            if (mir.on) {
                km = !(fb.ok && fmb.ok) || rs_tile_may_hit(fmb, mir.eye_m, r0, r1, r2, mir.pad_m);
                if (km && mir.side != 0.0f) {
                    const float ha = mir.side * vdot(vsub(f4xyz(r0), mir.p0), mir.pn);
                    if (__float_as_uint(r0.w) & RT_TAG_SPHERE) {
                        km = !(ha < -(fabsf(r1.x) + mir.pad_m));
                    } else {
                        const float hb = ha + mir.side * vdot(f4xyz(r1), mir.pn), hc = ha + mir.side * vdot(f4xyz(r2), mir.pn);
                        km = !(ha < -mir.pad_m && hb < -mir.pad_m && hc < -mir.pad_m);
                    }
                }
            }
        }
        const unsigned int bp = __ballot_sync(0xffffffffu, kp), bm = __ballot_sync(0xffffffffu, km);
        if (lane == 0) { s_ball[0][warp] = bp; s_ball[1][warp] = bm; }
        __syncthreads();
        int op = 0, om = 0;
#pragma unroll
        for (int w = 0; w < 8; w++)
            if (w < warp) { op += __popc(s_ball[0][w]); om += __popc(s_ball[1][w]); }
        const unsigned int below = (1u << lane) - 1u;
        if (kp) s_surv[0][op + __popc(bp & below)] = (unsigned short)j;
        if (km) s_surv[1][om + __popc(bm & below)] = (unsigned short)j;
        __syncthreads();
    }
    int ns0 = 0, ns1 = 0;
#pragma unroll
    for (int w = 0; w < 8; w++) { ns0 += __popc(s_ball[0][w]); ns1 += __popc(s_ball[1][w]); }

    const int bbx = cx * RS_CB_W + warp % RS_CB_W, bby = cy * RS_CB_H + warp / RS_CB_W;
    if (bbx >= bins_x || bby >= bins_y) return;
    const int tile = bby * bins_x + bbx;
    /* The work tiles (16x8) of this bin that hold at least one pixel to trace: all of them, or with a
     * screen-tile shard those with own pixels or pixels of the 1-pixel FXAA apron. */
    const int wx = bbx * (RT_SMALL_BIN_W / RT_SMALL_TILE_W) + lane % (RT_SMALL_BIN_W / RT_SMALL_TILE_W);
    const int wy = bby * (RT_SMALL_BIN_H / RT_SMALL_TILE_H) + lane / (RT_SMALL_BIN_W / RT_SMALL_TILE_W);
    bool need = lane < (RT_SMALL_BIN_W / RT_SMALL_TILE_W) * (RT_SMALL_BIN_H / RT_SMALL_TILE_H) && wx < work_tiles_x &&
                wy < work_tiles_y;
    if (need && f.n_shards > 1) {
        /* the work tile grown by one pixel, in shard-tile coordinates */
        const int ax = max(wx * RT_SMALL_TILE_W - 1, 0) / f.tile_w;
        const int bx2 = min(wx * RT_SMALL_TILE_W + RT_SMALL_TILE_W, f.imgw - 1) / f.tile_w;
        const int ay = max(wy * RT_SMALL_TILE_H - 1, 0) / f.tile_h;
        const int by2 = min(wy * RT_SMALL_TILE_H + RT_SMALL_TILE_H, f.imgh - 1) / f.tile_h;
        need = false;
        for (int ty = ay; ty <= by2 && !need; ty++)
            for (int tx = ax; tx <= bx2; tx++)
                if ((tx + ty * f.stride) % f.n_shards == f.shard_id) { need = true; break; }
    }
    const unsigned int need_mask = __ballot_sync(0xffffffffu, need);
    if (work_list && need_mask) {                /* sharded frames: append this bin's work tiles to the list */
        unsigned int at = 0;
        if (lane == 0) at = atomicAdd(work_count, (unsigned int)__popc(need_mask));
        at = __shfl_sync(0xffffffffu, at, 0) + __popc(need_mask & ((1u << lane) - 1u));
        if (need) work_list[at] = (unsigned int)(wy * work_tiles_x + wx);
    }
    if (need_mask == 0u) return;                 /* nothing of this bin is traced here: no list needed either */
    const int x0 = bbx * RT_SMALL_BIN_W, y0 = bby * RT_SMALL_BIN_H;
    const int x1 = min(x0 + RT_SMALL_BIN_W - 1, f.imgw - 1), y1 = min(y0 + RT_SMALL_BIN_H - 1, f.imgh - 1);
    vec3 corner[4];
    rs_tile_corners(f, x0, y0, x1, y1, corner);
    const RsFrustum fr = rs_frustum_from_corners(corner);
    unsigned short* __restrict__ out = tile_lists + (size_t)tile * RT_SMALL_LIST;
    int count = 0;
    for (int i0 = 0; i0 < ns0; i0 += 32) {
        const int i = i0 + lane;
        const int j = s_surv[0][min(i, ns0 - 1)];
        const bool keep = i < ns0 && (!fr.ok || rs_tile_may_hit(fr, eye, sc.prims + 3 * j, pad));
        const unsigned int m = __ballot_sync(0xffffffffu, keep);
        const int at = count + __popc(m & ((1u << lane) - 1u));
        if (keep && at < RT_SMALL_LIST - 1) out[1 + at] = (unsigned short)(prims_off + RS_LEAF * j);
        count += __popc(m);
    }
    if (lane == 0) out[0] = count < RT_SMALL_LIST ? (unsigned short)count : (unsigned short)0xFFFFu;
    if (!mirror_lists) return;
    unsigned short* __restrict__ mout = mirror_lists + (size_t)tile * RT_SMALL_LIST;
    if (!(mir.on && fr.ok)) {
        if (lane == 0) mout[0] = (unsigned short)0xFFFFu;       /* traverse */
        return;
    }
    const RsFrustum frm = rs_mirror_frustum(mir, corner);
    int mcount = 0;
    for (int i0 = 0; i0 < ns1; i0 += 32) {
        const int i = i0 + lane;
        const int j = s_surv[1][min(i, ns1 - 1)];
        const bool mkeep = i < ns1 && (!frm.ok || rs_tile_may_hit(frm, mir.eye_m, sc.prims + 3 * j, mir.pad_m));
        const unsigned int mm = __ballot_sync(0xffffffffu, mkeep);
        const int mat = mcount + __popc(mm & ((1u << lane) - 1u));
        if (mkeep && mat < RT_SMALL_LIST - 1) mout[1 + mat] = (unsigned short)(prims_off + RS_LEAF * j);
        mcount += __popc(mm);
    }
    if (lane == 0) mout[0] = mcount < RT_SMALL_LIST ? (unsigned short)mcount : (unsigned short)0xFFFFu;
}

#ifdef RS_TIMELINE
/* debug variant (tools/build_variant.sh tl -DRS_TIMELINE): %globaltimer stamps of the last 8 frames' kernels.
 * [slot][0] pre-pass first entry  [1] pre-pass last CTA done (before its final wait)  [2] tracer first CTA entry
 * [3] tracer first / [4] last CTA past griddepcontrol.wait  [5] last unit done  [6] last warp exit */
__device__ unsigned long long g_tl_frame[8][8];
__device__ unsigned int g_tl_frame_no;
__device__ __forceinline__ unsigned long long tl_now() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}
__device__ __forceinline__ unsigned long long* tl_slot() { return g_tl_frame[*(volatile unsigned int*)&g_tl_frame_no & 7u]; }
extern "C" cudaError_t rt_debug_timeline_small(unsigned long long* out, int reset) {
    cudaError_t e = cudaMemcpyFromSymbol(out, g_tl_frame, sizeof(g_tl_frame));
    if (e == cudaSuccess && reset) {
        unsigned long long init[8][8];
        for (int s = 0; s < 8; s++)
            for (int k = 0; k < 8; k++) init[s][k] = (k == 0 || k == 2 || k == 3) ? ~0ull : 0ull;
        const unsigned int zero = 0;
        e = cudaMemcpyToSymbol(g_tl_frame, init, sizeof(init));
        if (e == cudaSuccess) e = cudaMemcpyToSymbol(g_tl_frame_no, &zero, sizeof(zero));
    }
    return e;
}
#endif

/* The pre-pass reads the scene and writes lists nobody else is using, so -- when the host knows that the
 * scene was complete before the previous kernel on the stream (wait_first = 0: that kernel is the FXAA / copy
 * of this context's previous frame) -- it runs WHILE that kernel finishes.  Its last act is to wait for it:
 * the tracer, which waits for this kernel, then never overwrites the base image under a running FXAA. */
/* (Round 2 measured a second instantiation without the light cells -- no shared memory, 128-thread CTAs of 7 K
 * registers that become resident beside the previous frame's FXAA CTAs instead of queueing behind them: the pre-pass
 * alone got 2 us faster and the bench step 0.5 % SLOWER, 53.97 vs 54.25 Grays/s; profiles/r02_frame_timeline.txt.) */
__global__ void __launch_bounds__(256, 4)
rt_cull_tiles_kernel(const RtSceneDev sc, const RtFrameDev f, unsigned short* __restrict__ tile_lists, int bins_x,
                     int bins_y, unsigned int* __restrict__ work_list, unsigned int* __restrict__ work_count,
                     int work_tiles_x, int work_tiles_y, int wait_first, uint4* __restrict__ lgrid,
                     int n_cull_blocks, unsigned short* __restrict__ mirror_lists) {
#ifdef RS_TIMELINE
    if (threadIdx.x == 0) atomicMin(tl_slot() + 0, tl_now());
#endif
    if (wait_first) rs_griddep_wait();
    rs_griddep_launch();                         /* the tracer may start staging the scene */
    /* the CTAs behind the screen bins fill the light bins (lgrid != NULL), one coarse cell each */
    if ((int)blockIdx.x >= n_cull_blocks) rs_light_cell(sc, f, lgrid, (int)blockIdx.x - n_cull_blocks);
    else rs_cull_block(sc, f, tile_lists, bins_x, bins_y, work_list, work_count, work_tiles_x, work_tiles_y, mirror_lists,
                       (int)blockIdx.x);
#ifdef RS_TIMELINE
    __syncthreads();
    if (threadIdx.x == 0) atomicMax(tl_slot() + 1, tl_now());
#endif
    /* Every CTA waits, not just one: with the others gone early the pre-pass ends 4 us sooner at 3840x2160, but the
     * tracer's CTAs then find the SMs' shared memory in pieces more often (7..12 of 592 become resident only at
     * the frame's end) and the frame gets 0.2 % slower (profiles/r02_frame_timeline.txt). */
    rs_griddep_wait();
}

/* ---------------------------------------------------------------- ray queue (compaction)
 * RS_Q_CAP parked rays per CTA, structure of arrays of float4: [0] (o.xyz, pixel index) [1] (d.xyz, generation)
 * [2..5] the colours (phong.rgb, kR) of generations 0..3 waiting for the ray's result (s_pending of the lane that
 * parked it).  A stack under a CTA-wide spin lock taken by whole warps: pushes and pops are a few shared-memory
 * accesses per lane, so the lock is held for a few hundred cycles at most.  qctl[0] = lock, qctl[1] = count. */
#define RS_Q_CAP 64
#define RS_Q_FIELDS (2 + RTB_MAX_DEPTH)

__device__ __forceinline__ void rs_q_lock(unsigned int* qctl, int lane) {
    if (lane == 0)
        while (atomicCAS(qctl, 0u, 1u) != 0u) __nanosleep(32);
    __syncwarp();
    __threadfence_block();
}
__device__ __forceinline__ void rs_q_unlock(unsigned int* qctl, int lane) {
    __threadfence_block();
    __syncwarp();
    if (lane == 0) atomicExch(qctl, 0u);
}

/* All 32 lanes call.  Lanes with `parked` offer their ray; as many as fit are stored (lowest lanes first).  Returns
 * whether THIS lane's ray was stored; a lane that offered and got false keeps its ray. */
__device__ __forceinline__ bool rs_q_push(float4* __restrict__ q, unsigned int* qctl, int lane, bool parked,
                                          const RtRay& ray, int pixel, const float4* __restrict__ s_pending, int tid) {
    const unsigned int m = __ballot_sync(0xffffffffu, parked);
    if (m == 0u) return false;
    const int rank = __popc(m & ((1u << lane) - 1u));
    rs_q_lock(qctl, lane);
    int base = 0, take = 0;
    if (lane == 0) {
        const int cnt = (int)((volatile unsigned int*)qctl)[1];
        take = min(__popc(m), RS_Q_CAP - cnt);
        base = cnt;
        ((volatile unsigned int*)qctl)[1] = (unsigned int)(cnt + take);
    }
    base = __shfl_sync(0xffffffffu, base, 0);
    take = __shfl_sync(0xffffffffu, take, 0);
    const bool stored = parked && rank < take;
    if (stored) {
        const int slot = base + rank;
        q[slot] = make_float4(ray.o.x, ray.o.y, ray.o.z, __int_as_float(pixel));
        q[RS_Q_CAP + slot] = make_float4(ray.d.x, ray.d.y, ray.d.z, __int_as_float(ray.depth));
        for (int dd = 0; dd < ray.depth; dd++) q[(2 + dd) * RS_Q_CAP + slot] = s_pending[dd * RT_SMALL_THREADS + tid];
    }
    rs_q_unlock(qctl, lane);
    return stored;
}

/* All 32 lanes call.  Up to 32 parked rays leave the queue, one per lane from lane 0 up; returns whether this lane
 * got one (its waiting colours are then in this thread's s_pending column). */
__device__ __forceinline__ bool rs_q_pop(const float4* __restrict__ q, unsigned int* qctl, int lane, RtRay& ray,
                                         int& pixel, float4* __restrict__ s_pending, int tid) {
    rs_q_lock(qctl, lane);
    int base = 0, n = 0;
    if (lane == 0) {
        const int cnt = (int)((volatile unsigned int*)qctl)[1];
        n = min(32, cnt);
        base = cnt - n;
        ((volatile unsigned int*)qctl)[1] = (unsigned int)base;
    }
    base = __shfl_sync(0xffffffffu, base, 0);
    n = __shfl_sync(0xffffffffu, n, 0);
    const bool have = lane < n;
    if (have) {
        const int slot = base + lane;
        const float4 a = q[slot], b = q[RS_Q_CAP + slot];
        ray.o = v3(a.x, a.y, a.z);
        ray.d = v3(b.x, b.y, b.z);
        ray.depth = __float_as_int(b.w);
        ray.off_plane0 = false;          /* the candidate lists of the warp that parked it are not ours: traverse */
        pixel = __float_as_int(a.w);
        for (int dd = 0; dd < ray.depth; dd++) s_pending[dd * RT_SMALL_THREADS + tid] = q[(2 + dd) * RS_Q_CAP + slot];
    }
    rs_q_unlock(qctl, lane);
    return have;
}

/* ---------------------------------------------------------------- tracer */

/* The last warp to finish re-arms the work counter for the next launch on this stream: every warp's fetches
 * precede its arrival here (lane 0 of each warp calls this once).  (A few CTAs become resident only when the
 * frame's units have all been handed out -- the previous frame's FXAA CTAs leave an SM's shared memory in pieces; an
 * early exit for them before the staging cost two spilled registers and 0.9 % of the bench.) */
__device__ __forceinline__ void rs_warp_finished(unsigned int* tile_counter, const unsigned int* order_r) {
    __threadfence();
    if (atomicAdd(tile_counter + 1, 1u) == gridDim.x * RT_SMALL_WARPS - 1) {
        tile_counter[0] = 0;
        tile_counter[1] = 0;
        tile_counter[RS_CNT_TILES] = 0;
#ifdef RS_TIMELINE
        g_tl_frame_no++;
#endif
        if (order_r)                         /* this buffer is written by the next frame */
            for (int c = 0; c < RS_CLASSES; c++) ((unsigned int*)order_r)[c] = 0;
    }
}

template <bool COUNT, bool COMPACT>
__global__ void __launch_bounds__(RT_SMALL_THREADS, RT_SMALL_CTAS_PER_SM)
rt_trace_small_kernel(const RtSceneDev sc, const RtFrameDev f, unsigned int* __restrict__ image,
                      int* __restrict__ hit_ids, RtCountersDev* __restrict__ counters,
                      unsigned int* __restrict__ tile_counter, const unsigned short* __restrict__ tile_lists,
                      int tiles_x, unsigned int n_tiles_all, int bins_x,
                      const unsigned int* __restrict__ work_list, int force_linear,
                      const unsigned int* order_r, unsigned int* order_w, const uint4* __restrict__ lgrid,
                      const unsigned short* __restrict__ mirror_lists, unsigned int tiles_x_magic) {
    /* tiles_x_magic = floor(2^32 / tiles_x) + 1: __umulhi(t, magic) == t / tiles_x for every t with t * tiles_x < 2^32
     * (Granlund-Montgomery; the launcher checks the bound), one IMAD.HI instead of a ~20-instruction division per patch */
    /* fixed-size state first (compile-time shared addresses), the staged scene in dynamic shared memory */
    __shared__ float4 s_pending[RTB_MAX_DEPTH * RT_SMALL_THREADS];
    __shared__ short s_stack[(RT_SMALL_STACK + 2) * RT_SMALL_THREADS];   /* sentinel + entries + 1 scratch */
    __shared__ unsigned short s_list[RT_SMALL_WARPS][2 * RT_SMALL_LIST];   /* per-warp primary + mirror candidates */
    __shared__ __align__(16) float4 s_queue[COMPACT ? RS_Q_FIELDS * RS_Q_CAP : 1];   /* parked rays (ray compaction) */
    __shared__ unsigned int s_qctl[2];
    extern __shared__ __align__(16) unsigned char smem[];

    const int tid = threadIdx.x;
    const int warp = tid >> 5, lane = tid & 31;
    const int n = sc.n_bvh_prims, n_nodes = n > 1 ? n - 1 : 0;
    const int prims_off = 64 * n_nodes;
    /* dynamic shared memory: [planes, RS_PLANES_BYTES][nodes][leaf records][materials]; everything is addressed
     * relative to the nodes (`base`), the planes at a fixed negative offset */
    float4* s_planes = (float4*)smem;
    float4* s_nodes = (float4*)(smem + RS_PLANES_BYTES);
    float4* s_prims = s_nodes + 4 * n_nodes;
    float4* s_mats = s_prims + (RS_LEAF / 16) * n;

    unsigned int u0 = 0;
    if (lane == 0) u0 = atomicAdd(tile_counter, 1u);
    if (COMPACT && tid < 2) s_qctl[tid] = 0u;
#ifdef RS_TIMELINE
    unsigned long long* const tl_frame = tl_slot();
    if (tid == 0) atomicMin(tl_frame + 2, tl_now());
#endif

    /* stage the scene: links become byte offsets from the start of the nodes */
    for (int i = tid; i < 4 * n_nodes; i += RT_SMALL_THREADS) {
        float4 v = __ldg(sc.nodes + i);
        if ((i & 3) == 3) {
            const int c0 = __float_as_int(v.x), c1 = __float_as_int(v.y);
            v.x = __int_as_float(c0 >= 0 ? c0 * 64 : ~(prims_off + (~c0) * RS_LEAF));
            v.y = __int_as_float(c1 >= 0 ? c1 * 64 : ~(prims_off + (~c1) * RS_LEAF));
        }
        s_nodes[i] = v;
    }
    for (int i = tid; i < 3 * n; i += RT_SMALL_THREADS) s_prims[4 * (i / 3) + i % 3] = __ldg(sc.prims + i);
    for (int i = tid; i < n; i += RT_SMALL_THREADS) {
        const vec3 nrm = rt_normalize(rt_cross(f4xyz(__ldg(sc.prims + 3 * i + 1)), f4xyz(__ldg(sc.prims + 3 * i + 2))));
        s_prims[4 * i + 3] = make_float4(nrm.x, nrm.y, nrm.z, 0.0f);      /* (meaningless for a sphere; never read there) */
    }
    for (int i = tid; i < 2 * sc.n_objects; i += RT_SMALL_THREADS) s_mats[i] = __ldg(sc.mats + i);
    for (int i = tid; i < 2 * sc.n_planes; i += RT_SMALL_THREADS) s_planes[i] = __ldg(sc.planes + i);
    __syncthreads();                             /* the only CTA barrier of a working CTA */
    rs_griddep_wait();                           /* the pre-pass has finished: lists, work list, counts */
#ifdef RS_TIMELINE
    if (tid == 0) {
        const unsigned long long t = tl_now();
        const unsigned long long first = atomicMin(tl_frame + 3, t);
        atomicMax(tl_frame + 4, t);
        if (t > first && t - first > 20000ull) atomicAdd(tl_frame + 7, 1ull);   /* [7] CTAs that came 20 us late */
    }
#endif

    /* Work units: single 8x4 patches, fetched one at a time from an atomic counter, nothing reserved ahead.
     * Unit k is patch 3 - k % 4 of tile n_tiles - 1 - k / 4: the frame is walked from its LAST tile to its
     * first (screen-tile shards: the list of the pre-pass, count in tile_counter[RS_CNT_TILES], its own cache
     * line).  Patch costs are heavy-tailed (mean 11 us, p99 52 us, max 111 us in the Island view at 1080p,
     * tools/costmap_probe.py) and the expensive ones -- mirror chains between the sea and the icy slopes -- sit
     * in the lower half of the frame, the sky at the top costs 2 us a patch: bottom-up, the long patches
     * start early and the frame ends on cheap ones.  Measured against the alternatives:
     * whole tiles first / patches last in raster order +5..13 % time, reserving one unit ahead +7 % (a warp
     * inside a 100 us patch sits on its reserved unit), whole tiles over the expensive part +60 %. */
    const bool listed = f.n_shards > 1;          /* whole frames walk all tiles, no list */
    const unsigned int n_tiles = listed ? __ldcg(tile_counter + RS_CNT_TILES) : n_tiles_all;
    const unsigned int n_fetch = RT_SMALL_PATCHES * n_tiles;
    /* Order history (whole frames, order_r != NULL): the previous frame of this size filed every patch under
     * the class of its measured cost (order_w below); this frame walks the classes from the most expensive
     * down, so the longest patches start first wherever they are on the screen.  Lane j < 8 keeps the running
     * end of class 7 - j in walking order.  Results do not depend on the order; a history that does not
     * hold every patch exactly once (first frame, other size) is not used. */
    unsigned int cls_end = 0;
    bool by_cost = false;
    if (order_r) {
        unsigned int c = lane < RS_CLASSES ? __ldcg(order_r + (RS_CLASSES - 1 - lane)) : 0u;
#pragma unroll
        for (int o = 1; o < RS_CLASSES; o <<= 1) {
            const unsigned int v = __shfl_up_sync(0xffffffffu, c, o);
            if (lane >= o) c += v;
        }
        cls_end = c;
        by_cost = __shfl_sync(0xffffffffu, c, RS_CLASSES - 1) == n_fetch;
    }

    RtSmallShared sh;
    /* opaque moves: keep the two addresses in registers instead of re-deriving them in the hot loops */
    {
        const unsigned int b0 = (unsigned int)__cvta_generic_to_shared(smem) + RS_PLANES_BYTES;
        const unsigned int s0 = (unsigned int)__cvta_generic_to_shared(s_stack + tid);
        asm volatile("mov.u32 %0, %1;" : "=r"(sh.base) : "r"(b0));
        asm volatile("mov.u32 %0, %1;" : "=r"(sh.stack0) : "r"(s0));
    }
    sh.defer = false;
    sh.deep = force_linear || __ldg(sc.depth) > RT_SMALL_STACK;
    sh.list = s_list[warp];
    sh.mlist = mirror_lists ? s_list[warp] + RT_SMALL_LIST : nullptr;
    sh.lgrid = lgrid;
    sh.n_prims = n;
    sh.n_planes = sc.n_planes;
    sh.prims_off = prims_off;
    sh.root = n > 1 ? 0 : ~prims_off;
    RtSmallScene<COUNT> acc(sh);

    RtPixelCounters pc = {{0, 0, 0, 0, 0, 0, 0, 0, 0}, 0, 0, 0, 0};
    unsigned int n_pixels = 0;

    u0 = __shfl_sync(0xffffffffu, u0, 0);
#ifdef RS_TIMELINE
    unsigned long long tl_t0, tl_t1 = 0, tl_units = 0;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(tl_t0));
#endif
    if constexpr (!COMPACT) {
        /* every lane follows its own pixel's chain of rays to the end (the default: see RT_OPT_COMPACTION) */
        while (u0 < n_fetch) {
            unsigned int q = 0;                       /* patch id = 4 * tile + patch */
            if (by_cost) {
                const int j = __popc(__ballot_sync(0xffffffffu, lane < RS_CLASSES && u0 >= cls_end));
                const unsigned int start = __shfl_sync(0xffffffffu, cls_end, (j + 31) & 31);   /* end of the class before */
                const unsigned int idx = u0 - (j ? start : 0u);
                q = __ldcg(order_r + RS_ORDER_HEAD + (size_t)(RS_CLASSES - 1 - j) * n_fetch + idx);
            }
            else {
                const unsigned int slot = n_tiles - 1u - u0 / RT_SMALL_PATCHES;
                q = RT_SMALL_PATCHES * (listed ? __ldg(work_list + slot) : slot) + (RT_SMALL_PATCHES - 1u - u0 % RT_SMALL_PATCHES);
            }
            const unsigned int tile = q / RT_SMALL_PATCHES;
            const int p = (int)(q % RT_SMALL_PATCHES);
            const unsigned int unit_t0 = (unsigned int)clock();
            const int by = tiles_x > 1 ? (int)__umulhi(tile, tiles_x_magic) : (int)tile, bx = (int)tile - by * tiles_x;   /* tile / tiles_x, exactly */
            const int x0 = bx * RT_SMALL_TILE_W, y0 = by * RT_SMALL_TILE_H;
            const int bin = (y0 / RT_SMALL_BIN_H) * bins_x + x0 / RT_SMALL_BIN_W;
            const unsigned short entry = __ldg(tile_lists + (size_t)bin * RT_SMALL_LIST + lane);
            s_list[warp][lane] = entry;
            if (mirror_lists) s_list[warp][RT_SMALL_LIST + lane] = __ldg(mirror_lists + (size_t)bin * RT_SMALL_LIST + lane);
            __syncwarp();
            {
                const int x = x0 + (p % (RT_SMALL_TILE_W / 8)) * 8 + (lane & 7);
                const int y = y0 + (p / (RT_SMALL_TILE_W / 8)) * 4 + (lane >> 3);
    #ifdef RS_COSTMAP
                unsigned long long cm_t0;
                asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(cm_t0));
    #endif
                if (rt_patch_pixel_needed(f, x - (lane & 7), y - (lane >> 3), x, y, lane)) {
                    image[y * f.imgw + x] = rt_trace_pixel<RtSmallScene<COUNT>, COUNT>(acc, f, x, y, s_pending,
                                                                                      RT_SMALL_THREADS, tid, hit_ids, pc);
                    n_pixels++;
                }
    #ifdef RS_COSTMAP
                __syncwarp();   /* debug variant: ns spent on the patch, stored over the hit id of its first pixel */
                if (lane == 0 && hit_ids && x < f.imgw && y < f.imgh) {
                    unsigned long long cm_t1;
                    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(cm_t1));
                    hit_ids[y * f.imgw + x] = (int)(cm_t1 - cm_t0);
                }
    #endif
            }
            __syncwarp();                            /* the list is rewritten for the next unit */
    #ifdef RS_TIMELINE
            asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(tl_t1));
            tl_units++;
    #endif
            /* One unit per atomic.  (Round 2 measured a warp reserving 4 units after a cheap patch: the bookkeeping
             * cost two more spilled registers across the pixel trace and the bench 3.5 %, 49.3 vs 51.0 Grays/s, and
             * won nothing even at 4K -- the same-address atomic runs at half of what the L2 serves) */
            unsigned int u1 = 0;
            if (lane == 0) {
                u1 = atomicAdd(tile_counter, 1u);
                if (order_w) {                       /* file the patch under its cost class for the next frame */
                    const unsigned int k = ((unsigned int)clock() - unit_t0) >> 10;
                    const int cls = min(RS_CLASSES - 1, 31 - __clz((int)(k | 1u)));
                    const unsigned int at = atomicAdd(order_w + cls, 1u);
                    order_w[RS_ORDER_HEAD + (size_t)cls * n_fetch + at] = q;
                }
            }
            u0 = __shfl_sync(0xffffffffu, u1, 0);
        }
    } else {
        /* Every turn of the loop runs ONE batch of up to 32 rays on this warp: the primary rays of the next 8x4 patch, or
         * -- first, whenever a full warp's worth is waiting (and at the end whatever is left) -- 32 parked rays of the
         * CTA's queue.  Either way a lane follows its ray's chain until the pixel is finished or its next ray is one to
         * park; rays the queue has no room for are simply finished by the lane that holds them. */
        for (;;) {
            bool from_queue = false;
            {
                unsigned int waiting = 0;
                if (lane == 0) waiting = ((volatile unsigned int*)s_qctl)[1];
                waiting = __shfl_sync(0xffffffffu, waiting, 0);
                from_queue = waiting >= 32u || (u0 >= n_fetch && waiting > 0u);
            }
            if (!from_queue && u0 >= n_fetch) break;

            RtRay ray;
            int pixel = 0;
            bool live = false;                        /* this lane holds a ray */
            unsigned int q = 0;                       /* patch id = 4 * tile + patch */
            unsigned int unit_t0 = 0;
            if (from_queue) {
                live = rs_q_pop(s_queue, s_qctl, lane, ray, pixel, s_pending, tid);
            } else {
                if (by_cost) {
                    const int j = __popc(__ballot_sync(0xffffffffu, lane < RS_CLASSES && u0 >= cls_end));
                    const unsigned int start = __shfl_sync(0xffffffffu, cls_end, (j + 31) & 31);   /* end of the class before */
                    const unsigned int idx = u0 - (j ? start : 0u);
                    q = __ldcg(order_r + RS_ORDER_HEAD + (size_t)(RS_CLASSES - 1 - j) * n_fetch + idx);
                }
                else {
                    const unsigned int slot = n_tiles - 1u - u0 / RT_SMALL_PATCHES;
                    q = RT_SMALL_PATCHES * (listed ? __ldg(work_list + slot) : slot) + (RT_SMALL_PATCHES - 1u - u0 % RT_SMALL_PATCHES);
                }
                const unsigned int tile = q / RT_SMALL_PATCHES;
                const int p = (int)(q % RT_SMALL_PATCHES);
                unit_t0 = (unsigned int)clock();
                const int by = tiles_x > 1 ? (int)__umulhi(tile, tiles_x_magic) : (int)tile, bx = (int)tile - by * tiles_x;   /* tile / tiles_x, exactly */
                const int x0 = bx * RT_SMALL_TILE_W, y0 = by * RT_SMALL_TILE_H;
                const int bin = (y0 / RT_SMALL_BIN_H) * bins_x + x0 / RT_SMALL_BIN_W;
                const unsigned short entry = __ldg(tile_lists + (size_t)bin * RT_SMALL_LIST + lane);
                __syncwarp();                        /* nobody still reads the previous patch's lists */
                s_list[warp][lane] = entry;
                if (mirror_lists) s_list[warp][RT_SMALL_LIST + lane] = __ldg(mirror_lists + (size_t)bin * RT_SMALL_LIST + lane);
                __syncwarp();
                const int x = x0 + (p % (RT_SMALL_TILE_W / 8)) * 8 + (lane & 7);
                const int y = y0 + (p / (RT_SMALL_TILE_W / 8)) * 4 + (lane >> 3);
                if (rt_patch_pixel_needed(f, x - (lane & 7), y - (lane >> 3), x, y, lane)) {
                    live = true;
                    pixel = y * f.imgw + x;
                    ray.o = v3(f.pos[0], f.pos[1], f.pos[2]);
                    ray.d = rt_primary_dir(f, x, y);
                    ray.depth = 0;
                    ray.off_plane0 = false;
                    n_pixels++;
                }
            }
    #ifdef RS_COSTMAP
            unsigned long long cm_t0;
            asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(cm_t0));
            const int cm_pixel = pixel;
    #endif
            /* pass 0: chains run until they end or want to park (compact); what the queue cannot take is finished in
             * pass 1 by the lane that holds it */
    #pragma unroll 1
            for (int pass = 0; pass < 2; pass++) {
                sh.defer = pass == 0;
                bool parked = false;
                if (live) {
                    const int py = pixel / f.imgw, px = pixel - py * f.imgw;
                    unsigned int rgba;
                    if (rt_trace_chain<RtSmallScene<COUNT>, COUNT, true>(acc, f, px, py, ray, s_pending, RT_SMALL_THREADS, tid,
                                                                         hit_ids, pc, rgba)) {
                        image[pixel] = rgba;
                        live = false;
                    } else {
                        parked = true;
                    }
                }
                __syncwarp();
                if (!sh.defer) break;
                if (rs_q_push(s_queue, s_qctl, lane, parked, ray, pixel, s_pending, tid)) live = false;
                if (!__any_sync(0xffffffffu, live)) break;
            }
    #ifdef RS_COSTMAP
            __syncwarp();   /* debug variant: ns spent on the patch, stored over the hit id of its first pixel */
            if (!from_queue && lane == 0 && hit_ids) {
                unsigned long long cm_t1;
                asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(cm_t1));
                hit_ids[cm_pixel] = (int)(cm_t1 - cm_t0);
            }
    #endif
            if (from_queue) continue;
    #ifdef RS_TIMELINE
            asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(tl_t1));
            tl_units++;
    #endif
            unsigned int u1 = 0;
            if (lane == 0) {
                u1 = atomicAdd(tile_counter, 1u);
                if (order_w) {                       /* file the patch under its cost class for the next frame */
                    const unsigned int k = ((unsigned int)clock() - unit_t0) >> 10;
                    const int cls = min(RS_CLASSES - 1, 31 - __clz((int)(k | 1u)));
                    const unsigned int at = atomicAdd(order_w + cls, 1u);
                    order_w[RS_ORDER_HEAD + (size_t)cls * n_fetch + at] = q;
                }
            }
            u0 = __shfl_sync(0xffffffffu, u1, 0);
        }
    }
    rs_griddep_launch();                         /* FXAA may be launched; it waits for this grid to complete */
    if (COUNT) {
        /* all lanes take part: lanes without pixels add zeros */
        rt_flush_counters(pc, n_pixels, counters);
    }
#ifdef RS_TIMELINE
    if (hit_ids && lane == 0) {   /* debug variant: per-warp (start, last unit done, exit, patches) as 4 x u64 */
        unsigned long long tl_t2;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(tl_t2));
        unsigned long long* o = (unsigned long long*)hit_ids + 4 * (size_t)(blockIdx.x * RT_SMALL_WARPS + warp);
        o[0] = tl_t0; o[1] = tl_t1; o[2] = tl_t2; o[3] = tl_units;
    }
    if (lane == 0) {
        if (tl_units) atomicMax(tl_frame + 5, tl_t1);
        atomicMax(tl_frame + 6, tl_now());
    }
#endif
    if (lane == 0) rs_warp_finished(tile_counter, order_r);
}

/* dynamic shared memory: the staged scene */
extern "C" size_t rt_trace_small_smem_bytes(int n_bvh_prims, int n_objects, int n_planes) {
    const size_t n_nodes = n_bvh_prims > 1 ? (size_t)n_bvh_prims - 1 : 0;
    (void)n_planes;   /* the plane block has a fixed size (<= 16 planes) */
    return RS_PLANES_BYTES + 16 * (4 * n_nodes + (RS_LEAF / 16) * (size_t)n_bvh_prims + 2 * (size_t)n_objects);
}

/* scenes this kernel accepts: link offsets (64 n_nodes + RS_LEAF n) must fit the 16-bit stack entries */
extern "C" int rt_trace_small_supported(int n_bvh_prims, int n_objects, int n_planes) {
    return n_bvh_prims <= RT_SMALL_MAX_PRIMS && n_planes <= 16 &&
           rt_trace_small_smem_bytes(n_bvh_prims, n_objects, n_planes) <= RT_SMALL_DYN_SMEM_MAX;
}

/* per device, once (rt_create): opt in to the dynamic shared memory the largest accepted scene needs */
extern "C" cudaError_t rt_trace_small_init(void) {
    cudaError_t e = cudaFuncSetAttribute(rt_trace_small_kernel<false, false>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         RT_SMALL_DYN_SMEM_MAX);
    if (e == cudaSuccess)
        e = cudaFuncSetAttribute(rt_trace_small_kernel<true, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, RT_SMALL_DYN_SMEM_MAX);
    if (e == cudaSuccess)
        e = cudaFuncSetAttribute(rt_trace_small_kernel<false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, RT_SMALL_DYN_SMEM_MAX);
    if (e == cudaSuccess)
        e = cudaFuncSetAttribute(rt_trace_small_kernel<true, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, RT_SMALL_DYN_SMEM_MAX);
    /* Every kernel of the frame chain asks for the SAME shared-memory carve-out (the largest, what the tracer needs).
     * Two kernels with different carve-outs cannot share an SM, so a dependent launch could not become resident
     * under its predecessor's tail: see rt_fxaa_init. */
    if (e == cudaSuccess) e = cudaFuncSetAttribute(rt_cull_tiles_kernel, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(rt_trace_small_kernel<false, false>, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(rt_trace_small_kernel<true, false>, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(rt_trace_small_kernel<false, true>, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(rt_trace_small_kernel<true, true>, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
    return e;
}

/* bytes of tile-list storage a frame of this size needs */
extern "C" size_t rt_trace_small_list_bytes(int imgw, int imgh) {
    const size_t bins = (size_t)((imgw + RT_SMALL_BIN_W - 1) / RT_SMALL_BIN_W) * ((imgh + RT_SMALL_BIN_H - 1) / RT_SMALL_BIN_H);
    const size_t tiles = (size_t)((imgw + RT_SMALL_TILE_W - 1) / RT_SMALL_TILE_W) * ((imgh + RT_SMALL_TILE_H - 1) / RT_SMALL_TILE_H);
    return 2 * bins * RT_SMALL_LIST * sizeof(unsigned short) + tiles * sizeof(unsigned int);   /* bin lists + work list + mirror lists */
}

extern "C" size_t rt_trace_small_patches(int imgw, int imgh) {
    return (size_t)RT_SMALL_PATCHES * ((imgw + RT_SMALL_TILE_W - 1) / RT_SMALL_TILE_W) * ((imgh + RT_SMALL_TILE_H - 1) / RT_SMALL_TILE_H);
}

/* bytes of ONE order buffer (class counts + RS_CLASSES lists of all patches) for a frame of this size */
extern "C" size_t rt_trace_small_order_bytes(int imgw, int imgh) {
    const size_t tiles = (size_t)((imgw + RT_SMALL_TILE_W - 1) / RT_SMALL_TILE_W) * ((imgh + RT_SMALL_TILE_H - 1) / RT_SMALL_TILE_H);
    return sizeof(unsigned int) * (RS_ORDER_HEAD + (size_t)RS_CLASSES * RT_SMALL_PATCHES * tiles);
}

extern "C" cudaError_t rt_launch_trace_small(const RtSceneDev* sc, const RtFrameDev* f, unsigned int* image,
                                             int* hit_ids, RtCountersDev* counters, unsigned int* tile_counter,
                                             unsigned short* tile_lists, int count, int force_linear, int n_sm,
                                             const unsigned int* order_r, unsigned int* order_w,
                                             int after_own_post, uint4* lgrid, int build_lgrid, int mirror_bins,
                                             int compact, cudaStream_t stream) {
    /* order_r / order_w: order history read / written by this frame (NULL: none), see the kernel.
     * after_own_post: the previous operation on the stream is the FXAA / copy kernel of this context's previous
     * frame (nothing this frame reads is still being written): the pre-pass overlaps it, see rt_cull_tiles_kernel */
    const size_t smem = rt_trace_small_smem_bytes(sc->n_bvh_prims, sc->n_objects, sc->n_planes);
    const int tiles_x = (f->imgw + RT_SMALL_TILE_W - 1) / RT_SMALL_TILE_W;
    const int tiles = tiles_x * ((f->imgh + RT_SMALL_TILE_H - 1) / RT_SMALL_TILE_H);
    const int bins_x = (f->imgw + RT_SMALL_BIN_W - 1) / RT_SMALL_BIN_W;
    const int bins = bins_x * ((f->imgh + RT_SMALL_BIN_H - 1) / RT_SMALL_BIN_H);
    const int tiles_y = (f->imgh + RT_SMALL_TILE_H - 1) / RT_SMALL_TILE_H;
    /* behind the bin lists: the work list of sharded frames (filled by the pre-pass) */
    unsigned int* work_list = (unsigned int*)(tile_lists + (size_t)bins * RT_SMALL_LIST);
    /* behind the work list: the mirror lists of the bins */
    unsigned short* mirror_lists = mirror_bins ? (unsigned short*)(work_list + tiles) : (unsigned short*)nullptr;
    const bool sharded = f->n_shards > 1;
    cudaLaunchAttribute pdl[1];
    pdl[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    pdl[0].val.programmaticStreamSerializationAllowed = 1;
    cudaLaunchConfig_t cfg = {};
    cfg.stream = stream;
    cfg.attrs = pdl;
    cfg.numAttrs = 1;
    const bool with_lights = lgrid && build_lgrid;
    const int bins_y = (f->imgh + RT_SMALL_BIN_H - 1) / RT_SMALL_BIN_H;
    const int cull_blocks = ((bins_x + RS_CB_W - 1) / RS_CB_W) * ((bins_y + RS_CB_H - 1) / RS_CB_H);
    const int lit = f->n_lights < RT_LGRID_LIGHTS ? f->n_lights : RT_LGRID_LIGHTS;    /* lights with bins */
    /* build_lgrid = 0: the light bins in `lgrid` are still those of this scene and these lights (api.cu) */
    cfg.gridDim = dim3(cull_blocks + (with_lights ? lit * RS_LG_BLOCKS_PER_LIGHT : 0), 1, 1);
    cfg.blockDim = dim3(256, 1, 1);
    cfg.dynamicSmemBytes = 0;
    RT_LAUNCHED(2);                              /* pre-pass + tracer */
    cudaError_t e = cudaLaunchKernelEx(&cfg, rt_cull_tiles_kernel, *sc, *f, tile_lists, bins_x, bins_y,
                                       sharded ? work_list : (unsigned int*)nullptr, tile_counter + RS_CNT_TILES, tiles_x,
                                       tiles_y, after_own_post ? 0 : 1, lgrid, cull_blocks, mirror_lists);
    if (e != cudaSuccess) return e;
    int grid = n_sm * RT_SMALL_CTAS_PER_SM;
    const int units = tiles * RT_SMALL_PATCHES;
    if (grid > (units + RT_SMALL_WARPS - 1) / RT_SMALL_WARPS) grid = (units + RT_SMALL_WARPS - 1) / RT_SMALL_WARPS;
    cfg.gridDim = dim3(grid, 1, 1);
    cfg.blockDim = dim3(RT_SMALL_THREADS, 1, 1);
    cfg.dynamicSmemBytes = smem;
    const unsigned int* wl = work_list;
    const unsigned short* tl = tile_lists;
    const uint4* lg = lgrid;
    const unsigned short* ml = mirror_lists;
#define RS_LAUNCH(C, Q) cudaLaunchKernelEx(&cfg, rt_trace_small_kernel<C, Q>, *sc, *f, image, hit_ids, counters, tile_counter, tl, \
                                          tiles_x, (unsigned int)tiles, bins_x, wl, force_linear, order_r, order_w, lg, ml, tiles_x_magic)
    if ((unsigned long long)tiles * (unsigned long long)tiles_x >= (1ull << 32)) return cudaErrorInvalidValue;   /* > 1e9 pixels */
    const unsigned int tiles_x_magic = tiles_x > 1 ? (unsigned int)((1ull << 32) / (unsigned long long)tiles_x) + 1u : 0u;   /* one column: no division */
    if (compact) e = count ? RS_LAUNCH(true, true) : RS_LAUNCH(false, true);
    else         e = count ? RS_LAUNCH(true, false) : RS_LAUNCH(false, false);
#undef RS_LAUNCH
    if (e != cudaSuccess) return e;
    return cudaGetLastError();
}
