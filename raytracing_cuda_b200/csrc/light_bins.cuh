/*
 * light_bins.cuh -- the conservative cull geometry shared by the pre-pass of the shared-memory tracer
 * (trace_small.cu: light bins of small scenes) and the light grid of big scenes (light_grid.cu).
 * See the `light bins` comment below for what the bins are and why the answer is exact.
 */
#ifndef RTB_LIGHT_BINS_CUH
#define RTB_LIGHT_BINS_CUH

#include "trace_pixel.cuh"

#define RS_LG_DELTA 0.00390625f              /* 2^-8: |s|, |t| up to 1 + delta belong to a face */
/* bin = (s + 1 + delta) * scale, G bins per face edge */
#define RS_LG_SCALE_OF(G) ((float)(G) / (2.0f + 2.0f * RS_LG_DELTA))

/* world-space pad of the cull tests from the scene's extent and the apex (eye or light) position: 2^-13 of the
 * largest coordinate in play */
__device__ __forceinline__ float rs_cull_pad(const RtSceneDev& sc, const float* apex) {
    float m = fmaxf(fmaxf(fabsf(apex[0]), fabsf(apex[1])), fabsf(apex[2]));
    if (sc.n_bvh_prims > 1) {
        const float4 a = __ldg(sc.nodes), b = __ldg(sc.nodes + 1), c = __ldg(sc.nodes + 2);
        const float r[12] = {a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w, c.x, c.y, c.z, c.w};
#pragma unroll
        for (int k = 0; k < 2; k++)
#pragma unroll
            for (int q = 0; q < 3; q++) m = fmaxf(m, fabsf(r[6 * k + q]) + r[6 * k + 3 + q]);
    }
    return 1.220703125e-04f * m + 1e-6f;      /* 2^-13 */
}

/* ---------------------------------------------------------------- light bins
 * Every shadow ray of light L lies on a line through L (kernel.cu:177-182: origin P + 0.001 dir, dir towards L), so
 * which primitives it can meet depends on its DIRECTION SEEN FROM L only.  Per light, a cube map around L with
 * RT_LGRID_G x RT_LGRID_G direction bins per face holds the (at most 7) non-emissive primitives a line of that
 * bin can hit; the tracer looks its shadow ray up (major axis of P - L, two divisions) and tests that list instead
 * of traversing the hierarchy.  Triangles are one-sided (`det < 0.001`, kernel.cu:104): towards a light only
 * those facing AWAY from it can be hit, which halves the lists.  Conservative like the primary bins: cones grown
 * by 1 % of a bin, world-space pad on every plane test, faces overlapping by RS_LG_DELTA; the exact leaf tests
 * decide, so the answer equals the reference's loop over all objects.
 * Face f = 2 axis + (negative side); face coordinates (s, t) = components (axis+1)%3, (axis+2)%3 over |axis|. */

/* A primitive seen from one light in the coordinates of one cube face: (m, a, b) = (component along the face
 * axis, made positive towards the face; the two others).  A direction of the face is (1, s, t).
 *   sphere:   v[0] = centre, r = radius grown by the rounding noise of the leaf test and the pad
 *   triangle: v[0..2] = vertices, e[k] = plane through the light and edge k as (A, B, C): E_k(s, t) = A + B s + C t
 *             is >= 0 on the triangle's side, tol[k] its pad. */
struct RsLgPrim {
    float v[3][3];
    float e[3][3];
    float tol[3];
    float r;            /* > 0: sphere */
    float fill[2];      /* 96 B = 6 float4 */
};
static_assert(sizeof(RsLgPrim) % 16 == 0, "RsLgPrim is copied as float4s");

/* the bins [g0s, g1s) x [g0t, g1t) of a face, grown by 1 % of a bin: bounds in s, t and the reciprocal lengths of
 * the normals (-s, 1, 0), (-t, 0, 1) of the four side planes */
struct RsLgRect {
    float s0, s1, t0, t1, is0, is1, it0, it1;
};
template <int G>
__device__ __forceinline__ float rs_lg_coord(float g) { return g * (1.0f / RS_LG_SCALE_OF(G)) - (1.0f + RS_LG_DELTA); }
template <int G>
__device__ __forceinline__ RsLgRect rs_lg_rect(float g0s, float g0t, float g1s, float g1t) {
    RsLgRect r;
    r.s0 = rs_lg_coord<G>(g0s - 0.01f); r.s1 = rs_lg_coord<G>(g1s + 0.01f);
    r.t0 = rs_lg_coord<G>(g0t - 0.01f); r.t1 = rs_lg_coord<G>(g1t + 0.01f);
    r.is0 = rsqrtf(1.0f + r.s0 * r.s0); r.is1 = rsqrtf(1.0f + r.s1 * r.s1);
    r.it0 = rsqrtf(1.0f + r.t0 * r.t0); r.it1 = rsqrtf(1.0f + r.t1 * r.t1);
    return r;
}

/* world vector -> face coordinates */
__device__ __forceinline__ void rs_lg_to_face(int face, const vec3& w, float* o) {
    const float sg = (face & 1) ? -1.0f : 1.0f;
    const int axis = face >> 1;
    o[0] = sg * (axis == 0 ? w.x : (axis == 1 ? w.y : w.z));
    o[1] = axis == 0 ? w.y : (axis == 1 ? w.z : w.x);
    o[2] = axis == 0 ? w.z : (axis == 1 ? w.x : w.y);
}

/* Leaf record `rec` as an occluder of the light at `lp`: false if it can never be one (emissive, kernel.cu:189;
 * a triangle showing the light its front, `det < 0.001`).  d = -dir(ray) leaves the light: det = d.Ng > 0 needs
 * the light on the BACK side. */
__device__ __forceinline__ bool rs_lg_prim(int face, const vec3& lp, const float4* __restrict__ rec, float pad,
                                           RsLgPrim& p) {
    const float4 q0 = __ldg(rec), q1 = __ldg(rec + 1);
    const unsigned int tag = __float_as_uint(q0.w);
    if (tag & RT_TAG_EMISSIVE) return false;
    const vec3 a = vsub(f4xyz(q0), lp);
    if (tag & RT_TAG_SPHERE) {
        rs_lg_to_face(face, a, p.v[0]);
        /* d2 = L.L - tca^2 carries rounding noise of about 2^-21 |L|^2 (as in rs_tile_may_hit) */
        p.r = sqrtf(q1.y + 3.814697265625e-06f * vdot(a, a)) + pad;
        return true;
    }
    const float4 q2 = __ldg(rec + 2);
    const vec3 e1 = f4xyz(q1), e2 = f4xyz(q2);
    const vec3 ng = vcross(e1, e2);
    const float side = vdot(ng, a);
    if (-side > 1e-3f * sqrtf(vdot(ng, ng) * vdot(a, a)) + pad * sqrtf(vdot(ng, ng))) return false;
    const vec3 b = vadd(a, e1), c = vadd(a, e2);
    rs_lg_to_face(face, a, p.v[0]);
    rs_lg_to_face(face, b, p.v[1]);
    rs_lg_to_face(face, c, p.v[2]);
    p.r = 0.0f;
    /* planes through the light and the edges, oriented towards the third vertex.  A point y within `pad` of the
     * triangle has n.y >= -|n| pad; as y = tau (1, s, t) with |y| >= |n| / |edge| - pad (the edge's line is that
     * far from the light) and |(1, s, t)| <= 1.75, E(s, t) >= -1.75 pad |edge| / (1 - 1/8) = -2 pad |edge| */
    float T = 0.0f;
#pragma unroll
    for (int k = 0; k < 3; k++) {
        const float* u = p.v[k];
        const float* w = p.v[(k + 1) % 3];
        p.e[k][0] = u[1] * w[2] - u[2] * w[1];
        p.e[k][1] = u[2] * w[0] - u[0] * w[2];
        p.e[k][2] = u[0] * w[1] - u[1] * w[0];
        const float dx = w[0] - u[0], dy = w[1] - u[1], dz = w[2] - u[2];
        /* |n| / |edge| = distance of the edge's line from the light; closer than 8 pads: no test for this edge */
        const float el = sqrtf(dx * dx + dy * dy + dz * dz);
        const float nl2 = p.e[k][0] * p.e[k][0] + p.e[k][1] * p.e[k][1] + p.e[k][2] * p.e[k][2];
        p.tol[k] = nl2 > 64.0f * pad * pad * el * el ? 2.5f * pad * el : 3.0e38f;
        if (k == 0) T = p.e[0][0] * p.v[2][0] + p.e[0][1] * p.v[2][1] + p.e[0][2] * p.v[2][2];
    }
    const float la = vdot(a, a), lb = vdot(b, b), lc = vdot(c, c);
    const bool flat = !(T * T > 1e-10f * la * lb * lc);       /* light (nearly) in the triangle's plane: no edge tests */
    const float sg = flat ? 0.0f : (T > 0.0f ? 1.0f : -1.0f);
#pragma unroll
    for (int k = 0; k < 3; k++)
#pragma unroll
        for (int q = 0; q < 3; q++) p.e[k][q] *= sg;
    return true;
}

/* can a line through the light with a direction inside the rectangle hit the primitive? */
__device__ __forceinline__ bool rs_lg_may_hit(const RsLgPrim& p, const RsLgRect& r, float pad) {
    if (p.r > 0.0f) {
        const float m = p.v[0][0], a = p.v[0][1], b = p.v[0][2];
        if ((a - r.s0 * m) * r.is0 < -p.r || (r.s1 * m - a) * r.is1 < -p.r) return false;
        if ((b - r.t0 * m) * r.it0 < -p.r || (r.t1 * m - b) * r.it1 < -p.r) return false;
        return true;
    }
    bool o0 = true, o1 = true, o2 = true, o3 = true, ob = true;
#pragma unroll
    for (int k = 0; k < 3; k++) {
        const float m = p.v[k][0], a = p.v[k][1], b = p.v[k][2];
        o0 = o0 && (a - r.s0 * m) * r.is0 < -pad;
        o1 = o1 && (r.s1 * m - a) * r.is1 < -pad;
        o2 = o2 && (b - r.t0 * m) * r.it0 < -pad;
        o3 = o3 && (r.t1 * m - b) * r.it1 < -pad;
        ob = ob && m < -pad;
    }
    if (o0 || o1 || o2 || o3 || ob) return false;
#pragma unroll
    for (int k = 0; k < 3; k++) {
        const float big = p.e[k][0] + fmaxf(p.e[k][1] * r.s0, p.e[k][1] * r.s1) + fmaxf(p.e[k][2] * r.t0, p.e[k][2] * r.t1);
        if (big < -p.tol[k]) return false;
    }
    return true;
}

#endif
