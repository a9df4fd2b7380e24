/*
 * frustum.cuh -- the conservative screen-tile cull shared by the binning pre-pass of the shared-memory tracer
 * (trace_small.cu: candidate lists per 64x16-pixel bin) and the beam cull of the persistent HBM tracer
 * (trace_big.cu: candidate lists per 8x4-pixel patch, gathered by a cooperative walk of the LBVH).
 *
 * A tile's primary rays (reference kernel.cu:244-253) all lie in the cone spanned by the four corner directions of
 * the tile grown by one pixel; a primitive no ray of that cone can hit is dropped, everything else is decided by
 * the exact leaf tests, so the (t, index) minimum over the survivors equals the reference's loop over all objects.
 * Every plane test carries a world-space pad (rs_cull_pad, light_bins.cuh) orders of magnitude above the rounding of
 * the ray set-up and of the acceptance region of the leaf tests.
 */
#ifndef RTB_FRUSTUM_CUH
#define RTB_FRUSTUM_CUH

#include "light_bins.cuh"

/* un-normalised primary-ray target at a (fractional) pixel position: the bilinear blend of
 * reference kernel.cu:246-251 */
__device__ __forceinline__ vec3 rs_primary_target(const RtFrameDev& f, float px, float py) {
    const float partX = px / (float)(f.imgw - 1), partY = py / (float)(f.imgh - 1);
    vec3 t;
    {
        const float vd = f.LD[0] + partX * (f.RD[0] - f.LD[0]), vu = f.LU[0] + partX * (f.RU[0] - f.LU[0]);
        t.x = vu - partY * (vu - vd);
    }
    {
        const float vd = f.LD[1] + partX * (f.RD[1] - f.LD[1]), vu = f.LU[1] + partX * (f.RU[1] - f.LU[1]);
        t.y = vu - partY * (vu - vd);
    }
    {
        const float vd = f.LD[2] + partX * (f.RD[2] - f.LD[2]), vu = f.LU[2] + partX * (f.RU[2] - f.LU[2]);
        t.z = vu - partY * (vu - vd);
    }
    return t;
}

struct RsFrustum {
    vec3 n[4];     /* unit inward normals of the 4 side planes through the eye */
    bool ok;       /* false: degenerate / non-convex corner set -> keep every primitive */
};

/* the cone of all positive combinations of four corner directions c[0..3] (in order around the cone) */
__device__ __forceinline__ RsFrustum rs_frustum_from_corners(const vec3* c) {
    const vec3 mid = vadd(vadd(c[0], c[1]), vadd(c[2], c[3]));
    RsFrustum fr;
    fr.ok = true;
#pragma unroll
    for (int k = 0; k < 4; k++) {
        const vec3 a = c[k], b = c[(k + 1) & 3];
        vec3 n = vcross(a, b);
        const float len = sqrtf(vdot(n, n));
        const float s = (vdot(n, mid) < 0.0f ? -1.0f : 1.0f) / len;
        n = vscale(n, s);
        /* every ray of the cone is a convex combination of the four corners: it is inside plane k if the
         * two corners that do not span it are */
        const float e0 = vdot(n, c[(k + 2) & 3]), e1 = vdot(n, c[(k + 3) & 3]);
        if (!(len > 0.0f) || !(e0 > 0.0f) || !(e1 > 0.0f)) fr.ok = false;
        fr.n[k] = n;
    }
    return fr;
}

/* corner directions of the tile [x0, x1] x [y0, y1] grown by one pixel on every side */
__device__ __forceinline__ void rs_tile_corners(const RtFrameDev& f, int x0, int y0, int x1, int y1, vec3* c) {
    const float xa = (float)x0 - 1.0f, xb = (float)x1 + 1.0f, ya = (float)y0 - 1.0f, yb = (float)y1 + 1.0f;
    c[0] = rs_primary_target(f, xa, ya);
    c[1] = rs_primary_target(f, xb, ya);
    c[2] = rs_primary_target(f, xb, yb);
    c[3] = rs_primary_target(f, xa, yb);
}

/* can any ray of the tile (origin `eye`, inside `fr`) hit leaf record `rec`? */
__device__ __forceinline__ bool rs_tile_may_hit(const RsFrustum& fr, const vec3& eye, const float4 q0, const float4 q1,
                                                const float4 q2, float pad) {
    const unsigned int tag = __float_as_uint(q0.w);
    const vec3 a = vsub(f4xyz(q0), eye);
    if (tag & RT_TAG_SPHERE) {
        /* d2 = L.L - tca^2 carries rounding noise of about 2^-21 |L|^2: grow the radius accordingly */
        const float r = sqrtf(q1.y + 3.814697265625e-06f * vdot(a, a)) + pad;
#pragma unroll
        for (int k = 0; k < 4; k++)
            if (vdot(fr.n[k], a) < -r) return false;
        return true;
    }
    const vec3 e1 = f4xyz(q1), e2 = f4xyz(q2);
    /* det = -d.Ng for every ray: the eye on the back side means `det < 0.001` rejects them all */
    const vec3 ng = vcross(e1, e2);
    const float side = vdot(ng, a);
    if (side > 1e-3f * sqrtf(vdot(ng, ng) * vdot(a, a)) + pad * sqrtf(vdot(ng, ng))) return false;
    const vec3 b = vadd(a, e1), c = vadd(a, e2);
#pragma unroll
    for (int k = 0; k < 4; k++)
        if (vdot(fr.n[k], a) < -pad && vdot(fr.n[k], b) < -pad && vdot(fr.n[k], c) < -pad) return false;
    return true;
}
__device__ __forceinline__ bool rs_tile_may_hit(const RsFrustum& fr, const vec3& eye, const float4* __restrict__ rec,
                                                float pad) {
    return rs_tile_may_hit(fr, eye, __ldg(rec), __ldg(rec + 1), __ldg(rec + 2), pad);
}


/* can any ray of the cone (apex `eye`) meet the box of centre c and half-extent e?  Outside plane k when even the
 * box's farthest corner along n[k] is behind it.  (A cone of less than a half-space holds nothing behind its apex,
 * so boxes behind the eye fail some plane.) */
__device__ __forceinline__ bool rs_frustum_may_hit_box(const RsFrustum& fr, const vec3& eye, float cx, float cy, float cz,
                                                       float ex, float ey, float ez, float pad) {
    const float ax = cx - eye.x, ay = cy - eye.y, az = cz - eye.z;
#pragma unroll
    for (int k = 0; k < 4; k++) {
        const float far = fr.n[k].x * ax + fr.n[k].y * ay + fr.n[k].z * az +
                          fabsf(fr.n[k].x) * ex + fabsf(fr.n[k].y) * ey + fabsf(fr.n[k].z) * ez;
        if (far < -pad) return false;
    }
    return true;
}

#endif
