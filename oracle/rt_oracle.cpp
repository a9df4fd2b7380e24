/*
 * rt_oracle.cpp -- CPU ORACLE for the ray-tracing hot path.  TEST INFRASTRUCTURE.
 *
 * This file is NOT part of the product.  Only tests/, __graft_entry__.smoke()
 * and bench.py's cpu_baseline / --impl reference legs may build, load or call it.
 * The shipped path is the CUDA library under raytracing_cuda_b200/csrc and it
 * has no CPU fallback.
 *
 * What it is: a plain C++ restatement (OpenMP over rows) of the reference's
 * device code, primitive by primitive and ray by ray, with the reference's
 * LINEAR loop over all objects (no BVH), so it is an independent check of the
 * LBVH traversal as well as of the arithmetic:
 *
 *   pack_rgb()        reference kernel.cu:26-32
 *   check_hit()       reference kernel.cu:41-129   (sphere / plane / culled Moller-Trumbore)
 *   trace()           reference kernel.cu:131-225  (closest hit, sky, Phong + hard shadows, mirror to depth 4)
 *   primary ray       reference kernel.cu:241-257
 *   fxaa_pixel()      reference kernel.cu:262-403  (as a pure function of the 3x3 neighbourhood;
 *                                                   border pixels pass through, reference kernel.cu:330,399-402)
 *
 * Precision: the host flavour of the reference's helpers is followed
 * (structs.h:75-84: norm = sqrtf(dot), normalize = scale by (float)(1.0/norm));
 * double-typed literals are kept double (0.001, -0.01, 0.00001, PI=3.141592,
 * /12.0, max(0.0,.)).  __powf(x,y) is exp2f(y*log2f(x)); tex2D is a
 * point-sampled clamp-addressed fetch at floor(u*W), floor(v*H).  Built with
 * -ffp-contract=off so no FMA is formed.
 *
 * Parity pinning: oracle/build_ref.py compiles the reference's own kernel.cu
 * text (kernel.cu:9-403) for the host under shim macros into
 * oracle/_ref/libref_trace_cpu.so; tests/test_oracle.py requires
 * this restatement to match it bit-for-bit (image, pre-FXAA image) and
 * tests/golden/ holds vectors generated from that reference build.
 */
#include <cmath>
#include <cstdint>
#include <cstring>
#include <algorithm>

#include "../include/rtb_structs.h"

namespace {

struct V3 { float x, y, z; };

inline V3 mk(float x, float y, float z) { return V3{x, y, z}; }
inline V3 from(const rtb_float3& a) { return V3{a.x, a.y, a.z}; }
inline V3 add(V3 a, V3 b) { return mk(a.x + b.x, a.y + b.y, a.z + b.z); }
inline V3 sub(V3 a, V3 b) { return mk(a.x - b.x, a.y - b.y, a.z - b.z); }
inline float dot(V3 a, V3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
inline V3 scale(V3 a, float s) { return mk(a.x * s, a.y * s, a.z * s); }
inline V3 cross(V3 a, V3 b) { return mk(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x); }
inline V3 had(V3 a, V3 b) { return mk(a.x * b.x, a.y * b.y, a.z * b.z); }
inline float length(V3 v) { return sqrtf(dot(v, v)); }                       /* structs.h:79 */
inline V3 normalize(V3 v) { return scale(v, (float)(1.0 / length(v))); }     /* structs.h:82-84 */
inline float clampf(float x, float a, float b) { return std::max(a, std::min(b, x)); }

struct Ray { V3 origin, dir; };

struct Scene {
    const rtb_Object* objects;
    int n_objects;
    const rtb_Light* lights;
    int n_lights;
    V3 ambient;
    float skyVars[4];
    const unsigned char* tex[4];
    int texw, texh;
    float dayTime;
};

struct Counters {
    unsigned long long rays_primary, rays_shadow, rays_mirror, prim_tests, sky_lookups;
};

/* reference kernel.cu:26-32 */
inline unsigned int pack_rgb(float r, float g, float b) {
    r = clampf(r, 0.0f, 255.0f);
    g = clampf(g, 0.0f, 255.0f);
    b = clampf(b, 0.0f, 255.0f);
    return (unsigned int)(((int)b << 16) | ((int)g << 8) | (int)r);
}

/* reference kernel.cu:41-129 */
inline bool check_hit(const Scene& sc, const Ray& ray, int index, V3& hitPos, V3& hitNormal, float& hitDist) {
    const rtb_Object& o = sc.objects[index];
    if (o.type == RTB_SPHERE) {
        float sr = o.size.x;
        float sr2 = sr * sr;
        V3 L = sub(from(o.pos), ray.origin);
        float tca = dot(L, ray.dir);
        if (tca <= 0) return false;
        float d2 = dot(L, L) - tca * tca;
        if (d2 >= sr2 || d2 <= -0.01) return false;
        float thc = sqrtf(sr2 - d2);
        hitDist = tca - thc;                               /* near root only; may be negative */
        hitPos = add(ray.origin, scale(ray.dir, hitDist));
        hitNormal = normalize(sub(hitPos, from(o.pos)));
        return true;
    }
    if (o.type == RTB_PLANE) {
        V3 n = from(o.size);
        float denom = dot(ray.dir, n);
        if (denom * denom > 0.00001) {
            float t = dot(sub(from(o.pos), ray.origin), n) / denom;
            if (t < 0) return false;
            hitDist = t;
            hitPos = add(ray.origin, scale(ray.dir, t));
            hitNormal = n;
            return true;
        }
        return false;
    }
    if (o.type == RTB_TRIANGLE) {
        V3 v0 = from(o.pos), v1 = from(o.size), v2 = from(o.third);
        V3 e1 = sub(v1, v0);
        V3 e2 = sub(v2, v0);
        V3 pvec = cross(ray.dir, e2);
        float det = dot(e1, pvec);
        if (det < 0.001) return false;                     /* back-face + near-parallel cull */
        float invDet = 1 / det;
        V3 tvec = sub(ray.origin, v0);
        float u = dot(tvec, pvec) * invDet;
        if (u < 0 || u > 1) return false;
        V3 qvec = cross(tvec, e1);
        float v = dot(ray.dir, qvec) * invDet;
        if (v < 0 || u + v > 1) return false;
        float t = dot(e2, qvec) * invDet;
        if (t < 0) return false;
        hitDist = t;
        hitPos = add(ray.origin, scale(ray.dir, t));
        hitNormal = normalize(cross(e1, e2));
        return true;
    }
    return false;
}

inline const unsigned char* texel(const Scene& sc, int which, float u, float v) {
    int ix = (int)floorf(u * (float)sc.texw);
    int iy = (int)floorf(v * (float)sc.texh);
    ix = std::min(std::max(ix, 0), sc.texw - 1);
    iy = std::min(std::max(iy, 0), sc.texh - 1);
    return sc.tex[which] + 4 * ((size_t)iy * sc.texw + ix);
}

/* reference kernel.cu:154-164 with the uchar4 operators of structs.h:86-91 */
inline V3 sky(const Scene& sc, const Ray& ray) {
    const double PI = RTB_PI;
    float y = (float)(1 - (asinf(ray.dir.y) + PI / 2.0f) / PI);
    float x = fmodf((float)((atan2f(ray.dir.x, ray.dir.z) + PI) / (2.0f * PI) + sc.dayTime), 1);
    unsigned char acc[3] = {0, 0, 0};
    for (int k = 0; k < 4; k++) {
        const unsigned char* t = texel(sc, k, x, y);
        for (int c = 0; c < 3; c++) {
            unsigned char scaled = (unsigned char)(t[c] * sc.skyVars[k]);   /* truncating */
            acc[c] = (unsigned char)(acc[c] + scaled);                      /* wraps mod 256 */
        }
    }
    return scale(mk(acc[0], acc[1], acc[2]), 1.0f / 255.0f);
}

/* reference kernel.cu:131-225 */
V3 trace(const Scene& sc, const Ray& ray, int depth, int* primary_id, Counters& cnt) {
    if (depth > RTB_MAX_DEPTH) return mk(0, 0, 0);          /* kernel.cu:222-225 */
    if (depth == 0) cnt.rays_primary++; else cnt.rays_mirror++;

    float minHitDist = -1;
    V3 minHitPos = mk(0, 0, 0), minHitNormal = mk(0, 0, 0);
    int index = -1;
    V3 hitPos, hitNormal;
    float hitDist;

    for (int i = 0; i < sc.n_objects; i++) {
        cnt.prim_tests++;
        if (check_hit(sc, ray, i, hitPos, hitNormal, hitDist) && (hitDist < minHitDist || minHitDist == -1)) {
            minHitDist = hitDist;
            minHitPos = hitPos;
            minHitNormal = hitNormal;
            index = i;
        }
    }
    if (primary_id) *primary_id = index;

    if (index == -1) {
        cnt.sky_lookups++;
        return sky(sc, ray);
    }
    const rtb_Object& o = sc.objects[index];
    if (o.light) return from(o.color);

    V3 phong = had(from(o.color), sc.ambient);
    for (int i = 0; i < sc.n_lights; i++) {
        const rtb_Light& l = sc.lights[i];
        V3 vec = sub(from(l.pos), minHitPos);
        float shadowDist = length(vec);
        V3 shadowDir = normalize(vec);
        Ray shadow = {minHitPos, shadowDir};
        shadow.origin = add(shadow.origin, scale(shadow.dir, (float)0.001));
        cnt.rays_shadow++;

        float angle = (float)std::max(0.0, (double)dot(minHitNormal, shadowDir));
        for (int k = 0; k < sc.n_objects; k++) {
            if (sc.objects[k].light) continue;
            cnt.prim_tests++;
            if (check_hit(sc, shadow, k, hitPos, hitNormal, hitDist) && hitDist < shadowDist) {
                angle = 0;
                break;
            }
        }
        phong = add(phong, scale(had(from(o.color), from(l.color)), (angle * l.intensity) / 1));

        V3 spec = mk(0, 0, 0);
        if (o.shine > 0) {
            V3 lightDir = scale(shadowDir, -1);
            V3 specDir = normalize(sub(lightDir, scale(minHitNormal, 2 * dot(minHitNormal, lightDir))));
            float base = (float)std::max(0.0, (double)(-dot(specDir, ray.dir)));
            float p = exp2f(o.specular * log2f(base));      /* __powf */
            spec = scale(scale(scale(mk(1, 1, 1), p), o.shine), angle);
        }
        phong = add(phong, spec);
    }

    V3 refColor = mk(0, 0, 0);
    float kR = o.mirror;
    if (kR > 0) {
        Ray refl = {minHitPos, normalize(sub(ray.dir, scale(minHitNormal, 2 * dot(minHitNormal, ray.dir))))};
        refl.origin = add(refl.origin, scale(refl.dir, (float)0.001));
        refColor = trace(sc, refl, depth + 1, nullptr, cnt);
    }
    return add(scale(refColor, kR), scale(phong, 1 - kR));
}

inline void unpack(unsigned int c, int& r, int& g, int& b) {   /* kernel.cu:34-39 */
    b = (c >> 16) & 255;
    g = (c >> 8) & 255;
    r = c & 255;
}

inline float luminance(unsigned int c) {                        /* kernel.cu:293-298 */
    int r, g, b;
    unpack(c, r, g, b);
    const float c1 = 0.2126729f, c2 = 0.7151522f, c3 = 0.0721750f;
    return std::min(255.0f, r * c1 + g * c2 + b * c3) / 255.0f;
}

/* reference kernel.cu:330-402 for one pixel */
unsigned int fxaa_pixel(const unsigned int* img, int w, int h, int x, int y) {
    const unsigned int centre = img[(size_t)y * w + x];
    if (!(x + 1 < w && y + 1 < h && x > 0 && y > 0)) return centre;

    auto lum = [&](int dx, int dy) { return luminance(img[(size_t)(y + dy) * w + (x + dx)]); };
    float lumE = lum(1, 0), lumW = lum(-1, 0), lumN = lum(0, -1), lumS = lum(0, 1), lumM = lum(0, 0);
    float high = std::max(std::max(std::max(std::max(lumE, lumW), lumN), lumS), lumM);
    float low = std::min(std::min(std::min(std::min(lumE, lumW), lumN), lumS), lumM);
    float contrast = high - low;
    const float contrastThreshold = (float)0.0312, relativeThreshold = (float)0.063;
    float threshold = std::max(contrastThreshold, relativeThreshold * high);
    if (contrast < threshold) return centre;

    float lumNE = lum(1, -1), lumNW = lum(-1, -1), lumSE = lum(1, 1), lumSW = lum(-1, 1);
    float filter = (float)((2 * (lumE + lumW + lumS + lumN) + lumNE + lumNW + lumSE + lumSW) / 12.0);
    filter = std::min(1.0f, std::fabs(filter - lumM) / contrast);
    filter = filter * filter * (3.0f - 2.0f * filter);
    float b = filter;

    float hor = std::fabs(lumN + lumS - 2 * lumM) * 2 + std::fabs(lumNE + lumSE - 2 * lumE) +
                std::fabs(lumNW + lumSW - 2 * lumW);
    float ver = std::fabs(lumE + lumW - 2 * lumM) * 2 + std::fabs(lumNE + lumNW - 2 * lumN) +
                std::fabs(lumSE + lumSW - 2 * lumS);
    int dy = 0, dx = 0;
    if (hor >= ver) dy = std::fabs(lumN - lumM) >= std::fabs(lumS - lumM) ? -1 : 1;
    else            dx = std::fabs(lumE - lumM) >= std::fabs(lumW - lumM) ? 1 : -1;

    int pr, pg, pb, sr, sg, sb;
    unpack(centre, pr, pg, pb);
    unpack(img[(size_t)(y + dy) * w + (x + dx)], sr, sg, sb);
    float px = (float)pr, py = (float)pg, pz = (float)pb;
    return pack_rgb(sr * b + px * (1 - b), sg * b + py * (1 - b), sb * b + pz * (1 - b));
}

}  // namespace

extern "C" {

/* counters_out[5]: primary rays, shadow rays, mirror rays, primitive tests, sky lookups */
int oracle_render(const rtb_Object* objects, int n_objects, const rtb_Light* lights, int n_lights,
                  const rtb_Camera* cam, const float* ambient, const float* skyVars,
                  const unsigned char* const* tex, int texw, int texh, float dayTime, int alias,
                  int imgw, int imgh, int x_step, int y_step,
                  unsigned int* out, unsigned int* base_out, int* hit_ids,
                  unsigned long long* counters_out) {
    Scene sc;
    sc.objects = objects;
    sc.n_objects = n_objects;
    sc.lights = lights;
    sc.n_lights = n_lights;
    sc.ambient = mk(ambient[0], ambient[1], ambient[2]);
    for (int i = 0; i < 4; i++) { sc.skyVars[i] = skyVars[i]; sc.tex[i] = tex[i]; }
    sc.texw = texw;
    sc.texh = texh;
    sc.dayTime = dayTime;
    if (x_step < 1) x_step = 1;
    if (y_step < 1) y_step = 1;

    unsigned int* base = base_out;
    unsigned int* tmp = nullptr;
    if (!base) { tmp = new unsigned int[(size_t)imgw * imgh]; base = tmp; }
    if (x_step > 1 || y_step > 1) memset(base, 0, sizeof(unsigned int) * (size_t)imgw * imgh);

    unsigned long long c0 = 0, c1 = 0, c2 = 0, c3 = 0, c4 = 0;
    const V3 LD = from(cam->LD), LU = from(cam->LU), RD = from(cam->RD), RU = from(cam->RU);
#pragma omp parallel for schedule(dynamic, 4) reduction(+ : c0, c1, c2, c3, c4)
    for (int y = 0; y < imgh; y += y_step) {
        Counters cnt = {0, 0, 0, 0, 0};
        for (int x = 0; x < imgw; x += x_step) {
            /* reference kernel.cu:244-257 */
            float partX = (float)x / (float)(imgw - 1);
            float partY = (float)y / (float)(imgh - 1);
            V3 vd = add(LD, scale(sub(RD, LD), partX));
            V3 vu = add(LU, scale(sub(RU, LU), partX));
            V3 target = sub(vu, scale(sub(vu, vd), partY));
            Ray ray = {from(cam->pos), normalize(target)};
            int id = -1;
            V3 c = scale(trace(sc, ray, 0, &id, cnt), 255);
            base[(size_t)y * imgw + x] = pack_rgb(c.x, c.y, c.z);
            if (hit_ids) hit_ids[(size_t)y * imgw + x] = id;
        }
        c0 += cnt.rays_primary; c1 += cnt.rays_shadow; c2 += cnt.rays_mirror;
        c3 += cnt.prim_tests; c4 += cnt.sky_lookups;
    }
    if (counters_out) {
        counters_out[0] = c0; counters_out[1] = c1; counters_out[2] = c2;
        counters_out[3] = c3; counters_out[4] = c4;
    }
    if (out) {
        if (alias) {
#pragma omp parallel for schedule(static)
            for (int y = 0; y < imgh; y++)
                for (int x = 0; x < imgw; x++) out[(size_t)y * imgw + x] = fxaa_pixel(base, imgw, imgh, x, y);
        } else {
            memcpy(out, base, sizeof(unsigned int) * (size_t)imgw * imgh);
        }
    }
    delete[] tmp;
    return 0;
}

/* FXAA alone on a given pre-FXAA image (reference kernel.cu:262-403, alias=true) */
int oracle_fxaa(const unsigned int* in, unsigned int* out, int imgw, int imgh) {
#pragma omp parallel for schedule(static)
    for (int y = 0; y < imgh; y++)
        for (int x = 0; x < imgw; x++) out[(size_t)y * imgw + x] = fxaa_pixel(in, imgw, imgh, x, y);
    return 0;
}

/* one primitive test, for known-answer tests */
int oracle_check_hit(const rtb_Object* object, const float* origin, const float* dir,
                     float* hitPos, float* hitNormal, float* hitDist) {
    Scene sc;
    memset(&sc, 0, sizeof(sc));
    sc.objects = object;
    sc.n_objects = 1;
    Ray r = {mk(origin[0], origin[1], origin[2]), mk(dir[0], dir[1], dir[2])};
    V3 p = mk(0, 0, 0), n = mk(0, 0, 0);
    float t = 0;
    bool hit = check_hit(sc, r, 0, p, n, t);
    if (hit) {
        hitPos[0] = p.x; hitPos[1] = p.y; hitPos[2] = p.z;
        hitNormal[0] = n.x; hitNormal[1] = n.y; hitNormal[2] = n.z;
        *hitDist = t;
    }
    return hit ? 1 : 0;
}

/* the primary ray of pixel (x, y) exactly as oracle_render builds it (reference kernel.cu:244-253): near-tie
 * analysis of hit-id differences (tests, tools/near_ties.py) */
void oracle_primary_ray(const rtb_Camera* cam, int imgw, int imgh, float x, float y, float* origin, float* dir) {
    const V3 LD = from(cam->LD), LU = from(cam->LU), RD = from(cam->RD), RU = from(cam->RU);
    float partX = x / (float)(imgw - 1);       /* x, y may be fractional: sub-pixel probes around a pixel centre */
    float partY = y / (float)(imgh - 1);
    V3 vd = add(LD, scale(sub(RD, LD), partX));
    V3 vu = add(LU, scale(sub(RU, LU), partX));
    V3 d = normalize(sub(vu, scale(sub(vu, vd), partY)));
    origin[0] = cam->pos.x; origin[1] = cam->pos.y; origin[2] = cam->pos.z;
    dir[0] = d.x; dir[1] = d.y; dir[2] = d.z;
}

}  // extern "C"
