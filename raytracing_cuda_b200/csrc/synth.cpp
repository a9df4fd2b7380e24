/*
 * synth.cpp -- synthetic inputs with no reference counterpart.
 *
 *  rtb_terrain_build : BASELINE.json config 4 -- a tessellated fBm height field
 *      (~1M triangles at cells=708) over the built-in scene's footprint, to
 *      exercise the LBVH build and traversal at a size the reference's
 *      133-object __constant__ array (reference kernel.cu:14) cannot hold.
 *      Triangles are wound like the island top (reference scene.cpp:369-373)
 *      so they face up under the reference's back-face culling
 *      (reference kernel.cu:104); material = the mountain material
 *      (reference scene.cpp:338-342).
 *  rtb_sky_synthetic : asset-free equirectangular sky maps of any size, used
 *      where the reference's backgrounds/*.png (reference scene.cpp:626-632) are not
 *      available (the GPU box).  Integer arithmetic only: byte-identical everywhere.
 */
#include <cmath>
#include <cstdint>
#include <cstring>

#include "../../include/rtb_api.h"
#include "vecmath.h"

namespace {

inline uint32_t hash32(uint32_t x) {
    x ^= x >> 16; x *= 0x7feb352dU;
    x ^= x >> 15; x *= 0x846ca68bU;
    x ^= x >> 16;
    return x;
}
inline uint32_t hash3(uint32_t a, uint32_t b, uint32_t c) {
    return hash32(a * 0x9E3779B1U ^ hash32(b * 0x85EBCA77U ^ hash32(c + 0xC2B2AE3DU)));
}
inline float lattice(int ix, int iz, uint32_t seed) {
    return (float)(hash3((uint32_t)ix, (uint32_t)iz, seed) >> 8) * (1.0f / 16777216.0f);
}
inline float smooth(float t) { return t * t * (3.0f - 2.0f * t); }

float value_noise(float x, float z, uint32_t seed) {
    float fx = floorf(x), fz = floorf(z);
    int ix = (int)fx, iz = (int)fz;
    float tx = smooth(x - fx), tz = smooth(z - fz);
    float a = lattice(ix, iz, seed), b = lattice(ix + 1, iz, seed);
    float c = lattice(ix, iz + 1, seed), d = lattice(ix + 1, iz + 1, seed);
    float ab = a + (b - a) * tx, cd = c + (d - c) * tx;
    return ab + (cd - ab) * tz;
}

float terrain_height(float x, float z, uint32_t seed) {
    float amp = 0.5f, freq = 1.0f / 420.0f, sum = 0.0f;
    for (int o = 0; o < 5; o++) {
        sum += amp * value_noise(x * freq + 37.0f * o, z * freq - 19.0f * o, seed + 101u * o);
        amp *= 0.5f;
        freq *= 2.0f;
    }
    float n = sum / 0.96875f;           /* 0..1 */
    n = n * n * (3.0f - 2.0f * n);      /* sharpen ridges / flatten valleys */
    return -10.0f + 310.0f * n;
}

}  // namespace

extern "C" {

int rtb_terrain_build(int cells, unsigned int seed, rtb_Object* out, int capacity) {
    if (cells < 1) return 0;
    const int total = 2 * cells * cells + 3;
    if (!out) return total;
    if (capacity < total) return -1;

    const float x0 = -600.0f, x1 = 880.0f, z0 = -645.0f, z1 = 790.0f;
    const float dx = (x1 - x0) / cells, dz = (z1 - z0) / cells;
    const vec3 rock = vscale(vscale(v3(75, 111, 255), (float)(1.0 / 255)), 0.4f);
    const vec3 snow = v3(0.85f, 0.87f, 0.95f);

    memset(out, 0, sizeof(rtb_Object) * (size_t)total);
    rtb_Object& water = out[0];
    water.type = RTB_PLANE;
    water.color = vscale(vscale(v3(20, 143, 248), (float)(1.0 / 255)), 0.9f);
    water.mirror = 0.6f;
    water.specular = 256;
    water.pos = v3(0, -4.5f, 0);
    water.size = v3(0, 1, 0);

    int k = 1;
    for (int j = 0; j < cells; j++) {
        for (int i = 0; i < cells; i++) {
            float xa = x0 + dx * i, xb = x0 + dx * (i + 1);
            float za = z0 + dz * j, zb = z0 + dz * (j + 1);
            vec3 p0 = v3(xa, terrain_height(xa, za, seed), za);
            vec3 p1 = v3(xb, terrain_height(xb, za, seed), za);
            vec3 p2 = v3(xb, terrain_height(xb, zb, seed), zb);
            vec3 p3 = v3(xa, terrain_height(xa, zb, seed), zb);
            const vec3 tri[2][3] = {{p0, p2, p1}, {p0, p3, p2}};
            for (int t = 0; t < 2; t++) {
                rtb_Object& o = out[k++];
                o.type = RTB_TRIANGLE;
                o.shine = 0;
                o.specular = 256;
                o.mirror = 0;
                float hmean = (tri[t][0].y + tri[t][1].y + tri[t][2].y) * (1.0f / 3.0f);
                float s = clampf((hmean - 150.0f) * (1.0f / 100.0f), 0.0f, 1.0f);
                o.color = vadd(vscale(rock, 1.0f - s), vscale(snow, s));
                o.pos = tri[t][0];
                o.size = tri[t][1];
                o.third = tri[t][2];
            }
        }
    }
    /* emissive sun / moon discs at the 14:00 positions (reference scene.cpp:433-442, 758-771) */
    const vec3 discs[2] = {v3(-500, 500, 500), v3(-500, -500, 500)};
    const vec3 cols[2] = {v3(1, 0.8f, 0.05f), v3(0.9f, 0.9f, 1)};
    for (int d = 0; d < 2; d++) {
        rtb_Object& o = out[k++];
        o.type = RTB_SPHERE;
        o.color = cols[d];
        o.pos = discs[d];
        o.size = v3(50, 50, 50);
        o.light = 1;
    }
    return k;
}

void rtb_sky_synthetic(int which, int w, int h, unsigned char* out) {
    /* zenith / horizon colours per time of day */
    static const int zen[4][3] = {{90, 110, 190}, {40, 110, 230}, {70, 50, 120}, {4, 6, 22}};
    static const int hor[4][3] = {{250, 170, 150}, {170, 210, 250}, {255, 140, 60}, {20, 26, 60}};
    which &= 3;
#pragma omp parallel for schedule(static)
    for (int y = 0; y < h; y++) {
        /* y = 0 is straight up (reference kernel.cu:156: v = 1 - (asin(d.y)+PI/2)/PI) */
        int elev = (int)(((int64_t)(h / 2 - y) * 1024) / (h / 2 > 0 ? h / 2 : 1));   /* -1024..1024 */
        int a = elev < 0 ? 0 : elev;                                                 /* below horizon = horizon colour */
        unsigned char* row = out + (size_t)y * w * 4;
        for (int x = 0; x < w; x++) {
            uint32_t hcoarse = hash3((uint32_t)(x >> 6), (uint32_t)(y >> 6), 77u + which);
            uint32_t hfine = hash3((uint32_t)x, (uint32_t)y, 1234u + which);
            int cloud = (int)(hcoarse & 15) - 8;
            int grain = (int)(hfine & 3) - 1;
            int rgb[3];
            for (int c = 0; c < 3; c++) {
                int v = (hor[which][c] * (1024 - a) + zen[which][c] * a) >> 10;
                v += cloud + grain;
                rgb[c] = v < 0 ? 0 : (v > 255 ? 255 : v);
            }
            if (which == 3 && elev > 0 && (hfine >> 8) % 1499u == 0) {   /* stars */
                int br = 150 + (int)((hfine >> 20) & 105);
                rgb[0] = rgb[1] = rgb[2] = br;
            }
            row[4 * x + 0] = (unsigned char)rgb[0];
            row[4 * x + 1] = (unsigned char)rgb[1];
            row[4 * x + 2] = (unsigned char)rgb[2];
            row[4 * x + 3] = 255;
        }
    }
}

}  // extern "C"
