/*
 * ingest.cpp -- scene ingest: heightmaps and Wavefront OBJ meshes -> the reference's
 * Object array (structs.h:27-38: triangle = pos, size, third), ready for rt_set_scene.
 *
 * The reference builds its geometry in code only (scene.cpp:177-488); these loaders generalise that beyond
 * the built-in 133 primitives.  Triangles keep the reference's conventions: counter-clockwise seen from the
 * front (checkHit culls det < 0.001, kernel.cu:104), heightmap cells split like the island top
 * (scene.cpp:369-373: (p0, p2, p1), (p0, p3, p2)), material fields copied from a template object.
 */
#include <cstdlib>
#include <cstring>
#include <vector>

#include "../../include/rtb_api.h"

namespace {

rtb_Object default_material() {
    rtb_Object m;
    memset(&m, 0, sizeof(m));
    m.type = RTB_TRIANGLE;          /* the mountain material of the reference, scene.cpp:338-342 */
    m.shine = 0;
    m.specular = 256;
    m.mirror = 0;
    m.color.x = 0.5f; m.color.y = 0.5f; m.color.z = 0.55f;
    return m;
}

void emit_triangle(rtb_Object* out, int k, const rtb_Object& mat, const rtb_float3& a, const rtb_float3& b,
                   const rtb_float3& c) {
    rtb_Object o = mat;
    o.type = RTB_TRIANGLE;
    o.pos = a;
    o.size = b;
    o.third = c;
    out[k] = o;
}

/* one OBJ index token "v", "v/vt", "v//vn", "v/vt/vn" -> 0-based vertex index, or -1 */
long obj_index(const char* tok, long n_vertices) {
    char* end = nullptr;
    const long v = strtol(tok, &end, 10);
    if (end == tok || v == 0) return -1;
    const long idx = v > 0 ? v - 1 : n_vertices + v;      /* negative: relative to the last vertex read */
    return (idx >= 0 && idx < n_vertices) ? idx : -1;
}

}  // namespace

extern "C" {

int rtb_heightmap_objects(const float* heights, int nx, int nz, float x0, float x1, float z0, float z1,
                          const rtb_Object* material, rtb_Object* out, int capacity) {
    if (!heights || nx < 2 || nz < 2) return 0;
    const long long total = 2LL * (nx - 1) * (nz - 1);
    if (total > 0x7fffffffLL) return -1;
    if (!out) return (int)total;
    if (capacity < total) return -1;
    const rtb_Object mat = material ? *material : default_material();
    const float dx = (x1 - x0) / (float)(nx - 1), dz = (z1 - z0) / (float)(nz - 1);
    int k = 0;
    for (int j = 0; j + 1 < nz; j++) {
        for (int i = 0; i + 1 < nx; i++) {
            const float xa = x0 + dx * i, xb = x0 + dx * (i + 1);
            const float za = z0 + dz * j, zb = z0 + dz * (j + 1);
            const rtb_float3 p0 = {xa, heights[(size_t)j * nx + i], za};
            const rtb_float3 p1 = {xb, heights[(size_t)j * nx + i + 1], za};
            const rtb_float3 p2 = {xb, heights[(size_t)(j + 1) * nx + i + 1], zb};
            const rtb_float3 p3 = {xa, heights[(size_t)(j + 1) * nx + i], zb};
            emit_triangle(out, k++, mat, p0, p2, p1);
            emit_triangle(out, k++, mat, p0, p3, p2);
        }
    }
    return k;
}

int rtb_obj_parse(const char* text, size_t len, const rtb_Object* material, rtb_Object* out, int capacity) {
    if (!text) return 0;
    const rtb_Object mat = material ? *material : default_material();
    std::vector<rtb_float3> verts;
    std::vector<long> face;
    long long count = 0;
    size_t pos = 0;
    while (pos < len) {
        size_t eol = pos;
        while (eol < len && text[eol] != '\n') eol++;
        std::vector<char> line(text + pos, text + eol);
        line.push_back(0);
        pos = eol + 1;
        char* p = line.data();
        while (*p == ' ' || *p == '\t') p++;
        if (p[0] == 'v' && (p[1] == ' ' || p[1] == '\t')) {
            char* q = p + 1;
            rtb_float3 v;
            v.x = strtof(q, &q);
            v.y = strtof(q, &q);
            v.z = strtof(q, &q);
            verts.push_back(v);
        } else if (p[0] == 'f' && (p[1] == ' ' || p[1] == '\t')) {
            face.clear();
            char* q = p + 1;
            for (;;) {
                while (*q == ' ' || *q == '\t' || *q == '\r') q++;
                if (!*q) break;
                const long idx = obj_index(q, (long)verts.size());
                if (idx < 0) return -2;                   /* malformed or out-of-range index */
                face.push_back(idx);
                while (*q && *q != ' ' && *q != '\t' && *q != '\r') q++;
            }
            for (size_t t = 2; t < face.size(); t++) {    /* fan triangulation, winding kept */
                if (out) {
                    if (count >= capacity) return -1;
                    emit_triangle(out, (int)count, mat, verts[face[0]], verts[face[t - 1]], verts[face[t]]);
                }
                count++;
            }
        }
    }
    return count > 0x7fffffffLL ? -1 : (int)count;
}

}  // extern "C"
