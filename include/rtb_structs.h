/*
 * rtb_structs.h -- binary data model of the ray-tracing hot path.
 *
 * These PODs are the ABI of the drop-in boundary: a caller built against the
 * reference's structs.h (reference structs.h:8-51) hands arrays of them to
 * launchKernel(), so size, field order and offsets must be identical.  Only the
 * layouts are shared; all arithmetic on them lives in the package's own
 * csrc/vecmath.h.
 *
 *   Camera  108 B   reference structs.h:8-19
 *   Object   68 B   reference structs.h:27-38
 *   Ray      24 B   reference structs.h:40-44
 *   Light    28 B   reference structs.h:46-51
 *
 * Field reuse inside Object (reference kernel.cu:47-98):
 *   sphere   : pos = centre, size.x = radius
 *   plane    : pos = point on plane, size = normal
 *   triangle : pos, size, third = the three vertices (CCW = front face)
 */
#ifndef RTB_STRUCTS_H
#define RTB_STRUCTS_H

#include <stddef.h>

#if defined(__CUDACC__) || defined(__VECTOR_TYPES_H__)
#include <vector_types.h>
typedef float3 rtb_float3;
#else
typedef struct rtb_float3 { float x, y, z; } rtb_float3;
#endif

#ifdef __cplusplus
extern "C" {
#endif

/* reference structs.h:21-25 */
typedef enum rtb_Primitive {
    RTB_SPHERE   = 0,
    RTB_PLANE    = 1,
    RTB_TRIANGLE = 2
} rtb_Primitive;

/* reference structs.h:8-19.  Only pos, LD, LU, RD, RU are read on the device
 * (reference kernel.cu:246-251); left/right are never written by the reference. */
typedef struct rtb_Camera {
    rtb_float3 pos;
    float      horAngle;   /* degrees */
    float      verAngle;   /* degrees */
    float      fov;        /* degrees, vertical */
    rtb_float3 dir;
    rtb_float3 LD, LU, RD, RU;   /* frustum corner vectors, not normalised */
    rtb_float3 left;
    rtb_float3 right;
} rtb_Camera;

/* reference structs.h:27-38 */
typedef struct rtb_Object {
    int        type;       /* rtb_Primitive */
    float      shine;
    float      specular;
    float      mirror;
    rtb_float3 color;
    rtb_float3 pos;
    rtb_float3 size;
    rtb_float3 third;
    unsigned char light;   /* C++ bool in the reference: emissive flag */
    unsigned char _pad[3];
} rtb_Object;

/* reference structs.h:40-44 */
typedef struct rtb_Ray {
    rtb_float3 origin;
    rtb_float3 dir;
} rtb_Ray;

/* reference structs.h:46-51 */
typedef struct rtb_Light {
    rtb_float3 pos;
    rtb_float3 color;
    float      intensity;
} rtb_Light;

#ifdef __cplusplus
}
static_assert(sizeof(rtb_float3) == 12, "float3 must be 12 bytes");
static_assert(sizeof(rtb_Camera) == 108, "Camera ABI");
static_assert(sizeof(rtb_Object) == 68, "Object ABI");
static_assert(sizeof(rtb_Ray) == 24, "Ray ABI");
static_assert(sizeof(rtb_Light) == 28, "Light ABI");
static_assert(offsetof(rtb_Camera, horAngle) == 12 && offsetof(rtb_Camera, dir) == 24 &&
              offsetof(rtb_Camera, LD) == 36 && offsetof(rtb_Camera, LU) == 48 &&
              offsetof(rtb_Camera, RD) == 60 && offsetof(rtb_Camera, RU) == 72 &&
              offsetof(rtb_Camera, left) == 84 && offsetof(rtb_Camera, right) == 96, "Camera offsets");
static_assert(offsetof(rtb_Object, shine) == 4 && offsetof(rtb_Object, specular) == 8 &&
              offsetof(rtb_Object, mirror) == 12 && offsetof(rtb_Object, color) == 16 &&
              offsetof(rtb_Object, pos) == 28 && offsetof(rtb_Object, size) == 40 &&
              offsetof(rtb_Object, third) == 52 && offsetof(rtb_Object, light) == 64, "Object offsets");
static_assert(offsetof(rtb_Light, color) == 12 && offsetof(rtb_Light, intensity) == 24, "Light offsets");
#endif

/* Compile-time constants of the legacy boundary (reference scene.h:11-13). */
#define RTB_OBJECTS_NUMBER 133
#define RTB_LIGHTS_NUMBER  2
#define RTB_MAX_DEPTH      4          /* reference kernel.cu:11 */
#define RTB_PI             3.141592   /* reference kernel.cu:12, scene.h:5 (truncated on purpose) */

#endif /* RTB_STRUCTS_H */
