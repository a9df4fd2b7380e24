/*
 * scene_host.cpp -- host-side scene layer above the launchKernel boundary.
 *
 * Restates what the reference's scene.cpp does to produce the hot path's
 * inputs -- the 133-primitive built-in scene, the camera's frustum corners and
 * the time-of-day state -- with explicit state (struct rtb_scene) instead of
 * file-scope globals.  Every routine cites the reference lines it follows; the
 * float/double mixing of the original expressions is kept on purpose because
 * tests/test_scene_host.py compares the output bit-for-bit with the
 * reference's own scene.cpp compiled under oracle/_ref.
 *
 * Build with -ffp-contract=off (no FMA contraction on the host).
 */
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

#include "../../include/rtb_api.h"
#include "vecmath.h"

namespace {

/* reference scene.cpp:89-91: the PI/180 product is formed in double */
inline float to_rad(float angle) { return (float)((RTB_PI / 180.0f) * angle); }

/* reference transforms.h:7-40: row-dot 3x3 rotations, float cos/sin */
inline vec3 rot_rows(const vec3& v, const vec3& r0, const vec3& r1, const vec3& r2) {
    return v3(vdot(r0, v), vdot(r1, v), vdot(r2, v));
}
inline vec3 rot_y(const vec3& v, float a) {
    return rot_rows(v, v3(cosf(a), 0, sinf(a)), v3(0, 1, 0), v3(-sinf(a), 0, cosf(a)));
}
inline vec3 rot_z(const vec3& v, float a) {
    return rot_rows(v, v3(cosf(a), -sinf(a), 0), v3(sinf(a), cosf(a), 0), v3(0, 0, 1));
}

const float kInv255 = (float)(1.0 / 255);

}  // namespace

struct rtb_scene {
    rtb_Camera cam;
    float aspect;            /* reference scene.cpp:20 */
    float moveSpeed, camViewDelta, camViewLimit, runSpeedUp;
    bool play, antialiasing;
    float seaSpeed;
    char clock[8];
    float dayNightTime, dayNightSpeed, dayNightDistance, dayNightControlSpeed;
    float skyVars[4];
    std::vector<int> vecTree, vecMount, vecLight;
    vec3 ambient;
    vec3 matTree[4], matMount[4], matLake[4], matAmbient[4];
    std::vector<rtb_Object> objects;
    rtb_Light lights[RTB_LIGHTS_NUMBER];
};

namespace {

/* ---- camera: reference scene.cpp:100-126 ---- */
void camera_corners(rtb_scene* s) {
    rtb_Camera& cam = s->cam;
    float dirRad = to_rad(cam.horAngle);
    cam.dir = v3(cosf(dirRad), 0, sinf(dirRad));

    float a = to_rad((float)(cam.fov / 2.0));
    float h = tanf(a);
    float w = h * s->aspect;

    cam.LD = v3(1, -h, -w);
    cam.RD = v3(1, -h, w);
    cam.LU = v3(1, h, -w);
    cam.RU = v3(1, h, w);

    float av = to_rad(-cam.verAngle);
    cam.LD = rot_z(cam.LD, av);
    cam.RD = rot_z(cam.RD, av);
    cam.LU = rot_z(cam.LU, av);
    cam.RU = rot_z(cam.RU, av);

    float ah = to_rad(-cam.horAngle);
    cam.LD = rot_y(cam.LD, ah);
    cam.RD = rot_y(cam.RD, ah);
    cam.LU = rot_y(cam.LU, ah);
    cam.RU = rot_y(cam.RU, ah);
}

/* ---- builders: reference scene.cpp:177-442 ---- */
struct Builder {
    rtb_scene* s;
    void push(const rtb_Object& o) { s->objects.push_back(o); }
    int count() const { return (int)s->objects.size(); }

    static rtb_Object blank() {
        rtb_Object o;
        memset(&o, 0, sizeof(o));
        return o;
    }

    /* reference scene.cpp:177-187 */
    void sphere(vec3 color, float mirror, float specular, float shine, vec3 pos, float size) {
        rtb_Object o = blank();
        o.type = RTB_SPHERE;
        o.color = color;
        o.mirror = mirror;
        o.specular = specular;
        o.shine = shine;
        o.pos = pos;
        o.size = v3(size, size, size);
        o.light = 0;
        push(o);
    }

    /* reference scene.cpp:189-243 */
    void snowman(vec3 offset, float a) {
        vec3 color = vscale(v3(1, 1, 1), 0.8f);
        const float mirror = 0, specular = 1, shine = 0.05f;
        sphere(color, mirror, specular, shine, vadd(rot_y(v3(0, 0, 0), a), offset), 2);       /* belly */
        sphere(color, mirror, specular, shine, vadd(rot_y(v3(0, 3, 0), a), offset), 1.3f);    /* head */
        color = v3(0, 0, 0);
        const float mouthY2 = (float)(2.3 + 0.2);
        const struct { float x, y, z, r; } dots[] = {
            {0.35f, 3.2f, 1.15f, 0.2f},  {-0.35f, 3.2f, 1.15f, 0.2f},                        /* eyes */
            {0.2f, 2.3f, 1.05f, 0.1f},   {-0.2f, 2.3f, 1.05f, 0.1f},                          /* mouth */
            {0.55f, mouthY2, 1.05f, 0.1f}, {-0.55f, mouthY2, 1.05f, 0.1f},
            {0, 1, 1.6f, 0.2f}, {0, 0.3f, 1.85f, 0.2f}, {0, -0.5f, 1.8f, 0.2f},              /* buttons */
        };
        for (const auto& d : dots)
            sphere(color, mirror, specular, shine, vadd(rot_y(v3(d.x, d.y, d.z), a), offset), d.r);
    }

    /* reference scene.cpp:245-296: 4-triangle pyramid over an (almost) equilateral base */
    void pyramid(vec3 color, float mirror, float specular, float shine, vec3 pos, float base,
                 float height, float angle) {
        const float y = 0.86f, x = 0.5f;
        const float v = y * 1.0f / 3.0f;
        const float t = 0.5f;
        const vec3 c0 = v3(0, 0, 0), c1 = v3(1, 0, 0), c2 = v3(x, 0, y), apex = v3(x, t, v);
        vec3 tris[12] = {
            c0, c1, c2,        /* down  */
            c0, apex, c1,      /* front */
            c0, c2, apex,      /* left  */
            c2, c1, apex,      /* right */
        };
        for (int k = 0; k < 12; k++) {
            tris[k].x -= x;
            tris[k].z -= v;
        }
        for (int k = 0; k < 12; k++) {
            tris[k] = rot_y(tris[k], to_rad(angle));
            tris[k].x *= base;
            tris[k].y *= height;
            tris[k].z *= base;
            tris[k] = vadd(tris[k], pos);
        }
        for (int k = 0; k < 4; k++) {
            rtb_Object o = blank();
            o.type = RTB_TRIANGLE;
            o.shine = shine;
            o.specular = specular;
            o.mirror = mirror;
            o.color = color;
            o.pos = tris[k * 3];
            o.size = tris[k * 3 + 1];
            o.third = tris[k * 3 + 2];
            o.light = 0;
            push(o);
        }
    }

    void last_four_into(std::vector<int>& v) {
        int i = count();
        v.push_back(i - 1);
        v.push_back(i - 2);
        v.push_back(i - 3);
        v.push_back(i - 4);
    }

    /* reference scene.cpp:298-324 */
    void tree(vec3 offset, float angle) {
        vec3 crown = vscale(vscale(v3(100, 80, 200), 1.0f / 255.0f), 0.8f);
        vec3 trunk = v3(0.5f, 0, 0);
        const float mirror = 0.1f, specular = 1, shine = 0;
        pyramid(crown, mirror, specular, shine, vadd(v3(0, -1, 0), offset), 7, 19, angle);
        last_four_into(s->vecTree);
        pyramid(trunk, mirror, specular, shine, vadd(v3(0, -2, 0), offset), 4, 8, angle);
    }

    /* reference scene.cpp:326-336: the water plane; pos.y is the sea level */
    void ground(vec3 offset) {
        rtb_Object o = blank();
        o.type = RTB_PLANE;
        o.color = vscale(v3(0, 0, 30), 1.0f / 255.0f);
        o.mirror = 0.6f;
        o.specular = 256;
        o.shine = 0;
        o.pos = offset;
        o.size = v3(0, 1, 0);
        o.light = 0;
        push(o);
    }

    /* reference scene.cpp:338-350 */
    void mountain(vec3 offset, float size, float angle) {
        vec3 color = vscale(vscale(v3(18, 31, 60), 1.0f / 255.0f), 0.4f);
        pyramid(color, 0, 256, 0, offset, size, 1.5f * size, angle);
        last_four_into(s->vecMount);
    }

    /* reference scene.cpp:352-414: open-bottom slab, 10 triangles */
    void island(vec3 offset, float size, float d) {
        vec3 color = vscale(vscale(v3(100, 80, 200), 1.0f / 255.0f), 0.8f);
        const float mirror = 0.1f, specular = 1, shine = 0;
        const vec3 p[8] = {
            v3(0, 0, 0),  v3(1, 0, 0),  v3(1, 0, 1),  v3(0, 0, 1),
            v3(0, -d, 0), v3(1, -d, 0), v3(1, -d, 1), v3(0, -d, 1),
        };
        static const int idx[30] = {
            0, 2, 1, 0, 3, 2,   /* top    */
            4, 1, 5, 4, 0, 1,   /* front  */
            6, 3, 7, 6, 2, 3,   /* back   */
            5, 2, 6, 5, 1, 2,   /* right  */
            7, 0, 4, 7, 3, 0,   /* left   */
        };
        vec3 tris[30];
        for (int k = 0; k < 30; k++) {
            vec3 q = p[idx[k]];
            q.x = (float)(q.x - 0.5);
            q.z = (float)(q.z - 0.5);
            q.x *= size;
            q.y *= 1;
            q.z *= size;
            tris[k] = vadd(q, offset);
        }
        for (int k = 0; k < 10; k++) {
            s->vecTree.push_back(count());
            rtb_Object o = blank();
            o.type = RTB_TRIANGLE;
            o.shine = shine;
            o.specular = specular;
            o.mirror = mirror;
            o.color = color;
            o.pos = tris[k * 3];
            o.size = tris[k * 3 + 1];
            o.third = tris[k * 3 + 2];
            push(o);
        }
    }

    /* reference scene.cpp:416-431 */
    void igloo(vec3 offset, float size1, float size2) {
        vec3 color = vscale(v3(1, 1, 1), 0.8f);
        sphere(color, 0, 1, 0.05f, vadd(v3(0, 0, 0), offset), size1);
        sphere(color, 0, 1, 0.05f, vadd(v3(-6, 0, 6), offset), size2);
    }

    /* reference scene.cpp:433-442: emissive sun and moon discs */
    void light_objects() {
        sphere(v3(1, 0.8f, 0.05f), 0, 0, 0, s->lights[0].pos, 50);
        s->objects.back().light = 1;
        s->vecLight.push_back(count() - 1);
        sphere(v3(0.9f, 0.9f, 1), 0, 0, 0, s->lights[1].pos, 50);
        s->objects.back().light = 1;
        s->vecLight.push_back(count() - 1);
    }
};

/* reference scene.cpp:444-488 -- order fixes the object indices */
void build_objects(rtb_scene* s) {
    Builder b{s};
    s->objects.clear();
    s->objects.reserve(RTB_OBJECTS_NUMBER);
    const float level = -4.5f;

    b.ground(v3(0, level, 0));
    b.island(v3(0, -4, 0), 50, 2);

    b.snowman(v3(-4, -2, 17), to_rad(-50));
    b.snowman(v3(-15, -2, 5), to_rad(-20));

    b.tree(v3(-22, -2, -10), 90);
    b.tree(v3(-10, -2, -20), 90);
    b.tree(v3(0, -2, -20), 80);
    b.tree(v3(17, -2, 2), 90);
    b.tree(v3(20, -2, 9), 80);
    b.tree(v3(12, -2, 22), 70);

    const float d = 4;
    const struct { float x, z, size, angle; } m[12] = {
        {170, 0, 100, 0},   {90, -100, 110, 45}, {-35, -90, 100, 0}, {-100, 65, 100, 0},
        {25, 140, 100, 0},  {130, 90, 100, 0},   {100, 30, 70, 0},   {100, -40, 50, 30},
        {20, -100, 70, 0},  {-80, -40, 80, 0},   {-70, 100, 90, 0},  {60, 90, 50, 0},
    };
    for (const auto& q : m) b.mountain(vscale(v3(q.x, level, q.z), d), q.size * d, q.angle);

    b.igloo(v3(4, -4, -4), 10, 6);
    b.light_objects();
}

/* reference scene.cpp:666-672 */
vec3 color_by_time(const rtb_scene* s, const vec3* mats) {
    vec3 c = v3(0, 0, 0);
    for (int i = 0; i < 4; i++) c = vadd(c, vscale(mats[i], s->skyVars[i]));
    return c;
}

/* reference scene.cpp:674-687 */
void recolor_objects(rtb_scene* s) {
    for (int i : s->vecTree) s->objects[i].color = color_by_time(s, s->matTree);
    for (int i : s->vecMount) s->objects[i].color = color_by_time(s, s->matMount);
    s->objects[0].color = color_by_time(s, s->matLake);
    s->ambient = color_by_time(s, s->matAmbient);
}

/* reference scene.cpp:778-804: plateaus 6-8, 10-16, 18-20, 22-4 with 2 h linear cross-fades */
void calc_sky_vars(rtb_scene* s) {
    float* k = s->skyVars;
    for (int i = 0; i < 4; i++) k[i] = 0;
    const float d = s->dayNightTime;
    if (d >= 6 && d <= 8) k[0] = 1;
    if (d >= 10 && d <= 16) k[1] = 1;
    if (d >= 18 && d <= 20) k[2] = 1;
    if (d >= 22 || d <= 4) k[3] = 1;
    if (d > 8 && d < 10)  { k[1] = (d - 8) / 2;  k[0] = 1.0f - k[1]; }
    if (d > 16 && d < 18) { k[2] = (d - 16) / 2; k[1] = 1.0f - k[2]; }
    if (d > 20 && d < 22) { k[3] = (d - 20) / 2; k[2] = 1.0f - k[3]; }
    if (d > 4 && d < 6)   { k[0] = (d - 4) / 2;  k[3] = 1.0f - k[0]; }
}

/* Sun and moon (reference scene.cpp:758-776) as two ends of one arm: the arm turns once a day in a circle of radius
 * dayNightDistance that is tilted by -45 degrees about the vertical and centred on (-500, 0, 500); noon-ish phase
 * offset -120 degrees.  Both point lights shine with |height of the sun| / radius, and the two emissive spheres
 * ride along.  Float expression shapes are the reference's (x * -1 is an exact negation, (-p) + c == c + (-p)), so the
 * states stay bit-identical to its compiled scene.cpp (tests/test_scene_host.py). */
void move_lights(rtb_scene* s) {
    const vec3 centre = v3(-500.0f, 0.0f, 500.0f);
    const float phase = to_rad(fmodf((s->dayNightTime / 24.0f) * 360 - 120, 360));
    const vec3 arm = rot_y(vscale(v3(cosf(phase), sinf(phase), 0), s->dayNightDistance), to_rad(-45));
    const vec3 body[2] = {vadd(arm, centre), vadd(vscale(arm, -1), centre)};
    const float glow = fabsf(body[0].y) / s->dayNightDistance;
    for (int k = 0; k < 2; k++) {
        s->lights[k].pos = body[k];
        s->lights[k].color = v3(glow, glow, glow);
        s->objects[s->vecLight[k]].pos = body[k];
    }
}

/* "HH:MM\n" of the day clock; minutes from the truncated hundredths of an hour, as reference scene.cpp:732 prints */
void write_clock(rtb_scene* s) {
    const int hours = (int)s->dayNightTime;
    const int hundredths = (int)(s->dayNightTime * 100) % 100;
    const int minutes = (int)(hundredths / 100.0 * 60);
    snprintf(s->clock, sizeof(s->clock), "%02d:%02d\n", hours, minutes);
}

}  // namespace

/* ------------------------------------------------------------- C-ABI */

extern "C" {

rtb_scene* rtb_scene_create(void) {
    rtb_scene* s = new rtb_scene();
    /* globals of reference scene.cpp:13-67 */
    s->moveSpeed = 50;
    s->camViewDelta = 0.02f;
    s->camViewLimit = 44;
    s->runSpeedUp = 2;
    s->aspect = 1.7777f;
    s->play = true;
    s->antialiasing = true;
    s->seaSpeed = 2;
    memset(s->clock, 0, sizeof(s->clock));
    s->dayNightTime = 6;
    s->dayNightSpeed = (float)(0.5 * 1);
    s->dayNightDistance = 500;
    s->dayNightControlSpeed = 4;
    s->skyVars[0] = 0; s->skyVars[1] = 0; s->skyVars[2] = 0; s->skyVars[3] = 1;
    s->ambient = v3(0.1f, 0.2f, 0.4f);

    s->matTree[0] = vscale(v3(158, 114, 250), kInv255);
    s->matTree[1] = vscale(v3(218, 222, 255), kInv255);
    s->matTree[2] = vscale(v3(255, 166, 82), kInv255);
    s->matTree[3] = v3(0.31f, 0.25f, 0.62f);
    s->matMount[0] = vscale(v3(224, 205, 255), kInv255);
    s->matMount[1] = vscale(v3(75, 111, 255), kInv255);
    s->matMount[2] = vscale(v3(255, 230, 103), kInv255);
    s->matMount[3] = v3(0.02f, 0.04f, 0.09f);
    s->matLake[0] = vscale(v3(155, 4, 136), kInv255);
    s->matLake[1] = vscale(vscale(v3(20, 143, 248), kInv255), 0.9f);
    s->matLake[2] = vscale(v3(255, 20, 20), kInv255);
    s->matLake[3] = v3(0, 0, 0);
    s->matAmbient[0] = vscale(v3(139, 129, 197), kInv255);
    s->matAmbient[1] = vscale(vscale(v3(115, 136, 178), kInv255), 0.7f);
    s->matAmbient[2] = vscale(v3(164, 132, 121), kInv255);
    s->matAmbient[3] = v3(0.1f, 0.2f, 0.4f);

    /* initCamera, reference scene.cpp:165-173 */
    memset(&s->cam, 0, sizeof(s->cam));
    s->cam.pos = v3(-56, 2.2f, 72);
    s->cam.horAngle = 309;
    s->cam.verAngle = -7.07f;
    s->cam.fov = 40;
    camera_corners(s);

    /* initLights, reference scene.cpp:634-652 */
    for (int i = 0; i < RTB_LIGHTS_NUMBER; i++) {
        s->lights[i].color = v3(1, 1, 1);
        s->lights[i].intensity = 1;
        s->lights[i].pos = v3(-1000, 1000, 1000);
    }
    build_objects(s);
    return s;
}

void rtb_scene_destroy(rtb_scene* s) { delete s; }

rtb_Object* rtb_scene_objects(rtb_scene* s, int* n) {
    if (n) *n = (int)s->objects.size();
    return s->objects.data();
}
rtb_Light* rtb_scene_lights(rtb_scene* s, int* n) {
    if (n) *n = RTB_LIGHTS_NUMBER;
    return s->lights;
}
rtb_Camera* rtb_scene_camera(rtb_scene* s) { return &s->cam; }
float* rtb_scene_sky_vars(rtb_scene* s) { return s->skyVars; }
float* rtb_scene_ambient(rtb_scene* s) { return &s->ambient.x; }
float rtb_scene_time(rtb_scene* s) { return s->dayNightTime; }

void rtb_scene_set_time(rtb_scene* s, float hours) {
    s->dayNightTime = hours;
    calc_sky_vars(s);
    recolor_objects(s);
    move_lights(s);
    write_clock(s);
}

int rtb_scene_set_view(rtb_scene* s, int key) {
    if (key == 5) {          /* reference scene.cpp:736-741 */
        s->cam.pos = v3(-56, 2.2f, 72);
        s->cam.horAngle = 309;
        s->cam.verAngle = -7.07f;
    } else if (key == 6) {   /* reference scene.cpp:742-747 */
        s->cam.pos = v3(324.4f, 12.41f, -84);
        s->cam.horAngle = 141.2f;
        s->cam.verAngle = -12.65f;
    } else {
        return RT_ERR_ARG;
    }
    camera_corners(s);
    return RT_OK;
}

void rtb_scene_set_pose(rtb_scene* s, float px, float py, float pz, float horAngle, float verAngle) {
    s->cam.pos = v3(px, py, pz);
    s->cam.horAngle = horAngle;
    s->cam.verAngle = verAngle;
    camera_corners(s);
}

void rtb_scene_update_camera(rtb_scene* s, int imgw, int imgh) {
    s->aspect = (1.0f * imgw) / imgh;   /* reference scene.cpp:821 */
    camera_corners(s);
}

void rtb_scene_set_sea_level(rtb_scene* s, float y) { s->objects[0].pos.y = y; }
void rtb_scene_set_antialiasing(rtb_scene* s, int on) { s->antialiasing = on != 0; }

/* reference scene.cpp:128-140 */
void rtb_scene_mouse(rtb_scene* s, int deltaX, int deltaY) {
    s->cam.horAngle = fmodf(s->cam.horAngle + s->camViewDelta * deltaX + 360.0f, 360.0f);
    s->cam.verAngle = clampf(s->cam.verAngle + s->camViewDelta * deltaY, -s->camViewLimit, s->camViewLimit);
    camera_corners(s);
}

/* reference scene.cpp:806-816: moveCamera, controls, recolorObjects, calcSkyVars, moveLights --
 * in that order, so colours lag the sky weights by one tick exactly like the reference. */
void rtb_scene_animate(rtb_scene* s, float deltaTime, unsigned int keys) {
    auto key = [keys](unsigned int k) { return (keys & k) ? 1 : 0; };

    /* moveCamera, reference scene.cpp:142-163 */
    {
        vec3 camMove = v3(0, 0, 0);
        vec3 camForw = s->cam.dir;
        vec3 camUp = v3(0, 1, 0);
        vec3 camSide = v3(-s->cam.dir.z, 0, s->cam.dir.x);
        int verMove = key(RTB_KEY_D) - key(RTB_KEY_A);
        camMove = vadd(camMove, vscale(camSide, (float)verMove));
        int horMove = key(RTB_KEY_W) - key(RTB_KEY_S);
        camMove = vadd(camMove, vscale(camForw, (float)horMove));
        int upMove = key(RTB_KEY_Q) - key(RTB_KEY_E);
        camMove = vadd(camMove, vscale(camUp, (float)upMove));
        float run = key(RTB_KEY_SHIFT) ? s->runSpeedUp : 1;
        if (verMove || horMove || upMove) {
            camMove = vnormalize(camMove);
            s->cam.pos = vadd(s->cam.pos, vscale(vscale(camMove, s->moveSpeed * run), deltaTime));
        }
    }
    /* controls, reference scene.cpp:689-756 */
    {
        int timeControl = key(RTB_KEY_RIGHT) - key(RTB_KEY_LEFT);
        if (timeControl) {
            s->dayNightTime = fmodf(
                s->dayNightTime + s->dayNightSpeed * deltaTime * timeControl * s->dayNightControlSpeed + 24, 24);
        } else if (s->play) {
            s->dayNightTime = fmodf(s->dayNightTime + s->dayNightSpeed * deltaTime + 24, 24);
        }
        if (key(RTB_KEY_P)) s->play = true;
        if (key(RTB_KEY_O)) s->play = false;

        int seaControl = key(RTB_KEY_UP) - key(RTB_KEY_DOWN);
        s->objects[0].pos.y += seaControl * s->seaSpeed * deltaTime;

        bool change = false;
        if (key(RTB_KEY_1)) { s->dayNightTime = 6;  change = true; }
        if (key(RTB_KEY_2)) { s->dayNightTime = 14; change = true; }
        if (key(RTB_KEY_3)) { s->dayNightTime = 18; change = true; }
        if (key(RTB_KEY_4)) { s->dayNightTime = 1;  change = true; }
        if (timeControl || change || s->play) write_clock(s);

        /* the presets set the pose only; corners follow at the next mouse/update call */
        if (key(RTB_KEY_5)) {
            s->cam.pos = v3(-56, 2.2f, 72);
            s->cam.horAngle = 309;
            s->cam.verAngle = -7.07f;
        }
        if (key(RTB_KEY_6)) {
            s->cam.pos = v3(324.4f, 12.41f, -84);
            s->cam.horAngle = 141.2f;
            s->cam.verAngle = -12.65f;
        }
        if (key(RTB_KEY_B)) s->antialiasing = true;
        if (key(RTB_KEY_V)) s->antialiasing = false;
    }
    recolor_objects(s);
    calc_sky_vars(s);
    move_lights(s);
}

const char* rtb_scene_clock(rtb_scene* s) { return s->clock; }

void rtb_scene_frame(rtb_scene* s, int imgw, int imgh, rt_frame* out) {
    memset(out, 0, sizeof(*out));
    out->cam = s->cam;
    out->ambient[0] = s->ambient.x;
    out->ambient[1] = s->ambient.y;
    out->ambient[2] = s->ambient.z;
    for (int i = 0; i < 4; i++) out->skyVars[i] = s->skyVars[i];
    out->dayTime = s->dayNightTime / 24.0f;   /* reference scene.cpp:822 */
    out->alias = s->antialiasing ? 1 : 0;
    out->imgw = imgw;
    out->imgh = imgh;
}

}  // extern "C"
