/* ref_scene_wrap.cpp -- exposes the globals and entry points of the reference's
 * own scene.cpp (compiled from /root/reference/scene.cpp by oracle/build_ref.py,
 * Windows input replaced by ref_shims/windows.h) so tests can compare the
 * package's host scene layer with it field by field.  TEST INFRASTRUCTURE ONLY.
 *
 * The reference's launch() calls launchKernel (reference scene.cpp:823); here
 * that symbol forwards to a function pointer, so a test can route the
 * reference's own frame entry point into the package's drop-in library.
 */
#include <vector_types.h>
#include <structs.h>
#include <scene.h>
#include <vector>

float deltaTime = 1;   /* defined in the reference's main.cpp:53 */

extern Camera cam;
extern float aspect;
extern bool play;
extern bool antialiasing;
extern float dayNightTime;
extern float skyVars[4];
extern float3 ambient;
extern Object* objects;
extern Light* lights;
extern unsigned char *texture1, *texture2, *texture3, *texture4;
extern int texW, texH;

void cameraHelperAngles();
void initCamera();
void initLights();
void initObjects();
void recolorObjects();
void calcSkyVars();
void moveLights();

typedef void (*launch_fn)(unsigned int*, int, int, Camera, Object*, Light*, float3, float*,
                          unsigned char*, unsigned char*, unsigned char*, unsigned char*, int, int,
                          float, bool);
static launch_fn g_forward = nullptr;

extern "C" void launchKernel(unsigned int* result, int imgw, int imgh, Camera c, Object* o, Light* l,
                             float3 amb, float* sky, unsigned char* t1, unsigned char* t2,
                             unsigned char* t3, unsigned char* t4, int w, int h, float dayTime,
                             bool alias) {
    if (g_forward) g_forward(result, imgw, imgh, c, o, l, amb, sky, t1, t2, t3, t4, w, h, dayTime, alias);
}

extern "C" {
void ref_scene_set_forward(void* fn) { g_forward = (launch_fn)fn; }
void ref_scene_init() {   /* initScene() without srand/initTexture (reference scene.cpp:654-662) */
    initCamera();
    initLights();
    initObjects();
}
void ref_scene_set_time(float hours) {   /* steady state: SURVEY 3.2 harness rule */
    dayNightTime = hours;
    calcSkyVars();
    recolorObjects();
    moveLights();
}
void ref_scene_animate(float dt) { deltaTime = dt; animate(); }
void ref_scene_set_pose(float x, float y, float z, float hor, float ver) {
    cam.pos = float3{x, y, z};
    cam.horAngle = hor;
    cam.verAngle = ver;
    cameraHelperAngles();
}
void ref_scene_set_aspect(int w, int h) { aspect = (1.0f * w) / h; cameraHelperAngles(); }
void ref_scene_set_sea(float y) { objects[0].pos.y = y; }
void ref_scene_set_play(int p) { play = p != 0; }
void ref_scene_set_aa(int a) { antialiasing = a != 0; }
void ref_scene_set_textures(unsigned char* t1, unsigned char* t2, unsigned char* t3, unsigned char* t4, int w, int h) {
    texture1 = t1; texture2 = t2; texture3 = t3; texture4 = t4; texW = w; texH = h;
}
void ref_scene_mouse(int x, int y, int ww, int wh) { mouseMotion(x, y, ww, wh); }
void ref_scene_launch(unsigned int* out, int w, int h) { launch(out, w, h); }
Camera* ref_scene_camera() { return &cam; }
Object* ref_scene_objects() { return objects; }
Light* ref_scene_lights() { return lights; }
float* ref_scene_sky_vars() { return skyVars; }
float* ref_scene_ambient() { return &ambient.x; }
float ref_scene_time() { return dayNightTime; }
const char* ref_scene_clock() { return getTime(); }
}
