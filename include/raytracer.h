// include/raytracer.h — RayTracer of the reference (include/raytracer.h:17-47) with the same public interface, on top
// of the C ABI (rt_capi.h).  The OpenCL/OpenGL members (cl::Kernel, cl::ImageGL, GL texture) are replaced by one
// rt_ctx; the image lives in a headless device buffer instead of a GL texture shared with OpenCL.
#ifndef raytracer_h
#define raytracer_h

#include "camera.h"
#include "rt_capi.h"
#include "scene.h"

class Screen;   // presentation (src/screen.cpp) needs a graphics stack: unavailable on this box

class RayTracer {
private:
    int width, height;
    cl_uint sample_counter;
    rt_ctx* ctx;
    SceneCreator scene;

    void createCLBuffers();     // random table + scene + textures, src/raytracer.cpp:63-100
    [[noreturn]] void processError(const std::string& what) const;   // print + exit(-1), src/kernelgl.cpp:47-56

public:
    // kernel_path is accepted for source compatibility and ignored (the kernels are compiled into librtb200.so).
    // Like the reference the constructor loads "assets/scenes/scene.scene" relative to the working directory;
    // the environment variables RT_SCENE_FILE / RT_ASSET_ROOT / RT_TEXTURE_DIR / RT_TABLE_SEED / RT_MAX_DEPTH /
    // RT_ARITH (strict|fast) / RT_RNG (table|counter) / RT_DEVICE override the hard-coded choices.
    RayTracer(int w, int h, const char* kernel_path);
    ~RayTracer();

    void render(const Camera* camera);        // src/raytracer.cpp:127-144
    void renderAgain(const Camera* camera);   // src/raytracer.cpp:146-165
    void transferImage(Screen* screen, const char* shader_tex_id);   // GL only: a no-op here
    void setTime(float time);                 // declared but never defined in the reference: no-op
    void resize(int w, int h);                // declared but never defined in the reference: re-creates the buffers

    // headless additions
    void readImage(float* rgba_w_h_4) const;  // (sqrt(mean), 1), row 0 = bottom row on screen
    cl_uint sampleCounter() const { return sample_counter; }
    rt_ctx* context() const { return ctx; }
};

#endif /* raytracer_h */
