// host/raytracer.cpp — RayTracer (src/raytracer.cpp) on the C ABI.
#include "../../include/raytracer.h"

#include <cstdio>
#include <cstdlib>
#include <random>
#include <vector>

namespace {
const char* env_or(const char* name, const char* dflt) { const char* v = std::getenv(name); return (v && *v) ? v : dflt; }
}  // namespace

void RayTracer::processError(const std::string& what) const {
    std::fprintf(stderr, "%s\n", what.c_str());
    std::exit(-1);
}

RayTracer::RayTracer(int w, int h, const char* /*kernel_path*/) : width(w), height(h), sample_counter(0), ctx(nullptr) {
    if (rt_create(w, h, std::atoi(env_or("RT_DEVICE", "0")), &ctx) != RT_OK) processError(std::string("ERROR: ") + rt_last_error());
    try { createCLBuffers(); }
    catch (const SceneError& e) { processError(e.what()); }
}

RayTracer::~RayTracer() { rt_destroy(ctx); }

void RayTracer::createCLBuffers() {
    // random table, src/raytracer.cpp:67-93 (std::random_device unless RT_TABLE_SEED pins it)
    std::vector<float> table(RT_RANDOM_TABLE_FLOATS);
    const char* seed = std::getenv("RT_TABLE_SEED");
    rt_make_random_table(seed ? (uint32_t)std::strtoul(seed, nullptr, 10) : (uint32_t)std::random_device{}(), table.data());
    if (rt_set_random_table(ctx, table.data(), RT_RANDOM_TABLE_FLOATS) != RT_OK) processError(std::string("ERROR: ") + rt_last_error());

    rt_params prm;
    rt_get_params(ctx, &prm);
    prm.max_depth = (uint32_t)std::atoi(env_or("RT_MAX_DEPTH", "30"));
    prm.arith = std::string(env_or("RT_ARITH", "fast")) == "strict" ? RT_ARITH_STRICT : RT_ARITH_FAST;
    prm.rng_mode = std::string(env_or("RT_RNG", "table")) == "counter" ? RT_RNG_COUNTER : RT_RNG_TABLE;
    if (rt_set_params(ctx, &prm) != RT_OK) processError(std::string("ERROR: ") + rt_last_error());

    scene.setAssetRoot(env_or("RT_ASSET_ROOT", ""));
    scene.setTextureDir(env_or("RT_TEXTURE_DIR", ""));
    scene.loadScene(env_or("RT_SCENE_FILE", "assets/scenes/scene.scene"));   // hard-coded in the reference (:95)
    scene.loadTextures();
    scene.createScene(ctx);
}

void RayTracer::render(const Camera* camera) {
    sample_counter = 0;
    if (rt_render(ctx, camera->transferData()) != RT_OK) processError(std::string("ERROR: ") + rt_last_error());
}

void RayTracer::renderAgain(const Camera* camera) {
    sample_counter++;
    if (rt_render_again(ctx, camera->transferData()) != RT_OK) processError(std::string("ERROR: ") + rt_last_error());
}

void RayTracer::transferImage(Screen*, const char*) {}
void RayTracer::setTime(float) {}

void RayTracer::resize(int w, int h) {
    rt_destroy(ctx);
    ctx = nullptr;
    width = w; height = h; sample_counter = 0;
    if (rt_create(w, h, std::atoi(env_or("RT_DEVICE", "0")), &ctx) != RT_OK) processError(std::string("ERROR: ") + rt_last_error());
    scene = SceneCreator();
    try { createCLBuffers(); }
    catch (const SceneError& e) { processError(e.what()); }
}

void RayTracer::readImage(float* rgba) const {
    if (rt_read_image(ctx, rgba) != RT_OK) processError(std::string("ERROR: ") + rt_last_error());
}
