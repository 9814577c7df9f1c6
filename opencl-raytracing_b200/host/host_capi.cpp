// host/host_capi.cpp — C entry points over the C++ host classes (declared in include/rt_capi.h, "host-only helpers").
#include <cstdio>
#include <cstring>
#include <string>
#include <vector>

#include "../../include/camera.h"
#include "../../include/scene.h"

extern "C" int rt_set_error_message(int code, const char* msg);   // csrc/rt_capi.cpp

struct rt_scene { SceneCreator sc; };

extern "C" {

int rt_scene_load(const char* scene_path, const char* asset_root, const char* texture_dir, rt_scene** out) {
    if (!scene_path || !out) return rt_set_error_message(RT_ERR_INVALID, "rt_scene_load: null argument");
    *out = nullptr;
    rt_scene* s = new rt_scene();
    try {
        s->sc.setAssetRoot(asset_root ? asset_root : "");
        s->sc.setTextureDir(texture_dir ? texture_dir : "");
        s->sc.loadScene(scene_path);
        s->sc.loadTextures();
    } catch (const std::exception& e) {
        delete s;
        return rt_set_error_message(RT_ERR_INVALID, e.what());
    }
    *out = s;
    return RT_OK;
}
int rt_scene_get_desc(const rt_scene* s, rt_scene_desc* out) {
    if (!s || !out) return rt_set_error_message(RT_ERR_INVALID, "rt_scene_get_desc: null argument");
    *out = s->sc.desc();
    return RT_OK;
}
int rt_scene_get_textures(const rt_scene* s, const float** rgba, int* w, int* h, int* layers) {
    if (!s || !rgba || !w || !h || !layers) return rt_set_error_message(RT_ERR_INVALID, "rt_scene_get_textures: null argument");
    const std::vector<float>& t = s->sc.textures(w, h, layers);
    *rgba = t.empty() ? nullptr : t.data();
    return RT_OK;
}
int rt_scene_upload(const rt_scene* s, rt_ctx* ctx) {
    if (!s || !ctx) return rt_set_error_message(RT_ERR_INVALID, "rt_scene_upload: null argument");
    try { const_cast<rt_scene*>(s)->sc.createScene(ctx); }
    catch (const std::exception& e) { return rt_set_error_message(RT_ERR_INVALID, e.what()); }
    return RT_OK;
}
int rt_scene_free(rt_scene* s) { delete s; return RT_OK; }

int rt_camera_data(int fov, float aspect, const float pos[3], float yaw, float pitch, float out[12]) {
    if (!out) return rt_set_error_message(RT_ERR_INVALID, "rt_camera_data: null output");
    const Camera cam(fov, aspect, pos ? glm::vec3(pos[0], pos[1], pos[2]) : glm::vec3(0.0f), yaw, pitch);
    std::memcpy(out, cam.transferData(), 12 * sizeof(float));
    return RT_OK;
}

int rt_camera_run(int fov, float aspect, const float pos[3], float yaw, float pitch, const float* ops, int n_ops, float out[12]) {
    if (!out || n_ops < 0 || (n_ops > 0 && !ops)) return rt_set_error_message(RT_ERR_INVALID, "rt_camera_run: bad argument");
    Camera cam(fov, aspect, pos ? glm::vec3(pos[0], pos[1], pos[2]) : glm::vec3(0.0f), yaw, pitch);
    for (int i = 0; i < n_ops; ++i) {
        const int code = (int)ops[3 * i];
        const float a = ops[3 * i + 1], b = ops[3 * i + 2];
        switch (code) {
            case 0: cam.move(FORWARD, a); break;
            case 1: cam.move(BACK, a); break;
            case 2: cam.move(LEFT, a); break;
            case 3: cam.move(RIGHT, a); break;
            case 4: cam.rotate(a, b); break;
            case 5: cam.zoom(a); break;
            case 6: cam.setFasterSpeed(a != 0.0f); break;
            case 7: cam.setSlowerSpeed(a != 0.0f); break;
            case 8: cam.setSize(a); break;
            default: return rt_set_error_message(RT_ERR_INVALID, "rt_camera_run: unknown control code");
        }
    }
    std::memcpy(out, cam.transferData(), 12 * sizeof(float));
    return RT_OK;
}

int rt_write_tga(const char* path, const float* rgba, int w, int h) {
    if (!path || !rgba || w <= 0 || h <= 0 || w > 65535 || h > 65535) return rt_set_error_message(RT_ERR_INVALID, "rt_write_tga: bad argument");
    FILE* f = std::fopen(path, "wb");
    if (!f) return rt_set_error_message(RT_ERR_INVALID, "rt_write_tga: cannot open output file");
    const unsigned short hdr[9] = {0, 2, 0, 0, 0, 0, (unsigned short)w, (unsigned short)h, 24};   // main.cpp:280
    std::fwrite(hdr, sizeof hdr, 1, f);
    std::vector<unsigned char> row((size_t)w * 3);
    for (int y = 0; y < h; ++y) {                      // image row 0 is the bottom row, as TGA expects
        for (int x = 0; x < w; ++x) {
            const float* p = rgba + 4 * ((size_t)y * w + x);
            for (int c = 0; c < 3; ++c) {
                float v = p[2 - c];                    // BGR
                v = !(v >= 0.0f) ? 0.0f : (v > 1.0f ? 1.0f : v);   // NaN -> 0 (a float -> unsigned char cast of NaN is undefined)
                row[3 * x + c] = (unsigned char)(v * 255.0f + 0.5f);
            }
        }
        std::fwrite(row.data(), 1, row.size(), f);
    }
    std::fclose(f);
    return RT_OK;
}

}  // extern "C"
