// host/rt_headless.cpp — headless stand-in for the reference's application loop (main.cpp:62-117): the same
// objects (Camera, RayTracer), the same per-frame decision (camera moved -> render(), else renderAgain(),
// main.cpp:85-96), camera controls scripted from a file instead of GLFW callbacks, and the image written to a TGA
// (layout of main.cpp:262-289) and/or raw float32 instead of being presented through OpenGL.
//
//   rt_headless [--scene F] [--root DIR] [--textures DIR] [--width W] [--height H] [--spp N] [--depth D]
//               [--arith strict|fast] [--rng table|counter] [--table-seed S] [--fov DEG] [--pos x,y,z]
//               [--yaw DEG] [--pitch DEG] [--script FILE] [--out img.tga] [--raw img.f32]
//               [--checkpoint FILE] [--resume FILE]
//   checkpoint file: "RTC2" u32 width, height, samples, u64 fingerprint, then width*height*4 float32 of the (sqrt(mean),1)
//   image; --resume restores it (rt_restore_image) and continues with renderAgain, bit-identical to an uninterrupted run.
//   The fingerprint (FNV-1a) covers everything the continuation depends on — the 12 camera floats, max depth, arithmetic and
//   RNG mode, the scene file name and RT_TABLE_SEED — and a resume with a different fingerprint is refused.  The random table
//   is seeded from std::random_device unless RT_TABLE_SEED pins it (src/raytracer.cpp:73-74 does the same): without a pinned
//   seed a checkpoint cannot be continued bit-identically and --checkpoint / --resume say so.
//   script lines:  move forward|back|left|right DT | rotate DX DY | zoom SCROLL | fast 0|1 | slow 0|1 | frames N
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <sstream>
#include <string>
#include <vector>

#include "../../include/raytracer.h"

int main(int argc, char** argv) {
    int W = 1200, H = 800, spp = 16, fov = 60;   // main.cpp:10-11,65 (SCR_WIDTH x SCR_HEIGHT, fov 60)
    float pos[3] = {0, 0, 0}, yaw = 0, pitch = 0;
    std::string out_tga, out_raw, script, ckpt_out, ckpt_in;
    for (int i = 1; i < argc; ++i) {
        const std::string a = argv[i];
        auto val = [&]() -> const char* { if (i + 1 >= argc) { std::fprintf(stderr, "missing value for %s\n", a.c_str()); std::exit(2); } return argv[++i]; };
        if (a == "--scene") setenv("RT_SCENE_FILE", val(), 1);
        else if (a == "--root") setenv("RT_ASSET_ROOT", val(), 1);
        else if (a == "--textures") setenv("RT_TEXTURE_DIR", val(), 1);
        else if (a == "--depth") setenv("RT_MAX_DEPTH", val(), 1);
        else if (a == "--arith") setenv("RT_ARITH", val(), 1);
        else if (a == "--rng") setenv("RT_RNG", val(), 1);
        else if (a == "--table-seed") setenv("RT_TABLE_SEED", val(), 1);
        else if (a == "--width") W = std::atoi(val());
        else if (a == "--height") H = std::atoi(val());
        else if (a == "--spp") spp = std::atoi(val());
        else if (a == "--fov") fov = std::atoi(val());
        else if (a == "--yaw") yaw = (float)std::atof(val());
        else if (a == "--pitch") pitch = (float)std::atof(val());
        else if (a == "--pos") std::sscanf(val(), "%f,%f,%f", &pos[0], &pos[1], &pos[2]);
        else if (a == "--script") script = val();
        else if (a == "--out") out_tga = val();
        else if (a == "--raw") out_raw = val();
        else if (a == "--checkpoint") ckpt_out = val();
        else if (a == "--resume") ckpt_in = val();
        else { std::fprintf(stderr, "unknown option %s\n", a.c_str()); return 2; }
    }
    Camera camera(fov, (float)W / (float)H, glm::vec3(pos[0], pos[1], pos[2]), yaw, pitch);
    RayTracer tracer(W, H, "kernels/raytracer.cl");
    auto fingerprint = [&]() {
        uint64_t hsh = 1469598103934665603ull;
        auto mix = [&](const void* p, size_t n) { for (size_t k = 0; k < n; ++k) { hsh ^= ((const unsigned char*)p)[k]; hsh *= 1099511628211ull; } };
        mix(camera.transferData(), 48);
        rt_params prm;
        rt_get_params(tracer.context(), &prm);
        mix(&prm.max_depth, 4); mix(&prm.arith, 4); mix(&prm.rng_mode, 4); mix(&prm.mesh_mode, 4);
        for (const char* name : {"RT_SCENE_FILE", "RT_TABLE_SEED"}) { const char* v = std::getenv(name); if (v) mix(v, std::strlen(v)); mix("|", 1); }
        return hsh;
    };
    if ((!ckpt_in.empty() || !ckpt_out.empty()) && !std::getenv("RT_TABLE_SEED"))
        std::fprintf(stderr, "warning: no --table-seed: the random table of this run cannot be reproduced, a resumed run will not be bit-identical\n");

    bool camera_in_motion = true;                       // main.cpp:48: the first frame is a render()
    if (!ckpt_in.empty()) {
        std::ifstream f(ckpt_in, std::ios::binary);
        char magic[4] = {0};
        uint32_t hdr[3] = {0, 0, 0};
        uint64_t fp = 0;
        f.read(magic, 4); f.read((char*)hdr, sizeof hdr); f.read((char*)&fp, sizeof fp);
        if (!f || std::memcmp(magic, "RTC2", 4) != 0 || hdr[0] != (uint32_t)W || hdr[1] != (uint32_t)H || hdr[2] == 0) {
            std::fprintf(stderr, "cannot resume from %s (missing, damaged, or a different frame size)\n", ckpt_in.c_str());
            return 2;
        }
        if (fp != fingerprint()) {
            std::fprintf(stderr, "cannot resume from %s: it was written with a different camera, depth, arithmetic / RNG mode, scene or table seed\n", ckpt_in.c_str());
            return 2;
        }
        std::vector<float> img((size_t)W * H * 4);
        f.read((char*)img.data(), (std::streamsize)(img.size() * 4));
        if (!f || rt_restore_image(tracer.context(), img.data(), hdr[2]) != RT_OK) { std::fprintf(stderr, "resume failed: %s\n", rt_last_error()); return 2; }
        camera_in_motion = false;                       // the restored image belongs to this camera: continue with renderAgain
    }
    long frames = 0;
    const auto t0 = std::chrono::steady_clock::now();
    auto frame = [&]() {                                // main.cpp:91-95
        if (camera_in_motion) { tracer.render(&camera); camera_in_motion = false; }
        else tracer.renderAgain(&camera);
        ++frames;
    };
    if (script.empty()) {
        for (int k = 0; k < spp; ++k) frame();
    } else {
        std::ifstream f(script);
        if (!f) { std::fprintf(stderr, "cannot read script %s\n", script.c_str()); return 2; }
        std::string line;
        while (std::getline(f, line)) {
            std::istringstream ss(line);
            std::string cmd;
            if (!(ss >> cmd) || cmd[0] == '#') continue;
            if (cmd == "move") {
                std::string dir; float dt = 0; ss >> dir >> dt;
                camera.move(dir == "forward" ? FORWARD : dir == "back" ? BACK : dir == "left" ? LEFT : RIGHT, dt);
                camera_in_motion = true;
            } else if (cmd == "rotate") { float dx = 0, dy = 0; ss >> dx >> dy; camera.rotate(dx, dy); camera_in_motion = true; }
            else if (cmd == "zoom") { float s = 0; ss >> s; camera.zoom(s); camera_in_motion = true; }
            else if (cmd == "fast") { int b = 0; ss >> b; camera.setFasterSpeed(b != 0); }
            else if (cmd == "slow") { int b = 0; ss >> b; camera.setSlowerSpeed(b != 0); }
            else if (cmd == "frames") { int n = 0; ss >> n; for (int k = 0; k < n; ++k) frame(); }
            else { std::fprintf(stderr, "unknown script command %s\n", cmd.c_str()); return 2; }
        }
    }
    const double secs = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    rt_stats st;
    rt_get_stats(tracer.context(), &st);
    std::vector<float> img((size_t)W * H * 4);
    tracer.readImage(img.data());
    if (!out_tga.empty() && rt_write_tga(out_tga.c_str(), img.data(), W, H) != RT_OK) { std::fprintf(stderr, "%s\n", rt_last_error()); return 1; }
    if (!ckpt_out.empty()) {
        std::ofstream o(ckpt_out, std::ios::binary);
        const uint32_t hdr[3] = {(uint32_t)W, (uint32_t)H, st.samples_done};
        const uint64_t fp = fingerprint();
        o.write("RTC2", 4); o.write((const char*)hdr, sizeof hdr); o.write((const char*)&fp, sizeof fp); o.write((const char*)img.data(), (std::streamsize)(img.size() * 4));
    }
    if (!out_raw.empty()) { std::ofstream o(out_raw, std::ios::binary); o.write((const char*)img.data(), (std::streamsize)(img.size() * 4)); }
    std::printf("{\"frames\": %ld, \"samples_in_image\": %u, \"width\": %d, \"height\": %d, \"rays\": %llu, \"seconds\": %.4f, \"mrays_per_s\": %.1f}\n",
                frames, st.samples_done, W, H, (unsigned long long)st.rays_total, secs, st.rays_total / secs / 1e6);
    return 0;
}
