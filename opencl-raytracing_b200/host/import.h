// host/import.h — minimal importers standing in for Assimp (OBJ/MTL) and stb_image (PNG) at the two call sites of the
// reference: src/scene.cpp:195 (ReadFile(path, aiProcess_Triangulate | aiProcess_FlipUVs)) and :158 (stbi_loadf).
#pragma once
#include <cstdint>
#include <string>
#include <vector>

namespace rthost {

struct ObjMesh {                       // what SceneCreator::processMesh consumes from one aiMesh
    std::vector<float> positions;      // 3 per vertex, one vertex per face corner (Assimp's OBJ importer does not join)
    std::vector<float> uvs;            // 2 per vertex (v already flipped), empty if any corner lacks a vt
    std::vector<uint32_t> indices;     // 3 per face after fan triangulation (0,1,2),(0,2,3),...
    std::string diffuse_texture;       // map_Kd of the mesh's material ("" if none)
};

// Throws std::runtime_error("ERROR: Assimp: ...") on failure, like the reference's message at src/scene.cpp:198.
std::vector<ObjMesh> load_obj(const std::string& path);

struct Image8 { int width = 0, height = 0, channels = 0; std::vector<uint8_t> data; };
// 8-bit, non-interlaced PNG of colour type 0/2/4/6.  Returns false if the file cannot be read or decoded.
bool load_png(const std::string& path, Image8& out, std::string* why);

// stbi_loadf's LDR -> float conversion: powf(c / 255.0f, 2.2f) for colour channels, c / 255.0f for alpha
// (alpha = the last channel when the channel count is even).  Row 0 = top row, no flip.
std::vector<float> ldr_to_hdr(const Image8& im);

}  // namespace rthost
