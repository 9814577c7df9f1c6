// host/import.cpp — see import.h.
#include "import.h"

#include <zlib.h>

#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <map>
#include <sstream>
#include <stdexcept>

namespace rthost {

namespace {

std::string dir_of(const std::string& p) {
    const size_t s = p.find_last_of('/');
    return s == std::string::npos ? std::string() : p.substr(0, s + 1);
}

std::map<std::string, std::string> load_mtl(const std::string& path) {   // material name -> map_Kd
    std::map<std::string, std::string> out;
    std::ifstream f(path);
    std::string line, cur;
    while (std::getline(f, line)) {
        if (!line.empty() && line.back() == '\r') line.pop_back();
        size_t b = line.find_first_not_of(" \t");
        if (b == std::string::npos) continue;
        line = line.substr(b);
        if (line.compare(0, 7, "newmtl ") == 0) { cur = line.substr(7); out[cur] = ""; }
        else if (line.compare(0, 7, "map_Kd ") == 0 && !cur.empty()) out[cur] = line.substr(7);
    }
    return out;
}

int resolve_index(long i, size_t n) { return (int)(i > 0 ? i - 1 : (long)n + i); }

}  // namespace

std::vector<ObjMesh> load_obj(const std::string& path) {
    std::ifstream f(path);
    if (!f) throw std::runtime_error("ERROR: Assimp: Unable to open file \"" + path + "\".");
    std::vector<float> pos, tex;
    std::map<std::string, std::string> mtl;
    std::vector<ObjMesh> meshes;
    std::vector<bool> has_uv;
    std::string cur_mtl, line;
    auto start_mesh = [&]() {
        meshes.emplace_back();
        has_uv.push_back(true);
        meshes.back().diffuse_texture = mtl.count(cur_mtl) ? mtl[cur_mtl] : std::string();
    };
    while (std::getline(f, line)) {
        if (!line.empty() && line.back() == '\r') line.pop_back();
        const size_t hash = line.find('#');
        if (hash != std::string::npos) line.erase(hash);
        std::istringstream ss(line);
        std::string tag;
        if (!(ss >> tag)) continue;
        if (tag == "v") {
            std::string a, b, c; ss >> a >> b >> c;
            pos.push_back(std::strtof(a.c_str(), nullptr)); pos.push_back(std::strtof(b.c_str(), nullptr)); pos.push_back(std::strtof(c.c_str(), nullptr));
        } else if (tag == "vt") {
            std::string a, b; ss >> a >> b;
            tex.push_back(std::strtof(a.c_str(), nullptr)); tex.push_back(std::strtof(b.c_str(), nullptr));
        } else if (tag == "mtllib") {
            std::string name; std::getline(ss, name);
            const size_t b = name.find_first_not_of(" \t");
            if (b != std::string::npos) { auto m = load_mtl(dir_of(path) + name.substr(b)); mtl.insert(m.begin(), m.end()); }
        } else if (tag == "usemtl") {
            ss >> cur_mtl;
            if (meshes.empty() || !meshes.back().indices.empty()) start_mesh();
            else meshes.back().diffuse_texture = mtl.count(cur_mtl) ? mtl[cur_mtl] : std::string();
        } else if (tag == "o" || tag == "g") {
            if (meshes.empty() || !meshes.back().indices.empty()) start_mesh();
        } else if (tag == "f") {
            if (meshes.empty()) start_mesh();
            ObjMesh& m = meshes.back();
            std::vector<uint32_t> corners;
            std::string c;
            while (ss >> c) {
                long vi = 0, ti = 0;
                bool have_t = false;
                const size_t s1 = c.find('/');
                vi = std::strtol(c.substr(0, s1).c_str(), nullptr, 10);
                if (s1 != std::string::npos) {
                    const size_t s2 = c.find('/', s1 + 1);
                    const std::string t = c.substr(s1 + 1, s2 == std::string::npos ? std::string::npos : s2 - s1 - 1);
                    if (!t.empty()) { ti = std::strtol(t.c_str(), nullptr, 10); have_t = true; }
                }
                const int v = resolve_index(vi, pos.size() / 3);
                if (v < 0 || (size_t)v >= pos.size() / 3) throw std::runtime_error("ERROR: Assimp: OBJ: vertex index out of range");
                corners.push_back((uint32_t)(m.positions.size() / 3));
                m.positions.insert(m.positions.end(), pos.begin() + 3 * v, pos.begin() + 3 * v + 3);
                if (have_t) {
                    const int t = resolve_index(ti, tex.size() / 2);
                    if (t < 0 || (size_t)t >= tex.size() / 2) throw std::runtime_error("ERROR: Assimp: OBJ: texture index out of range");
                    m.uvs.push_back(tex[2 * t]);
                    m.uvs.push_back(1.0f - tex[2 * t + 1]);            // aiProcess_FlipUVs
                } else {
                    has_uv.back() = false;
                    m.uvs.push_back(0.0f); m.uvs.push_back(0.0f);
                }
            }
            for (size_t k = 1; k + 1 < corners.size(); ++k) {          // aiProcess_Triangulate, convex polygon fan
                m.indices.push_back(corners[0]); m.indices.push_back(corners[k]); m.indices.push_back(corners[k + 1]);
            }
        }
    }
    std::vector<ObjMesh> out;
    for (size_t i = 0; i < meshes.size(); ++i) {
        if (meshes[i].indices.empty()) continue;
        if (!has_uv[i]) meshes[i].uvs.clear();
        out.push_back(std::move(meshes[i]));
    }
    if (out.empty()) throw std::runtime_error("ERROR: Assimp: OBJ: no faces in \"" + path + "\"");
    return out;
}

// ------------------------------------------------------------------------------------------------ PNG
namespace {
uint32_t be32(const uint8_t* p) { return ((uint32_t)p[0] << 24) | ((uint32_t)p[1] << 16) | ((uint32_t)p[2] << 8) | p[3]; }
int paeth(int a, int b, int c) {
    const int p = a + b - c, pa = std::abs(p - a), pb = std::abs(p - b), pc = std::abs(p - c);
    return (pa <= pb && pa <= pc) ? a : (pb <= pc ? b : c);
}
}  // namespace

bool load_png(const std::string& path, Image8& out, std::string* why) {
    auto fail = [&](const char* m) { if (why) *why = m; return false; };
    std::ifstream f(path, std::ios::binary);
    if (!f) return fail("cannot open file");
    std::vector<uint8_t> buf((std::istreambuf_iterator<char>(f)), std::istreambuf_iterator<char>());
    static const uint8_t sig[8] = {0x89, 'P', 'N', 'G', 0x0D, 0x0A, 0x1A, 0x0A};
    if (buf.size() < 33 || std::memcmp(buf.data(), sig, 8) != 0) return fail("not a PNG");
    size_t p = 8;
    uint32_t w = 0, h = 0;
    int depth = 0, ctype = 0, interlace = 0;
    std::vector<uint8_t> idat;
    bool first_chunk = true;
    while (p + 12 <= buf.size()) {
        const uint32_t len = be32(&buf[p]);
        const char* type = (const char*)&buf[p + 4];
        if ((uint64_t)p + 12u + len > buf.size()) return fail("truncated chunk");
        const uint8_t* d = &buf[p + 8];
        if (first_chunk != !std::memcmp(type, "IHDR", 4)) return fail("IHDR must be the first chunk, once");
        first_chunk = false;
        if (!std::memcmp(type, "IHDR", 4)) {
            if (len != 13) return fail("bad IHDR length");
            w = be32(d); h = be32(d + 4); depth = d[8]; ctype = d[9]; interlace = d[12];
        }
        else if (!std::memcmp(type, "IDAT", 4)) idat.insert(idat.end(), d, d + len);
        else if (!std::memcmp(type, "IEND", 4)) break;
        p += 12 + len;
    }
    if (!w || !h) return fail("missing IHDR");
    if ((uint64_t)w * h > (1ull << 28)) return fail("image larger than 2^28 pixels");
    if (depth != 8 || interlace != 0) return fail("only 8-bit non-interlaced PNG is supported");
    int ch;
    switch (ctype) { case 0: ch = 1; break; case 2: ch = 3; break; case 4: ch = 2; break; case 6: ch = 4; break; default: return fail("palette PNG is not supported"); }
    const size_t stride = (size_t)w * ch;
    // deflate expands at most ~1032 : 1: a data stream too short for the declared size is refused before allocating
    if ((uint64_t)idat.size() * 1100u + 64u < (uint64_t)(stride + 1) * h) return fail("zlib inflate failed");
    std::vector<uint8_t> raw((stride + 1) * h);
    uLongf raw_len = raw.size();
    if (uncompress(raw.data(), &raw_len, idat.data(), idat.size()) != Z_OK || raw_len != raw.size()) return fail("zlib inflate failed");
    out.width = (int)w; out.height = (int)h; out.channels = ch;
    out.data.assign(stride * h, 0);
    for (uint32_t y = 0; y < h; ++y) {
        const uint8_t* src = &raw[(stride + 1) * y];
        const int ft = src[0];
        uint8_t* cur = &out.data[stride * y];
        const uint8_t* up = y ? &out.data[stride * (y - 1)] : nullptr;
        for (size_t x = 0; x < stride; ++x) {
            const int a = x >= (size_t)ch ? cur[x - ch] : 0, b = up ? up[x] : 0, c = (up && x >= (size_t)ch) ? up[x - ch] : 0;
            int v = src[1 + x];
            switch (ft) { case 0: break; case 1: v += a; break; case 2: v += b; break; case 3: v += (a + b) >> 1; break; case 4: v += paeth(a, b, c); break; default: return fail("bad filter type"); }
            cur[x] = (uint8_t)v;
        }
    }
    return true;
}

std::vector<float> ldr_to_hdr(const Image8& im) {
    float lut_c[256], lut_a[256];
    for (int i = 0; i < 256; ++i) { lut_a[i] = (float)i / 255.0f; lut_c[i] = powf(lut_a[i], 2.2f); }
    const int ch = im.channels, n_col = (ch & 1) ? ch : ch - 1;
    std::vector<float> out(im.data.size());
    for (size_t i = 0; i < im.data.size(); ++i) out[i] = ((int)(i % ch) < n_col ? lut_c : lut_a)[im.data[i]];
    return out;
}

}  // namespace rthost
