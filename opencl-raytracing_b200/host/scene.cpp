// host/scene.cpp — SceneCreator: the reference's scene builder and .scene file parser (src/scene.cpp) restated.
// The grammar is the one the reference's regular expressions define (src/scene.cpp:314,407,423,445,455),
// implemented here with a hand-written scanner; tokens are consumed strictly left to right (the reference relies
// on clang's left-to-right evaluation of function arguments, src/scene.cpp:351,355,380-392).
#include "../../include/scene.h"

#include <cmath>
#include <cstdlib>
#include <fstream>
#include <sstream>

#include "import.h"

namespace {

[[noreturn]] void scene_error(const std::string& m) { throw SceneError(m); }

bool is_space(char c) { return c == ' ' || c == '\t' || c == '\n' || c == '\r' || c == '\f' || c == '\v'; }
bool is_digit(char c) { return c >= '0' && c <= '9'; }

// Split on commas that are not inside a parenthesised group: a comma is a delimiter unless a ')' follows
// before any '(' (the reference's ",(?![^(]*\))").  No empty trailing token (std::sregex_token_iterator, -1).
std::vector<std::string> split_records(const std::string& s) {
    std::vector<std::string> out;
    size_t begin = 0;
    for (size_t i = 0; i < s.size(); ++i) {
        if (s[i] != ',') continue;
        bool inside = false;
        for (size_t j = i + 1; j < s.size(); ++j) {
            if (s[j] == '(') break;
            if (s[j] == ')') { inside = true; break; }
        }
        if (inside) continue;
        out.push_back(s.substr(begin, i - begin));
        begin = i + 1;
    }
    if (begin < s.size() || out.empty()) out.push_back(s.substr(begin));
    return out;
}

// num = [-+]? digits* ( '.' digits+ )?   whose first character is a digit, '.', '+' or '-'
size_t scan_number(const std::string& s, size_t p) {
    if (p >= s.size() || !(is_digit(s[p]) || s[p] == '.' || s[p] == '+' || s[p] == '-')) return std::string::npos;
    if (s[p] == '+' || s[p] == '-') ++p;
    while (p < s.size() && is_digit(s[p])) ++p;
    if (p + 1 < s.size() && s[p] == '.' && is_digit(s[p + 1])) {
        ++p;
        while (p < s.size() && is_digit(s[p])) ++p;
    }
    return p;
}
size_t skip_ws(const std::string& s, size_t p) { while (p < s.size() && is_space(s[p])) ++p; return p; }

float to_float(const std::string& s, const std::string& whole, const char* what) {
    const char* b = s.c_str();
    char* e = nullptr;
    const float v = std::strtof(b, &e);
    if (e == b) scene_error(std::string("ERROR: SCENE: IMPROPER ") + what + ": " + whole);   // std::stof would throw here
    return v;
}

struct Cursor {
    std::vector<std::string> tok;
    size_t i = 0;
    const std::string& next() {
        if (i >= tok.size()) scene_error("ERROR: SCENE: NOT ENOUGH PARAMETERS");
        return tok[i];
    }
};

// "\s*( num,\s*num,\s*num )\s*"
void get_vec(Cursor& c, float out[3]) {
    const std::string& w = c.next();
    size_t p = skip_ws(w, 0);
    bool ok = p < w.size() && w[p] == '(';
    size_t num_begin[3] = {0, 0, 0}, num_end[3] = {0, 0, 0};
    if (ok) ++p;
    for (int k = 0; ok && k < 3; ++k) {
        num_begin[k] = p;
        p = scan_number(w, p);
        if (p == std::string::npos) { ok = false; break; }
        num_end[k] = p;
        if (k < 2) { ok = p < w.size() && w[p] == ','; if (ok) p = skip_ws(w, p + 1); }
    }
    if (ok) { ok = p < w.size() && w[p] == ')'; if (ok) ok = skip_ws(w, p + 1) == w.size(); }
    if (!ok) scene_error("ERROR: SCENE: IMPROPER VECTOR: " + w);
    for (int k = 0; k < 3; ++k) out[k] = to_float(w.substr(num_begin[k], num_end[k] - num_begin[k]), w, "VECTOR");
    ++c.i;
}
cl_float3 get_vec3(Cursor& c) { float v[3]; get_vec(c, v); return cl_float3{v[0], v[1], v[2], 0.0f}; }

// "\s* num \s*"
float get_float(Cursor& c) {
    const std::string& w = c.next();
    size_t p = skip_ws(w, 0);
    const size_t e = scan_number(w, p);
    if (e == std::string::npos || skip_ws(w, e) != w.size()) scene_error("ERROR: SCENE: IMPROPER FLOAT: " + w);
    const float v = to_float(w, w, "FLOAT");
    ++c.i;
    return v;
}

// "\s*\d\s*" — exactly one digit (material ids 0-9), src/scene.cpp:455
cl_uint get_uint(Cursor& c) {
    const std::string& w = c.next();
    const size_t p = skip_ws(w, 0);
    if (p >= w.size() || !is_digit(w[p]) || skip_ws(w, p + 1) != w.size()) scene_error("ERROR: SCENE: IMPROPER UNSIGNED INT: " + w);
    ++c.i;
    return (cl_uint)(w[p] - '0');
}

// "\s*\".*?\"\s*"; the reference extracts substr(first_quote + 1, last_quote - 2), src/scene.cpp:407-414
std::string get_path(Cursor& c) {
    const std::string& w = c.next();
    const size_t p = skip_ws(w, 0);
    size_t e = w.size();
    while (e > p && is_space(w[e - 1])) --e;
    if (!(e >= p + 2 && w[p] == '"' && w[e - 1] == '"')) scene_error("ERROR: SCENE: IMPROPER PATH: " + w);
    const size_t first = w.find('"'), last = w.rfind('"');
    ++c.i;
    const std::string word = w;
    return word.substr(first + 1, last >= 2 ? last - 2 : 0);
}

bool file_exists(const std::string& p) { std::ifstream f(p); return (bool)f; }
std::string join_path(const std::string& root, const std::string& p) {
    if (root.empty() || (!p.empty() && p[0] == '/')) return p;
    return root.back() == '/' ? root + p : root + "/" + p;
}

}  // namespace

// src/scene.cpp:110-120
void SceneCreator::addMaterial(MatType type, const cl_float3& color, cl_float extra_data) { materials.push_back(Material(type, color, extra_data)); }
void SceneCreator::addSphere(const cl_float3& pos, cl_float r, cl_uint mat_ID) { spheres.push_back(Sphere(pos, r, mat_ID)); }
void SceneCreator::addPlane(const cl_float3& pos, const cl_float3& normal, cl_uint mat_ID) { planes.push_back(Plane(pos, normal, mat_ID)); }

// src/scene.cpp:122-143: p1 = pos + axis * sqrt(r1^2 - h^2), p2 = pos - axis * sqrt(r2^2 - h^2)
void SceneCreator::addLens(const cl_float3& pos, const cl_float3& normal, cl_float r1, cl_float r2, cl_float h, cl_uint mat_ID) {
    if (!(r1 >= h && r2 >= h)) scene_error("ERROR: LENS: r1 >= h && r2 >= h DOES NOT HOLD");   // assert() in the reference
    Lens lens;
    lens.pos = pos;
    const float t1 = std::sqrt(r1 * r1 - h * h), t2 = std::sqrt(r2 * r2 - h * h);
    lens.p1 = cl_float3{pos.x + normal.x * t1, pos.y + normal.y * t1, pos.z + normal.z * t1, 0.0f};
    lens.p2 = cl_float3{pos.x - normal.x * t2, pos.y - normal.y * t2, pos.z - normal.z * t2, 0.0f};
    lens.r1 = r1; lens.r2 = r2; lens.mat_ID = mat_ID;
    lenses.push_back(lens);
}

// src/scene.cpp:192-295 (loadModel, processNode, processMesh, transformVertex)
void SceneCreator::loadModel(const std::string& path, cl_uint mat_ID, const glm::mat4& t) {
    std::vector<rthost::ObjMesh> obj;
    try { obj = rthost::load_obj(join_path(asset_root, path)); }
    catch (const std::runtime_error& e) { scene_error(e.what()); }
    if (materials.size() <= mat_ID) scene_error("ERROR: MATERIAL OF ID: " + std::to_string(mat_ID) + " DOES NOT EXIST");
    for (const rthost::ObjMesh& m : obj) {
        const cl_uint index_anchor = (cl_uint)indices.size(), vertex_anchor = (cl_uint)vertices.size();
        const size_t nv = m.positions.size() / 3;
        for (size_t i = 0; i < nv; ++i) {
            const float x = m.positions[3 * i], y = m.positions[3 * i + 1], z = m.positions[3 * i + 2];
            cl_float3 v;                                                   // v' = M * (v, 1), column-major (:226-232)
            v.x = t[0][0] * x + t[1][0] * y + t[2][0] * z + t[3][0];
            v.y = t[0][1] * x + t[1][1] * y + t[2][1] * z + t[3][1];
            v.z = t[0][2] * x + t[1][2] * y + t[2][2] * z + t[3][2];
            v.w = 0.0f;
            vertices.push_back(v);
            if (!m.uvs.empty()) texture_uv.push_back(cl_float2{m.uvs[2 * i], m.uvs[2 * i + 1]});   // only if present (:246-252)
        }
        indices.insert(indices.end(), m.indices.begin(), m.indices.end());
        cl_uint texture_ID = (cl_uint)-1;
        if (materials[mat_ID].type == t_textured) {                        // :264-292
            if (m.diffuse_texture.empty()) scene_error("ERROR: MESH HAS NO TEXTURE APPLIED, USE A DIFFERENT MATERIAL");
            size_t j = 0;
            while (j < texture_paths.size() && texture_paths[j] != m.diffuse_texture) ++j;
            if (j == texture_paths.size()) texture_paths.push_back(m.diffuse_texture);
            texture_ID = (cl_uint)j;
        }
        meshes.push_back(Mesh(vertex_anchor, index_anchor, (cl_uint)(m.indices.size() / 3), texture_ID));
    }
    models.push_back(Model(mesh_count_total, (cl_uint)obj.size(), mat_ID));
    mesh_count_total += (cl_uint)obj.size();
}

// The .mtl files of the reference carry absolute paths of the author's machine (assets/cube/cube.mtl:13):
// use the path as given if it exists, else its basename under the texture directory.
std::string SceneCreator::resolveTexture(const std::string& mtl_path) const {
    if (file_exists(mtl_path)) return mtl_path;
    if (file_exists(join_path(asset_root, mtl_path))) return join_path(asset_root, mtl_path);
    const size_t s = mtl_path.find_last_of("/\\");
    const std::string base = s == std::string::npos ? mtl_path : mtl_path.substr(s + 1);
    const std::string dir = texture_dir.empty() ? join_path(asset_root, "assets/textures") : texture_dir;
    return join_path(dir, base);
}

// src/scene.cpp:145-190
void SceneCreator::loadTextures() {
    texture_data.clear();
    tex_layers = 0; tex_width = tex_height = 1;
    if (models.empty()) return;                                            // 1x1x1 dummy array (:188)
    if (texture_paths.empty()) scene_error("ERROR: TEXTURE COUNT = 0");
    for (size_t id = 0; id < texture_paths.size(); ++id) {
        rthost::Image8 im;
        std::string why;
        bool ok = false;
        try { ok = rthost::load_png(resolveTexture(texture_paths[id]), im, &why); }
        catch (const std::exception&) { ok = false; }                      // e.g. bad_alloc on a damaged header: the reference's error path, not terminate()
        if (!ok) scene_error("ERROR: STBimage: COULD NOT FIND THE TEXTURE");
        if (id == 0) { tex_width = im.width; tex_height = im.height; }
        else if (im.width != tex_width || im.height != tex_height)
            scene_error("ERROR: TEXTURES HAVE DIFFERENT SIZES: TEMPLATE: " + std::to_string(tex_width) + " x " + std::to_string(tex_height) +
                        ", TEXTURE ID(" + std::to_string(id) + "): " + std::to_string(im.width) + " x " + std::to_string(im.height));
        if (im.channels != 4) scene_error("ERROR: STBimage: TEXTURE HAS A WRONG FORMAT: " + std::to_string(im.channels) + " INSTEAD OF 4 (RGBA)");
        const std::vector<float> f = rthost::ldr_to_hdr(im);
        texture_data.insert(texture_data.end(), f.begin(), f.end());
        ++tex_layers;
    }
}

// src/scene.cpp:297-403
void SceneCreator::loadScene(const std::string& path) {
    std::ifstream file(join_path(asset_root, path));
    if (!file) scene_error("ERROR: SCENE: NOT SUCCESFULLY READ: " + path);
    std::stringstream data;
    data << file.rdbuf();
    enum { l_none, l_materials, l_spheres, l_planes, l_lenses, l_models } mode = l_none;
    glm::mat4 model(1.0f);
    std::string line;
    while (std::getline(data, line)) {
        const size_t hash = line.find('#');
        if (hash != std::string::npos) line.erase(hash);
        if (line.empty()) continue;                                        // a blank-only line is NOT empty (:323)
        const size_t colon = line.find(':');
        if (colon != std::string::npos) {
            const std::string word = line.substr(0, colon);
            if (word == "MATERIALS") { mode = l_materials; continue; }
            if (word == "SPHERES") { mode = l_spheres; continue; }
            if (word == "PLANES") { mode = l_planes; continue; }
            if (word == "LENSES") { mode = l_lenses; continue; }
            if (word == "MODELS") { mode = l_models; continue; }
            if (mode != l_models) scene_error("ERROR: SCENE: OPERATION " + word + " DOES NOT EXIST");
            Cursor c;
            c.tok = split_records(line.substr(colon + 1));
            if (word == "translate") { float v[3]; get_vec(c, v); model = glm::translate(model, glm::vec3(v[0], v[1], v[2])); }
            else if (word == "rotate") {
                const float deg = get_float(c);
                float v[3]; get_vec(c, v);
                model = glm::rotate(model, glm::radians(deg), glm::vec3(v[0], v[1], v[2]));
            }
            else if (word == "scale") { float v[3]; get_vec(c, v); model = glm::scale(model, glm::vec3(v[0], v[1], v[2])); }
            else if (word == "load") {
                const std::string p = get_path(c);
                const cl_uint id = get_uint(c);
                loadModel(p, id, model);
                model = glm::mat4(1.0f);
            }                                                              // other words are silently ignored (:348-357)
            continue;
        }
        Cursor c;
        c.tok = split_records(line);
        switch (mode) {
            case l_materials: {
                const std::string word = c.next();
                MatType type = t_light;
                if (word == "reflective") type = t_reflective;
                else if (word == "refractive") type = t_refractive;
                else if (word == "diffuse") type = t_diffuse;
                else if (word == "dielectric") type = t_dielectric;
                else if (word == "light") type = t_light;
                else if (word == "textured") type = t_textured;
                else scene_error("ERROR: SCENE: MATERIAL: " + word + " DOES NOT EXIST");
                ++c.i;
                const cl_float3 col = get_vec3(c);
                const float extra = get_float(c);
                addMaterial(type, col, extra);
                break;
            }
            case l_spheres: {
                const cl_float3 pos = get_vec3(c);
                const float r = get_float(c);
                const cl_uint id = get_uint(c);
                addSphere(pos, r, id);
                break;
            }
            case l_planes: {
                const cl_float3 pos = get_vec3(c);
                const cl_float3 n = get_vec3(c);
                const cl_uint id = get_uint(c);
                addPlane(pos, n, id);
                break;
            }
            case l_lenses: {
                const cl_float3 pos = get_vec3(c);
                const cl_float3 n = get_vec3(c);
                const float r1 = get_float(c), r2 = get_float(c), h = get_float(c);
                const cl_uint id = get_uint(c);
                addLens(pos, n, r1, r2, h, id);
                break;
            }
            default:
                scene_error("ERROR: SCENE: OPERATION NOT SPECIFIED");
        }
    }
}

rt_scene_desc SceneCreator::desc() const {
    rt_scene_desc d{};
    d.materials = materials.data(); d.n_materials = (uint32_t)materials.size();
    d.spheres = spheres.data(); d.n_spheres = (uint32_t)spheres.size();
    d.planes = planes.data(); d.n_planes = (uint32_t)planes.size();
    d.lenses = lenses.data(); d.n_lenses = (uint32_t)lenses.size();
    d.vertices = vertices.data(); d.n_vertices = (uint32_t)vertices.size();
    d.uvs = texture_uv.data(); d.n_uvs = (uint32_t)texture_uv.size();
    d.indices = indices.data(); d.n_indices = (uint32_t)indices.size();
    d.meshes = meshes.data(); d.n_meshes = (uint32_t)meshes.size();
    d.models = models.data(); d.n_models = (uint32_t)models.size();
    return d;
}

void SceneCreator::createScene(rt_ctx* ctx) {
    if (tex_layers > 0 && rt_upload_textures(ctx, texture_data.data(), tex_width, tex_height, tex_layers) != RT_OK)
        scene_error(std::string("ERROR: TEXTURE UPLOAD: ") + rt_last_error());
    const rt_scene_desc d = desc();
    if (rt_upload_scene(ctx, &d) != RT_OK) scene_error(std::string("ERROR: SCENE UPLOAD: ") + rt_last_error());
}
