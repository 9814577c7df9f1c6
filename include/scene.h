// include/scene.h — SceneCreator of the reference (include/scene.h:30-153, src/scene.cpp) with the same host data
// model and the same public building interface.  The OpenCL plumbing (cl::Buffer members, createKernel,
// setupBuffers, setKernelArgs, createScene(cl::Context&, cl::Device&)) is replaced by createScene(rt_ctx*), which
// hands the host arrays to the C ABI (rt_upload_textures + rt_upload_scene).  Assimp and stb_image are replaced
// by a small OBJ/MTL reader and a zlib-based PNG reader that reproduce what the reference consumes from them.
#ifndef scene_h
#define scene_h

#include <stdexcept>
#include <string>
#include <vector>

#include "rt_capi.h"
#include "rt_host_types.h"

enum MatType { t_refractive, t_reflective, t_dielectric, t_diffuse, t_textured, t_light };   // include/scene.h:30

struct Material {                 // 48 bytes, include/scene.h:32-39
    MatType type;
    cl_float3 color;
    cl_float extra_data;          // reflect attenuation / refractive index / diffuse albedo scale
    Material(MatType type, const cl_float3& color, cl_float extra_data) : type(type), color(color), extra_data(extra_data) {}
    Material() {}
};
struct Sphere {                   // 32 bytes
    cl_float3 pos;
    cl_float r;
    cl_uint mat_ID;
    Sphere(const cl_float3& pos, cl_float r, cl_uint mat_ID) : pos(pos), r(r), mat_ID(mat_ID) {}
};
struct Plane {                    // 48 bytes
    cl_float3 pos;
    cl_float3 normal;
    cl_uint mat_ID;
    Plane(const cl_float3& pos, const cl_float3& normal, cl_uint mat_ID) : pos(pos), normal(normal), mat_ID(mat_ID) {}
};
struct Lens {                     // 64 bytes
    cl_float3 pos, p1, p2;        // centre and the centres of the two spherical caps
    cl_float r1, r2;
    cl_uint mat_ID;
};
struct Mesh {                     // 16 bytes
    cl_uint vertex_anchor, index_anchor, face_count, texture_ID;
    Mesh(cl_uint va, cl_uint ia, cl_uint fc, cl_uint tid) : vertex_anchor(va), index_anchor(ia), face_count(fc), texture_ID(tid) {}
};
struct Model {                    // 12 bytes
    cl_uint mesh_anchor, mesh_count, mat_ID;
    Model(cl_uint ma, cl_uint mc, cl_uint mid) : mesh_anchor(ma), mesh_count(mc), mat_ID(mid) {}
};
static_assert(sizeof(Material) == 48 && sizeof(Sphere) == 32 && sizeof(Plane) == 48 && sizeof(Lens) == 64 &&
              sizeof(Mesh) == 16 && sizeof(Model) == 12, "layouts of include/scene.h:32-81");

// The reference prints the message and calls exit(-1) (src/scene.cpp:29-32).  Library code throws this instead;
// RayTracer and the headless driver keep the print-and-exit behaviour at their boundary.
struct SceneError : std::runtime_error {
    explicit SceneError(const std::string& m) : std::runtime_error(m) {}
};

class SceneCreator {
private:
    std::vector<Material> materials;
    std::vector<Sphere> spheres;
    std::vector<Plane> planes;
    std::vector<Lens> lenses;
    std::vector<Model> models;
    std::vector<cl_float3> vertices;
    std::vector<cl_float2> texture_uv;
    std::vector<cl_uint> indices;
    std::vector<Mesh> meshes;
    std::vector<std::string> texture_paths;

    std::vector<float> texture_data;          // RGBA f32, [layer][row][col]
    int tex_width = 1, tex_height = 1, tex_layers = 0;
    cl_uint mesh_count_total = 0;             // function-static in the reference (src/scene.cpp:193)
    std::string asset_root;                   // the reference resolves paths against the working directory
    std::string texture_dir;                  // where .mtl texture basenames are looked up (see resolveTexture)

    std::string resolveTexture(const std::string& mtl_path) const;

public:
    void setAssetRoot(const std::string& root) { asset_root = root; }
    void setTextureDir(const std::string& dir) { texture_dir = dir; }

    void addMaterial(MatType type, const cl_float3& color, cl_float extra_data);
    void addSphere(const cl_float3& pos, cl_float r, cl_uint mat_ID);
    void addPlane(const cl_float3& pos, const cl_float3& normal, cl_uint mat_ID);
    void addLens(const cl_float3& pos, const cl_float3& normal, cl_float r1, cl_float r2, cl_float h, cl_uint mat_ID);
    void loadModel(const std::string& path, cl_uint mat_ID, const glm::mat4& transform = glm::mat4(1.0f));

    void loadTextures();                      // src/scene.cpp:145-190 (host half)
    void loadScene(const std::string& path);  // src/scene.cpp:297-403

    // SceneCreator::setupBuffers + createScene (src/scene.cpp:46-87): upload through the C ABI.
    void createScene(rt_ctx* ctx);

    rt_scene_desc desc() const;
    const std::vector<float>& textures(int* w, int* h, int* layers) const { *w = tex_width; *h = tex_height; *layers = tex_layers; return texture_data; }
    const std::vector<std::string>& texturePaths() const { return texture_paths; }
};

#endif /* scene_h */
