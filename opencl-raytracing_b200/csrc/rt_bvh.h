// csrc/rt_bvh.h — GPU-built LBVH over the finite primitives of a scene (spheres, lenses, triangles).
// The reference is brute force only (kernels/raytracer.cl:322-360, O(N) per ray); north_star asks for a GPU-built
// LBVH "that must return the same nearest hit as brute force".  Planes are unbounded and stay in the brute list.
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

namespace rtb {

// Traversal node: both children's boxes live in the parent, so one 64-byte fetch decides two subtrees.
// child >= 0: internal node index; child < 0: leaf, slot = ~child.
struct alignas(16) BvhNode {
    float lo0[3]; int32_t child0;
    float hi0[3]; int32_t child1;
    float lo1[3]; uint32_t flags;     // bit 0 / 1: the subtree of child 0 / 1 holds a sphere or a lens (needs the inflated box test)
    float hi1[3]; uint32_t rho;       // two binary16, child 0 low / child 1 high: an upper bound of the radius of any sphere inside the
                                      // child's box (0.5005 x its smallest extent + 1e-6, rounded up) — the thin-flare bound of slab<true>
};
static_assert(sizeof(BvhNode) == 64, "one node = four 16-byte words");

// One push per level of the binary tree; a Karras tree over 63-bit codes + 32-bit positions is at most 95 levels deep.
constexpr int kBvhStack = 96;

struct DevBvh {
    const BvhNode* nodes;     // n_leaves - 1 internal nodes, root = 0 (n_leaves >= 2)
    const float4* leaves;     // per leaf slot (Morton order) 4 x float4: the primitive's test data as in the stream (sphere: 1,
                              // lens: 2, triangle: 3 words), then (hit id = kind << 28 | reference index, stream index, 0, 0)
    uint32_t n_leaves;        // 0 = no BVH (brute-force path)
    int32_t root;             // 0, or ~0 when the tree is a single leaf
};

struct BvhBuildInput {        // device pointers into the scene buffers
    const float4* stream;     // test stream (see rt_device.h)
    uint32_t off_len, off_tri;
    uint32_t n_sph, n_len, n_tri;
    const uint32_t* tri_face_id;   // reference-global face number per stream triangle
};

struct BvhBuffers {           // owned by the context; (re)allocated by build_lbvh
    BvhNode* nodes = nullptr;
    float4* leaves = nullptr;
    uint32_t n_leaves = 0;
};

// Morton codes -> radix sort (CUB) -> Karras hierarchy -> bottom-up refit -> traversal nodes; all on `stream`.
cudaError_t build_lbvh(const BvhBuildInput& in, BvhBuffers& out, cudaStream_t stream);
void free_lbvh(BvhBuffers& b);

}  // namespace rtb
