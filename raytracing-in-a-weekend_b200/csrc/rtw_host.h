// rtw_host.h — host-side flattened scene: what rtw_scene_create builds from an rtw_scene_desc
// before uploading. Replaces Scene::new / AABB::new / QuadAABB::new / IAABB::new
// (Rust/src/viewport.rs:120-135, objects/aabb/aabb.rs:20-71, qaabb.rs:127-163, iaabb.rs:80-116):
// one two-level SAH BVH instead of three random-axis median-split trees plus two per instance.
#pragma once
#include <string>
#include <vector>
#include "rtw_device.h"
#include "rtw_lbvh.h"
#include "../../include/rtw.h"

namespace rtw {

struct HostScene {
    std::vector<Node> nodes;
    std::vector<Node16> nodes16;
    std::vector<float4> spheres, sphere_vel, sphere_col, sphere_box;
    std::vector<uint2> sphere_ids;
    std::vector<Quad> quads;
    std::vector<uint2> quad_ids;
    std::vector<Instance> instances;
    std::vector<float4> one_sphere;
    std::vector<uint32_t> moving;
    std::vector<Material> materials;
    std::vector<Texture> textures;                 // obj / texels pointers are filled by the uploader
    std::vector<std::vector<float4>> texels;       // per texture, RGBA f32 (A = 0)
    std::vector<Noise> noises;
    Level top;
    // Enclosing volumes: a constant-density instance made of one sphere only scatters rays
    // that enter it from outside (instance.rs:280-300: a ray that starts inside meets the exit face,
    // probes from there outwards and finds nothing => None). When such a sphere contains every other
    // primitive of the scene, and the camera, no ray of a render ever starts outside it, so it can
    // never return a hit: `top_inside` is the top level built without those instances. The fast
    // kernels walk it when the camera test passes (rtw_api.cu); the strict ones walk `top`.
    struct Enclosing { float c[3]; float r; float mov_far, mov_speed; };
    std::vector<Enclosing> enclosing;
    Level top_inside;            // valid when !enclosing.empty()
    // Baked instances. The reference walks every instance as a level of its own: ray -> local frame, two inner BVHs,
    // hit -> world (instance.rs:250-310); in the kernels that is an instance-entry and a level-exit event per
    // instance the ray's box test passes - the dearest, most divergent steps of the extend kernel (14 % of its warp
    // instructions at 5 of 32 lanes on C5, profiles/r1l_wf_extend_by_line.txt). Where the instance's frame change
    // is a rigid motion (the Euler map of vec3.rs:161-181 is orthogonal, e.g. any pure y rotation: C4's boxes,
    // C5's 1000-sphere cluster) a sphere of the instance IS a sphere in world space and a quad a quad, with the
    // same ray parameter t. The builder then puts world-space copies of the instance's primitives into a second
    // pair of top levels (`top_baked`, `top_baked_inside`), each copy remembering (instance, inner primitive): the
    // fast kernels walk those, find the same closest hit up to rounding (t from the world-space copy, within an
    // ulp or two of the local one) and evaluate the Hit through the instance's frame as before. Not baked:
    // constant-density volumes (they need the probe walk), instances holding moving spheres, non-rigid Euler maps.
    // The strict kernels always walk the reference's structure (`top`). RTW_NO_BAKE=1 keeps the fast ones on it too.
    Level top_baked, top_baked_inside;
    bool has_baked = false;
    uint32_t n_spheres_orig = 0, n_quads_orig = 0;       // spheres / quads beyond these are baked copies
    std::vector<uint2> baked_sphere_src, baked_quad_src;
    uint32_t n_lights = 0;
    std::vector<float4> lights;                          // 2 per light: (box middle, bits(ref)) (emitted, 0); Scene::lights
    uint32_t n_walked_instances = 0, n_walked_instances_baked = 0;   // instance leaves walked as levels in top / top_baked
    float background[3];
    // the leaves of every level, for the optional on-device LBVH rebuild (rtw_lbvh.h)
    std::vector<LbvhLeaf> lbvh_leaves;
    std::vector<LbvhLevelDesc> lbvh_levels;
    rtw_scene_info info;
    uint32_t top_depth = 0, inner_depth = 0;   // deepest node chain of any top level / of any instance (-> stack_words)
    // the same for what the fast kernels walk: the baked top levels when there are any, and only the instances
    // that are still walked as levels there; fast_levels = is any instance walked as a level at all
    uint32_t fast_top_depth = 0, fast_inner_depth = 0;
    bool fast_levels = false, any_levels = false;
    std::vector<uint8_t> inst_baked;           // per instance: baked into the fast kernels' top level (not walked as a level there)

    // A Scene whose pointers address the vectors above (host memory): what the CPU-side unit
    // tests of the traversal code walk. The product uploads the same arrays to the device.
    Scene host_view();
};

// Validates `d`, flattens it and builds the BVH. Returns RTW_OK or RTW_E_*; `err` explains.
int build_host_scene(const rtw_scene_desc* d, HostScene& out, std::string& err);

// Viewport::new (viewport.rs:308-401)
void camera_init(rtw_camera* cam, uint32_t width, float aspect_ratio, float vfov_deg, const float origin[3],
                 const float direction[3], const float vup[3], float lens_radius);

// Integrator constants + derived launch parameters. Returns RTW_OK or RTW_E_INVALID.
struct RenderParams;
int make_render_params(const rtw_camera* cam, const rtw_render_opts* o, RenderParams& P, uint32_t& total_samples,
                       std::string& err);

}  // namespace rtw
