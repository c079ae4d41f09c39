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
    float background[3];
    // the leaves of every level, for the optional on-device LBVH rebuild (rtw_lbvh.h)
    std::vector<LbvhLeaf> lbvh_leaves;
    std::vector<LbvhLevelDesc> lbvh_levels;
    rtw_scene_info info;
    uint32_t top_depth = 0, inner_depth = 0;   // deepest node chain of the top level / of any instance (-> stack_words)

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
