// rtw_shade.h — "generate" and "shade/scatter": camera rays, the Hit of the winning primitive
// (normal, UV, texel, Perlin), the uber-material and the integrators, as an iterative path state
// machine.
//
// Restates, for one path:
//   render_row / Viewport::render / render_no_rand loop bodies   Rust/src/viewport.rs:287-300,455-470,502-508
//   Sphere / Quad::collision_normal after the root is accepted    objects/sphere.rs:124-146, quad.rs:61-80
//   Instance::collision_normal hit mapping + volume scatter       objects/instance.rs:280-307
//   ImageTexture::color_at, PerlinNoise::noise / perlin_interp    texture.rs:86-108,154-179,259-267
//   Material::on_hit, refract, reflectance, diffuse               objects/materials.rs:89-154,213-228
//   ray_color_gradient / ray_color_bg_color / ray_color_d         viewport/ray_color.rs:12-92, main.rs:17-46
// The recursion L = emitted + col_mod * L(next) becomes rad += thr*emitted; thr *= col_mod.
#pragma once
#include "rtw_trace.h"
#include "../../include/rtw.h"

namespace rtw {

// Per-launch constants (kernel parameter).
struct RenderParams {
    rtw_camera cam;
    uint32_t integrator, mode;
    uint32_t depth;
    uint32_t s_root;          // STRATIFIED: ceil(sqrt(samples))
    float mint, maxt;
    float time0;              // frame / fps
    float shutter_speed;
    uint32_t k0, k1;          // Philox key
    uint32_t row_begin, row_count;
    uint32_t sample_begin, sample_count;
    float light_weight;       // RTW_INT_LIGHT_BIASED: biased_weight
    uint32_t row_band, row_stride;   // interleaved row bands (rtw_render_opts); 0 / 0 = contiguous rows
};
// output (compact) row -> image row, and image pixel index -> position in the output buffer
RTW_HD uint32_t image_row(const RenderParams& P, uint32_t jr) {
    if (P.row_stride == 0u) return P.row_begin + jr;
    const uint32_t b = jr / P.row_band;
    return P.row_begin + b * P.row_stride + (jr - b * P.row_band);
}
RTW_HD uint32_t output_pixel(const RenderParams& P, uint32_t pixel) {
    const uint32_t W = P.cam.width;
    if (P.row_stride == 0u) return pixel - P.row_begin * W;
    const uint32_t j = pixel / W, rel = j - P.row_begin, b = rel / P.row_stride;
    return (b * P.row_band + (rel - b * P.row_stride)) * W + (pixel - j * W);
}

struct Hit {   // objects.rs:16-22
    float t;
    V3 normal, point, col_mod;
    float metallicness, opacity, ir;
    V3 emitted;
};

// ------------------------------------------------------------------ texture.rs
// perlin_interp (texture.rs:86-108)
RTW_HD float perlin_interp(const V3 c[2][2][2], float u, float v, float w) {
    float uu = u * u * (3.0f - 2.0f * u);
    float vv = v * v * (3.0f - 2.0f * v);
    float ww = w * w * (3.0f - 2.0f * w);
    float accum = 0.0f;
#pragma unroll
    for (int ii = 0; ii < 2; ++ii)
#pragma unroll
        for (int jj = 0; jj < 2; ++jj)
#pragma unroll
            for (int kk = 0; kk < 2; ++kk) {
                float i = (float)ii, j = (float)jj, k = (float)kk;
                V3 weight_v = v3(u - i, v - j, w - k);
                accum += (i * uu + (1.0f - i) * (1.0f - uu)) * (j * vv + (1.0f - j) * (1.0f - vv)) *
                         (k * ww + (1.0f - k) * (1.0f - ww)) * dot(c[ii][jj][kk], weight_v);
            }
    return accum;
}
RTW_HD int floor_to_int(float f) {
#if defined(__CUDA_ARCH__)
    return __float2int_rd(f);
#else
    float fl = floorf(f);
    if (!(fl == fl)) return 0;
    if (fl >= 2147483648.0f) return 2147483647;
    if (fl <= -2147483648.0f) return (int)0x80000000;
    return (int)fl;
#endif
}
// PerlinNoise::noise (texture.rs:154-179). `nz` may point at shared memory (staged tables).
RTW_HD float perlin_noise(const Noise* nz, V3 p) {
    float u = p.x - floorf(p.x), v = p.y - floorf(p.y), w = p.z - floorf(p.z);
    int i = floor_to_int(p.x), j = floor_to_int(p.y), k = floor_to_int(p.z);
    V3 c[2][2][2];
#pragma unroll
    for (int di = 0; di < 2; ++di)
#pragma unroll
        for (int dj = 0; dj < 2; ++dj)
#pragma unroll
            for (int dk = 0; dk < 2; ++dk) {
                uint32_t idx = (uint32_t)(nz->perm[0][(i + di) & 255] ^ nz->perm[1][(j + dj) & 255] ^
                                          nz->perm[2][(k + dk) & 255]) & 255u;
                float4 rv = nz->ranvec[idx];
                c[di][dj][dk] = v3(rv.x, rv.y, rv.z);
            }
    return perlin_interp(c, u, v, w);
}

// `f32 as usize` saturates: negative / NaN -> 0 (sphere.rs:137-138)
RTW_HD uint32_t to_usize(float f) {
    if (!(f > 0.0f)) return 0u;
    if (f >= 4294967040.0f) return 0xFFFFFFFFu;
    return (uint32_t)f;
}

// ImageTexture::color_at (texture.rs:259-267)
RTW_HD V3 texture_color_at(const Scene& sc, const Noise* noise_tables, const Texture& tx, uint32_t x, uint32_t y, V3 p) {
    float noise_mult = 1.0f;
    if (tx.noise_id >= 0) noise_mult = perlin_noise(&noise_tables[tx.noise_id], p / tx.noise_scale);
    V3 texel;
    if (tx.n_texels == 1u) {
        texel = v3(tx.solid[0], tx.solid[1], tx.solid[2]);
    } else {
        // idx = y*row + x, kept inside the image (the reference would panic past the end)
        uint64_t idx = (uint64_t)y * tx.width + x;
        if (idx >= tx.n_texels) idx = tx.n_texels - 1u;
#if defined(__CUDA_ARCH__)
        uint32_t yy = (uint32_t)(idx / tx.width), xx = (uint32_t)(idx - (uint64_t)yy * tx.width);
        float4 t = tex2D<float4>(tx.obj, (float)xx + 0.5f, (float)yy + 0.5f);
#else
        float4 t = tx.texels[idx];
#endif
        texel = v3(t.x, t.y, t.z);
    }
    return texel * noise_mult;
}

// ------------------------------------------------------------------ Hit of the winning primitive
RTW_HD void load_material(const Scene& sc, uint32_t mat_id, Hit& h) {
    const float4* mp = reinterpret_cast<const float4*>(&sc.materials[mat_id]);
    float4 m0 = RTW_LDG(mp), m1 = RTW_LDG(mp + 1);
    h.metallicness = m0.x; h.opacity = m0.y; h.ir = m0.z;
    h.emitted = v3(m1.x, m1.y, m1.z);
}

// sphere.rs:124-146
RTW_HD void eval_sphere(const Scene& sc, const Noise* nt, uint32_t idx, const Ray& r, float t, Hit& h) {
    float4 s = RTW_LDG(&sc.spheres[idx]);
    float4 col = RTW_LDG(&sc.sphere_col[idx]);
    uint2 ids = RTW_LDG(&sc.sphere_ids[idx]);
    V3 c = v3(s.x, s.y, s.z);
    if (f2u(col.w) & 1u) {
        float4 vel = RTW_LDG(&sc.sphere_vel[idx]);
        c = c + v3(vel.x, vel.y, vel.z) * r.time;
    }
    V3 p = r.o + r.d * t;
    V3 normal = unit(p - c);
    const Texture& tx = sc.textures[ids.y];
    uint32_t tex_x = 0u, tex_y = 0u;
    if (tx.n_texels != 1u) {
        const float PI = 3.14159265358979323846f, FRAC_1_PI = 0.31830988618379067154f;
        float u = (atan2f(-normal.z, normal.x) + PI) * FRAC_1_PI * 0.5f;
        float v = 1.0f - (FRAC_1_PI * acosf(-normal.y));
        tex_x = to_usize(floorf(u * (float)(tx.width - 1u)));
        tex_y = to_usize(floorf(v * (float)(tx.height - 1u)));
    }
    h.t = t; h.normal = normal; h.point = p;
    h.col_mod = texture_color_at(sc, nt, tx, tex_x, tex_y, p) * v3(col.x, col.y, col.z);
    load_material(sc, ids.x, h);
}

// quad.rs:61-80
RTW_HD void eval_quad(const Scene& sc, const Noise* nt, uint32_t idx, const Ray& r, float t, Hit& h) {
    const Quad* qp = &sc.quads[idx];
    float4 q0 = RTW_LDG(&qp->q0), q1 = RTW_LDG(&qp->q1), q2 = RTW_LDG(&qp->q2), q3 = RTW_LDG(&qp->q3);
    uint2 ids = RTW_LDG(&sc.quad_ids[idx]);
    V3 p = r.o + r.d * t;
    const Texture& tx = sc.textures[ids.y];
    uint32_t tex_x = 0u, tex_y = 0u;
    if (tx.n_texels != 1u) {
        V3 planar = p - v3(q1.x, q1.y, q1.z);
        V3 u = v3(q1.w, q2.x, q2.y), v = v3(q2.z, q2.w, q3.x), w = v3(q3.y, q3.z, q3.w);
        float alfa = dot(w, cross(planar, v));
        float beta = dot(w, cross(u, planar));
        tex_x = alfa != 1.0f ? to_usize(floorf(alfa * (float)tx.width)) : tx.width - 1u;
        tex_y = beta != 1.0f ? to_usize(floorf(beta * (float)tx.height)) : tx.height - 1u;
    }
    h.t = t; h.normal = v3(q0.x, q0.y, q0.z); h.point = p;
    h.col_mod = texture_color_at(sc, nt, tx, tex_x, tex_y, p);
    load_material(sc, ids.x, h);
}

RTW_HD void eval_prim_split(const Scene& sc, const Noise* nt, uint32_t ref, const Ray& r, float t, Hit& h) {
    if (ref_kind(ref) == KIND_SPHERE) eval_sphere(sc, nt, ref_index(ref), r, t, h);
    else eval_quad(sc, nt, ref_index(ref), r, t, h);
}

// The full Hit for the result of trace(), megakernel flavour (see eval_hit below).
RTW_HD void eval_hit_split(const TraceCtx& cx, const Noise* nt, const Best& b, const Ray& r, Hit& h) {
    const Scene& sc = *cx.sc;
    if (ref_kind(b.ref) != KIND_INSTANCE) { eval_prim_split(sc, nt, b.ref, r, b.t, h); return; }
    // instance.rs:257-307
    const Instance& in = sc.instances[ref_index(b.ref)];
    Ray l = to_local(in, r);
    eval_prim_split(sc, nt, b.inner, l, b.t, h);
    if (in.flags & INST_VOLUME) {
        float distance = volume_distance(in, cx);
        if (distance >= 0.0f) {
            // scatter inside the medium: entry-face t, col_mod and material are kept (instance.rs:295-299)
            h.point = h.point + l.d * distance;
            h.normal = random_unit_vec_stream(cx, cx.bounce << 16, 1u + in.index);
        }
    }
    if (!(in.flags & INST_ROT_IDENTITY)) {
        h.point = rot_apply(in.fwd_rows, h.point);
        h.normal = rot_apply(in.fwd_rows, h.normal);
    }
    h.point = h.point + v3(in.t[0], in.t[1], in.t[2]);
}

// The Hit of the winning primitive (sphere.rs:124-146, quad.rs:61-80), written so that a warp whose lanes hit different
// kinds of primitive diverges only where the kinds differ — the loads of the primitive, the normal, the quad's planar
// coordinates — and walks the texture lookup, the Perlin noise and the material load ONCE, together (as two functions
// that each ended in their own texel fetch and material load, inlined again behind the instance frame change, a warp
// ran up to four copies of that tail one after the other: segment_end held 64 % of wf_shade's warp instructions at 10
// lanes, profiles/r2n_wf_shade_by_line.txt). Same operations on the same values per path as the split flavour above.
// wf_shade: 3 320 -> 2 368 instructions, its launches 14 % shorter, C5 +4.3 %, C4 +2 %; the megakernel, whose lanes
// arrive here with fewer kinds per warp and which sits at its register cap, lost with it (C1 -10 %, C3 -2 %) and
// keeps the split flavour (profiles/r2v_ab_convergent_hit.txt): kOneTail of segment_end chooses.
RTW_HD void eval_prim(const Scene& sc, const Noise* nt, uint32_t ref, const Ray& r, float t, Hit& h) {
    const bool is_sphere = ref_kind(ref) == KIND_SPHERE;
    const uint32_t idx = ref_index(ref);
    const V3 p = r.o + r.d * t;
    V3 normal, col = v3(1.0f, 1.0f, 1.0f);
    uint2 ids;
    float alfa = 0.0f, beta = 0.0f;                 // quad.rs:66-67 (used only by a textured quad)
    if (is_sphere) {
        float4 s = RTW_LDG(&sc.spheres[idx]);
        float4 cl = RTW_LDG(&sc.sphere_col[idx]);
        ids = RTW_LDG(&sc.sphere_ids[idx]);
        V3 c = v3(s.x, s.y, s.z);
        if (f2u(cl.w) & 1u) {
            float4 vel = RTW_LDG(&sc.sphere_vel[idx]);
            c = c + v3(vel.x, vel.y, vel.z) * r.time;
        }
        normal = unit(p - c);
        col = v3(cl.x, cl.y, cl.z);
    } else {
        const Quad* qp = &sc.quads[idx];
        float4 q0 = RTW_LDG(&qp->q0), q1 = RTW_LDG(&qp->q1), q2 = RTW_LDG(&qp->q2), q3 = RTW_LDG(&qp->q3);
        ids = RTW_LDG(&sc.quad_ids[idx]);
        V3 planar = p - v3(q1.x, q1.y, q1.z);
        V3 u = v3(q1.w, q2.x, q2.y), v = v3(q2.z, q2.w, q3.x), w = v3(q3.y, q3.z, q3.w);
        alfa = dot(w, cross(planar, v));
        beta = dot(w, cross(u, planar));
        normal = v3(q0.x, q0.y, q0.z);
    }
    const Texture& tx = sc.textures[ids.y];
    uint32_t tex_x = 0u, tex_y = 0u;
    if (tx.n_texels != 1u) {
        if (is_sphere) {
            const float PI = 3.14159265358979323846f, FRAC_1_PI = 0.31830988618379067154f;
            float u = (atan2f(-normal.z, normal.x) + PI) * FRAC_1_PI * 0.5f;
            float v = 1.0f - (FRAC_1_PI * acosf(-normal.y));
            tex_x = to_usize(floorf(u * (float)(tx.width - 1u)));
            tex_y = to_usize(floorf(v * (float)(tx.height - 1u)));
        } else {
            tex_x = alfa != 1.0f ? to_usize(floorf(alfa * (float)tx.width)) : tx.width - 1u;
            tex_y = beta != 1.0f ? to_usize(floorf(beta * (float)tx.height)) : tx.height - 1u;
        }
    }
    h.t = t; h.normal = normal; h.point = p;
    const V3 texel = texture_color_at(sc, nt, tx, tex_x, tex_y, p);
    h.col_mod = is_sphere ? texel * col : texel;    // the quad has no colour of its own (quad.rs:76)
    load_material(sc, ids.x, h);
}

// The full Hit for the result of trace(): Scene::collision_normal's return value.
RTW_HD void eval_hit(const TraceCtx& cx, const Noise* nt, const Best& b, const Ray& r, Hit& h) {
    const Scene& sc = *cx.sc;
    const bool inst = ref_kind(b.ref) == KIND_INSTANCE;
    // instance.rs:257-307: the primitive's Hit in the instance's frame, then back into the world
    const Instance* in = inst ? &sc.instances[ref_index(b.ref)] : nullptr;
    Ray l = r;
    if (inst) l = to_local(*in, r);
    eval_prim(sc, nt, inst ? b.inner : b.ref, l, b.t, h);
    if (!inst) return;
    if (in->flags & INST_VOLUME) {
        float distance = volume_distance(*in, cx);
        if (distance >= 0.0f) {
            // scatter inside the medium: entry-face t, col_mod and material are kept (instance.rs:295-299)
            h.point = h.point + l.d * distance;
            h.normal = random_unit_vec_stream(cx, cx.bounce << 16, 1u + in->index);
        }
    }
    if (!(in->flags & INST_ROT_IDENTITY)) {
        h.point = rot_apply(in->fwd_rows, h.point);
        h.normal = rot_apply(in->fwd_rows, h.normal);
    }
    h.point = h.point + v3(in->t[0], in->t[1], in->t[2]);
}

// ------------------------------------------------------------------ materials.rs
RTW_HD V3 refract(V3 uv, V3 n, float etai_over_etat) {           // materials.rs:89-97
    float cos_theta = dot(-uv, n);
    if (cos_theta > 1.0f) cos_theta = 1.0f;
    V3 r_out_perp = (uv + n * cos_theta) * etai_over_etat;
    V3 r_out_parallel = n * -sqrtf(fabsf(1.0f - length2(r_out_perp)));
    return r_out_perp + r_out_parallel;
}
RTW_HD float reflectance(float cosine, float ref_idx) {          // materials.rs:98-103
    float r0 = (1.0f - ref_idx) / (1.0f + ref_idx);
    r0 = r0 * r0;
    float x = 1.0f - cosine;
    float x2 = x * x;
    float x4 = x2 * x2;
    return r0 + (1.0f - r0) * (x * x4);
}

// Material::on_hit (materials.rs:105-154): next direction and cos_theta.
RTW_HD V3 on_hit(const Hit& h, const Ray& r, PathRng& rng, float& cos_theta) {
    if (h.opacity > 0.0f) {
        V3 n; bool front_face;
        if (dot(r.d, h.normal) > 0.0f) { n = -h.normal; front_face = false; }
        else { n = h.normal; front_face = true; }
        float refraction_ratio = front_face ? 1.0f / h.ir : h.ir;
        V3 unit_direction = unit(r.d);
        float ct = dot(-unit_direction, n);
        if (ct > 1.0f) ct = 1.0f;
        float sin_theta = sqrtf(1.0f - ct * ct);
        bool cannot_refract = refraction_ratio * sin_theta > 1.0f;
        float refl = reflectance(ct, refraction_ratio);
        float coin = rng.next4().x;
        cos_theta = 0.0f;
        return (cannot_refract || refl > coin) ? reflect(unit_direction, n)
                                               : refract(unit_direction, n, refraction_ratio);
    }
    // diffuse (materials.rs:213-228)
    V3 target = h.normal + random_unit_vec(rng);
    V3 sc = close_to_zero(target) ? h.normal : target;
    V3 refl = reflect(unit(r.d), h.normal);
    cos_theta = h.metallicness != 1.0f ? dot(sc, h.normal) : 0.0f;
    return refl * h.metallicness + sc * (1.0f - h.metallicness);
}

// ------------------------------------------------------------------ camera (viewport.rs)
RTW_HD V3 ld3(const float* p) { return v3(p[0], p[1], p[2]); }

RTW_HD V3 sky_gradient(V3 d) {   // ray_color.rs:38-40
    V3 ud = unit(d);
    float t = 0.5f * (ud.y + 1.0f);
    return v3((1.0f - t) + t * 0.5f, (1.0f - t) + t * 0.7f, 1.0f);
}

// One path in flight. The Philox key lives in RenderParams (uniform); the counter here.
struct Path {
    Ray ray;         // the segment about to be traced (dead while a Trav carries it)
    V3 thr, rad;
    uint32_t pixel, sample, block;   // Philox counter words 0/1 and the next stream-0 block
    uint32_t depth_left;
    bool poison;     // pdf == 0 -> colour*0/0 = NaN all the way up
};
RTW_HD PathRng path_rng(const RenderParams& P, const Path& p) {
    PathRng r; r.k0 = P.k0; r.k1 = P.k1; r.pixel = p.pixel; r.sample = p.sample; r.block = p.block; return r;
}
RTW_HD TraceCtx path_ctx(const Scene& sc, const RenderParams& P, const Path& p, Work* work) {
    TraceCtx cx; cx.sc = &sc; cx.k0 = P.k0; cx.k1 = P.k1; cx.pixel = p.pixel; cx.sample = p.sample;
    cx.bounce = P.depth - p.depth_left; cx.mint = P.mint; cx.maxt = P.maxt; cx.work = work;
    return cx;
}

// Loop bodies of render_row (viewport.rs:287-300), Viewport::render (:455-470) and
// render_no_rand (:502-508): the camera ray of sample `s` of pixel (i, j).
RTW_HD void generate(const RenderParams& P, uint32_t i, uint32_t j, uint32_t s, Path& p) {
    p.pixel = j * P.cam.width + i; p.sample = s; p.block = 0u;
    PathRng rng = path_rng(P, p);
    V3 origin = ld3(P.cam.origin), p00 = ld3(P.cam.pixel00), du = ld3(P.cam.delta_u), dv = ld3(P.cam.delta_v);
    if (P.mode == RTW_MODE_NO_RAND) {
        p.ray.o = origin;
        p.ray.d = p00 + du * (float)i + dv * (float)j;
        p.ray.time = 0.0f;
    } else {
        float4 c = rng.next4();
        V3 rp = v3(0.f, 0.f, 0.f);
        if (P.cam.lens_radius != 0.0f) rp = random_in_unit_disk(rng);
        p.ray.o = origin + (ld3(P.cam.u) * rp.x + ld3(P.cam.v) * rp.y) * P.cam.lens_radius;
        if (P.mode == RTW_MODE_RANDOM) {
            p.ray.d = p00 + du * ((float)i + c.x) + dv * ((float)j + c.y);
            p.ray.time = P.time0 + P.shutter_speed * c.z;
        } else {
            uint32_t x = s / P.s_root, y = s % P.s_root;
            float sr = (float)P.s_root;
            p.ray.d = p00 + du * ((float)i + (((float)x + c.x) / sr)) + dv * ((float)j + (((float)y + c.y) / sr));
            p.ray.time = 0.0f;
        }
    }
    p.block = rng.block;
    p.thr = v3(1.f, 1.f, 1.f); p.rad = v3(0.f, 0.f, 0.f);
    p.depth_left = P.depth; p.poison = false;
}

// integrators that look at one segment only, whatever the depth
RTW_HD bool first_hit_integrator(uint32_t ig) { return ig == RTW_INT_NORMAL || ig == RTW_INT_ALBEDO || ig == RTW_INT_LIGHT_CAST; }
RTW_HD bool light_integrator(uint32_t ig) { return ig == RTW_INT_LIGHT_CAST || ig == RTW_INT_LIGHT_BIASED; }

// ------------------------------------------------------------------ Rust2's light-biased estimators
// Material::material_pdf (Rust2/src/objects/material.rs:17) for the uber-material of Rust/src: its diffuse branch
// answers like Lambertian (material.rs:44-63: clamp(+-cos, 0, 1) / pi, the sign by the side the incoming ray is on);
// a mirror (metallicness == 1, Mirror material.rs:93-101) and the dielectric (opacity > 0, MirrorGlass :199-235)
// answer 0 for every direction but their own reflection / refraction, i.e. for every light.
RTW_HD float material_pdf(const Hit& h, const Ray& in, V3 to_light) {
    if (h.opacity > 0.0f || h.metallicness == 1.0f) return 0.0f;
    float c = dot(unit(to_light), unit(h.normal));
    if (dot(in.d, h.normal) >= 0.0f) c = -c;
    c = c < 0.0f ? 0.0f : (c > 1.0f ? 1.0f : c);
    return c * 0.31830988618379067154f;
}
// The light loop of light_biased_ray_cast / light_biased_ray_color (Rust2/src/viewport/ray_color.rs:67-92, 130-152):
// one ray per light from the hit point towards the middle of the light's box (Ray::new: time 0); the light counts
// when the scene's closest hit along it is the light itself (`hr.0 == hl`). Returns sum(emitted * pdf / distance^2
// [* weight]) and, in `count`, the number (cast) or summed weight (biased) of the lights that counted.
// Needs a traversal stack: runs in the megakernel (and the host emulator), where the lane's own search is over.
RTW_HD_NOINL V3 light_gather(const TraceCtx& cx0, const RenderParams& P, const Hit& h, const Ray& in, const Stack& S,
                             bool biased, float& count, uint32_t& rays) {
    const Scene& sc = *cx0.sc;
    V3 color = v3(0.f, 0.f, 0.f);
    for (uint32_t l = 0; l < sc.n_lights; ++l) {
        const float4 lm = RTW_LDG(&sc.lights[2u * l]), le = RTW_LDG(&sc.lights[2u * l + 1u]);
        Ray r; r.o = h.point; r.d = unit(v3(lm.x, lm.y, lm.z) - h.point); r.time = 0.0f;
        TraceCtx cx = cx0;
        cx.bounce = 0x8000u | (cx0.bounce & 0x7FFFu);          // a volume on the way draws from its own block of the stream
        const Best b = trace(cx, r, S);
        ++rays;
        if (b.ref != f2u(lm.w)) continue;
        const V3 emitted = v3(le.x, le.y, le.z);
        const float pdf = material_pdf(h, in, r.d);
        const float distance2 = b.t * b.t * length2(r.d);
        if (biased) {
            const float mx = fmaxf(fmaxf(emitted.x, emitted.y), emitted.z);
            if (pdf <= 1.0f / (255.0f * mx)) continue;          // ray_color.rs:141-143
            count += P.light_weight;
            color = color + emitted * pdf / distance2 * P.light_weight;
        } else {
            count += 1.0f;
            color = color + emitted * pdf / distance2;
        }
    }
    return color;
}

// A path segment is: segment_begin (depth check, ray_color.rs:14-16; start the search) -> the search
// (rtw_trace.h, driven by the caller) -> segment_end (miss / Hit / scatter). segment_begin returns
// false when the path is over before tracing; segment_end returns true when the path is finished
// (p.rad final), else p.ray is the next segment.
RTW_HD bool segment_begin(const TraceCtx& cx, const RenderParams& P, const Path& p, Trav& T, const Stack& S) {
    if (!first_hit_integrator(P.integrator) && p.depth_left < 1u) return false;
    trav_begin(cx, p.ray, T, S);
    return true;
}

template <bool kOneTail = true>
RTW_HD bool segment_end(const TraceCtx& cx, const Noise* nt, const RenderParams& P, Path& p, const Trav& T,
                        const Stack* S = nullptr, uint32_t* shadow_rays = nullptr) {
    const Scene& sc = *cx.sc;
    const uint32_t ig = P.integrator;
    Ray ray; ray.o = T.o; ray.d = T.d; ray.time = p.ray.time;      // back in the world frame
    const Best b = trav_result(sc, T);
    if (b.ref == REF_NONE) {                                        // miss
        V3 bg;
        if (ig == RTW_INT_BG_COLOR || ig == RTW_INT_ALBEDO || light_integrator(ig)) bg = ld3(sc.background);   // ray_color.rs:91; Rust2 ray_color.rs:107,162
        else if (ig == RTW_INT_GLASS_TEST) bg = v3(0.f, 0.f, 1.f);                     // glass_tests.rs:54
        else bg = sky_gradient(ray.d);
        p.rad = p.rad + p.thr * bg;
        return true;
    }
    RTW_COUNT(cx, hits);
    Hit h;
    if (kOneTail) eval_hit(cx, nt, b, ray, h); else eval_hit_split(cx, nt, b, ray, h);
    if (ig == RTW_INT_LIGHT_CAST) {                                 // Rust2 ray_color.rs:61-105
        float count = 0.0f; uint32_t nr = 0u;
        V3 color = S ? light_gather(cx, P, h, ray, *S, false, count, nr) : v3(0.f, 0.f, 0.f);
        if (shadow_rays) *shadow_rays += nr;
        p.rad = (count != 0.0f ? color * h.col_mod / count : v3(0.f, 0.f, 0.f)) + h.emitted;
        return true;
    }
    if (first_hit_integrator(ig)) {                                 // viewport.rs:524-554
        if (ig == RTW_INT_NORMAL)
            p.rad = v3(0.5f * (h.normal.x + 1.0f), 0.5f * (h.normal.y + 1.0f), 0.5f * (h.normal.z + 1.0f));
        else p.rad = h.col_mod;
        return true;
    }
    if (ig == RTW_INT_GLASS_TEST && h.metallicness != 1.0f) {       // glass_tests.rs:17-19
        p.rad = p.rad + p.thr * v3(1.f, 1.f, 0.f);
        return true;
    }
    bool front = !(dot(ray.d, h.normal) > 0.0f);
    float cos_theta;
    PathRng rng = path_rng(P, p);
    V3 nd = on_hit(h, ray, rng, cos_theta);
    p.block = rng.block;
    if (close_to_zero(nd)) nd = front ? h.normal : h.normal * -1.0f;   // ray_color.rs:62-64
    if (ig == RTW_INT_BG_COLOR) {                                   // ray_color.rs:67-88
        p.rad = p.rad + p.thr * h.emitted;
        if (h.metallicness != 1.0f) {
            float pdf = cos_theta > 0.0f ? cos_theta * 0.31830988618379067154f : 0.0f;   // materials.rs:5-13
            if (!(pdf > 0.0f)) p.poison = true;
        }
    }
    if (ig == RTW_INT_LIGHT_BIASED) {                               // Rust2 ray_color.rs:124-160, iteratively:
        // L = emitted + col_mod / count * (L(reflected) + sum_l w * E_l * pdf_l / d_l^2), count = 1 + w * #lights that count
        float count = 1.0f; uint32_t nr = 0u;
        const V3 direct = S ? light_gather(cx, P, h, ray, *S, true, count, nr) : v3(0.f, 0.f, 0.f);
        if (shadow_rays) *shadow_rays += nr;
        p.rad = p.rad + p.thr * h.emitted;
        p.thr = p.thr * h.col_mod / count;
        p.rad = p.rad + p.thr * direct;
        p.ray.o = h.point; p.ray.d = nd;
        p.depth_left -= 1u;
        if (p.depth_left < 1u) { p.rad = p.rad + p.thr * ld3(sc.background); return true; }   // `depth <= 0 => bg_color` (:118-120)
        return false;
    }
    p.thr = p.thr * h.col_mod;
    p.ray.o = h.point; p.ray.d = nd;
    p.depth_left -= 1u;
    return false;
}

// Scalar form of one segment (the host emulator): returns true when the path is finished.
// `rays` counts Scene::collision_normal equivalents.
template <bool kOneTail = true>
RTW_HD bool path_step(const Scene& sc, const Noise* nt, const RenderParams& P, Path& p, uint32_t& rays,
                      const Stack& S, Work* work = nullptr) {
    TraceCtx cx = path_ctx(sc, P, p, work);
    Trav T;
    if (!segment_begin(cx, P, p, T, S)) return true;
    while (T.cur != CUR_DONE) {
        if (T.cur < NODE_LIMIT) trav_step_node<true>(cx, T, S);
        else trav_step_other(cx, p.ray.time, T, S);
    }
    ++rays;
    return segment_end<kOneTail>(cx, nt, P, p, T, &S, &rays);
}

RTW_HD V3 path_result(const Path& p) {
    if (p.poison) { float n = u2f(0x7FC00000u); return v3(n, n, n); }
    return p.rad;
}

}  // namespace rtw
