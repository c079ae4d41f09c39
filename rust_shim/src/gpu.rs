//! `impl Viewport { pub fn render_gpu(..) }` — the drop-in for Viewport::render / render_no_rand /
//! async_render (Rust/src/viewport.rs:215-305,430-516) on top of include/rtw.h.
//!
//! SOURCE ONLY: never compiled in this repository's build image (no Rust toolchain there). It is
//! written against the reference's types as they are at commit 66beb02c. It must live in child
//! modules of the modules that own the private fields it reads (Viewport's camera basis,
//! viewport.rs:56-67; Instance's contents, instance.rs:30-35; ImageTexture.img, texture.rs:22;
//! PerlinNoise tables, texture.rs:64-68) or those fields need `pub(crate)` getters — INTEGRATION.md
//! lists the five one-line accessors this file assumes (`raw_*`).
use crate::objects::{instance::Instance, quad::Quad, sphere::Sphere};
use crate::texture::texture::ImageTexture;
use crate::vec3::vec3::Vec3;
use crate::viewport::{Img, Scene, Viewport};
use image::Rgb;
use std::ffi::CStr;

use super::ffi::*;

/// The closures the reference passes as `ray_color`, closed into an enum (a GPU cannot run a Rust closure).
#[derive(Clone, Copy)]
pub enum Integrator { Gradient = 0, D = 1, BgColor = 2, Normal = 3, Albedo = 4, GlassTest = 5, LightCast = 6, LightBiased = 7 }
#[derive(Clone, Copy)]
pub enum Mode { Random = 0, Stratified = 1, NoRand = 2 }

fn f3(v: Vec3) -> [f32; 3] { [v.x, v.y, v.z] }

/// Flattened scene: owns the arrays the rtw_scene_desc points into.
pub struct FlatScene {
    spheres: Vec<rtw_sphere>, quads: Vec<rtw_quad>, instances: Vec<rtw_instance>,
    materials: Vec<rtw_material>, textures: Vec<rtw_texture>, texels: Vec<Vec<f32>>, noises: Vec<Box<rtw_noise>>,
    n_top_spheres: u32, n_top_quads: u32, background: [f32; 3],
}

impl FlatScene {
    fn texture(&mut self, t: &ImageTexture) -> u32 {
        // ImageTexture owns its texels (texture.rs:22); every Sphere clone deep-copies them, so
        // identical textures are de-duplicated here by content pointer + size
        let img: &[Vec3] = t.raw_img();
        let mut rgb = Vec::with_capacity(img.len() * 3);
        for p in img { rgb.extend_from_slice(&[p.x, p.y, p.z]); }
        let noise_id = match &t.noise {
            None => -1,
            Some(n) => {
                let mut z = Box::new(rtw_noise { ranvec: [[0.0; 3]; 256], perm_x: [0; 256], perm_y: [0; 256], perm_z: [0; 256] });
                for k in 0..256 {
                    z.ranvec[k] = f3(n.raw_ranvec()[k]);
                    z.perm_x[k] = n.raw_perm_x()[k] as u32; z.perm_y[k] = n.raw_perm_y()[k] as u32; z.perm_z[k] = n.raw_perm_z()[k] as u32;
                }
                self.noises.push(z);
                (self.noises.len() - 1) as i32
            }
        };
        self.texels.push(rgb);
        let rgb_ptr = self.texels.last().unwrap().as_ptr();
        self.textures.push(rtw_texture { width: t.row as u32, height: t.col as u32, rgb: rgb_ptr, noise_id, noise_scale: t.noise_scale });
        (self.textures.len() - 1) as u32
    }
    fn material(&mut self, m: &crate::objects::materials::Material) -> u32 {
        self.materials.push(rtw_material { metallicness: m.metallicness, opacity: m.opacity, ir: m.ir, emitted: f3(m.emmited) });
        (self.materials.len() - 1) as u32
    }
    fn sphere(&mut self, s: &Sphere) {
        let (mat_id, tex_id) = (self.material(&s.mat), self.texture(&s.texture));
        self.spheres.push(rtw_sphere { origin: f3(s.origin), radius: s.radius, col_mod: f3(s.col_mod), velocity: f3(s.velocity), mat_id, tex_id });
    }
    fn quad(&mut self, q: &Quad) {
        let (mat_id, tex_id) = (self.material(&q.mat), self.texture(&q.texture));
        self.quads.push(rtw_quad { origin: f3(q.origin), u: f3(q.u), v: f3(q.v), velocity: f3(q.velocity), mat_id, tex_id, shape: RTW_SHAPE_PARALLELOGRAM });
    }

    /// Scene (viewport.rs:79-88) -> the arrays of rtw_scene_desc. Top-level primitives first, then
    /// each instance's own (instance.rs:30-33).
    pub fn from_scene(scene: &Scene) -> FlatScene {
        let mut f = FlatScene { spheres: vec![], quads: vec![], instances: vec![], materials: vec![], textures: vec![],
                                texels: vec![], noises: vec![], n_top_spheres: scene.spheres.len() as u32,
                                n_top_quads: scene.quads.len() as u32, background: f3(scene.background_color) };
        for s in &scene.spheres { f.sphere(s); }
        for q in &scene.quads { f.quad(q); }
        for i in &scene.instances {
            let (fs, fq) = (f.spheres.len() as u32, f.quads.len() as u32);
            for s in i.raw_spheres() { f.sphere(s); }
            for q in i.raw_quads() { f.quad(q); }
            // dist_fn is an opaque &'static dyn Fn(f32)->f32 (instance.rs:19,37). `surface` always
            // returns exactly -1.0 (instance.rs:21-23); anything else is const_density (the only
            // other function the reference defines, instance.rs:24-26).
            let is_surface = (i.dist_fn)(1.0) == -1.0 && (i.dist_fn)(2.0) == -1.0;
            f.instances.push(rtw_instance { first_sphere: fs, n_spheres: f.spheres.len() as u32 - fs,
                                            first_quad: fq, n_quads: f.quads.len() as u32 - fq,
                                            translation: f3(i.gett()), rotation: f3(i.getr()),
                                            density_kind: if is_surface { 0 } else { 1 }, density: i.density });
        }
        f
    }
    fn desc(&self) -> rtw_scene_desc {
        // the texture rgb pointers were taken while `texels` was still growing: refresh them
        rtw_scene_desc { abi_version: RTW_ABI_VERSION,
            n_spheres: self.spheres.len() as u32, n_top_spheres: self.n_top_spheres,
            n_quads: self.quads.len() as u32, n_top_quads: self.n_top_quads, n_instances: self.instances.len() as u32,
            n_materials: self.materials.len() as u32, n_textures: self.textures.len() as u32, n_noises: self.noises.len() as u32,
            spheres: self.spheres.as_ptr(), quads: self.quads.as_ptr(), instances: self.instances.as_ptr(),
            materials: self.materials.as_ptr(), textures: self.textures.as_ptr(),
            noises: if self.noises.is_empty() { std::ptr::null() } else { &*self.noises[0] as *const rtw_noise },
            background: self.background, n_lights: 0, lights: std::ptr::null() }
    }
}

/// Device-resident scene; Drop releases the GPU memory.
pub struct GpuScene { h: *mut rtw_scene, _flat: FlatScene }
impl Drop for GpuScene { fn drop(&mut self) { unsafe { rtw_scene_destroy(self.h) } } }

fn check(rc: i32) -> Result<(), String> {
    if rc == RTW_OK { Ok(()) } else { Err(unsafe { CStr::from_ptr(rtw_last_error()) }.to_string_lossy().into_owned()) }
}

impl GpuScene {
    /// Replaces Scene::new / new_sphere / new_quad's BVH builds (viewport.rs:89-135).
    pub fn new(scene: &Scene, device: i32) -> Result<GpuScene, String> {
        let mut flat = FlatScene::from_scene(scene);
        for k in 0..flat.textures.len() { flat.textures[k].rgb = flat.texels[k].as_ptr(); }
        // noises are boxed one by one; the ABI wants them contiguous
        let contiguous: Vec<rtw_noise> = flat.noises.drain(..).map(|b| *b).collect();
        let mut d = flat.desc();
        d.noises = if contiguous.is_empty() { std::ptr::null() } else { contiguous.as_ptr() };
        let mut h: *mut rtw_scene = std::ptr::null_mut();
        check(unsafe { rtw_scene_create(&d, device, &mut h) })?;
        Ok(GpuScene { h, _flat: flat })
    }
}

/// The scene replicated on a list of CUDA devices of this process (include/rtw.h rtw_scene_multi): what
/// `presentation_image` / `test_run` (main.rs:76-80,412-414) use to put every GPU of the box behind one call.
pub struct GpuSceneMulti { h: *mut rtw_scene_multi, _flat: FlatScene }
impl Drop for GpuSceneMulti { fn drop(&mut self) { unsafe { rtw_scene_multi_destroy(self.h) } } }
#[derive(Clone, Copy)]
pub enum Split { Samples = 0, Tiles = 1 }

impl GpuSceneMulti {
    pub fn new(scene: &Scene, devices: &[i32]) -> Result<GpuSceneMulti, String> {
        let mut flat = FlatScene::from_scene(scene);
        for k in 0..flat.textures.len() { flat.textures[k].rgb = flat.texels[k].as_ptr(); }
        let contiguous: Vec<rtw_noise> = flat.noises.drain(..).map(|b| *b).collect();
        let mut d = flat.desc();
        d.noises = if contiguous.is_empty() { std::ptr::null() } else { contiguous.as_ptr() };
        let mut h: *mut rtw_scene_multi = std::ptr::null_mut();
        check(unsafe { rtw_scene_multi_create(&d, devices.as_ptr(), devices.len() as i32, &mut h) })?;
        Ok(GpuSceneMulti { h, _flat: flat })
    }
}

impl Viewport {
    fn gpu_opts(&self, integrator: Integrator, mode: Mode, seed: u64) -> rtw_render_opts {
        let mut o: rtw_render_opts = unsafe { std::mem::zeroed() };
        unsafe { rtw_render_opts_default(&mut o) };
        o.samples = self.samples as u32; o.depth = self.depth as u32; o.gamma = self.gamma;
        o.integrator = integrator as u32; o.mode = mode as u32;
        o.shutter_speed = self.shutter_speed; o.fps = self.fps; o.frame = self.frame as u32; o.seed = seed;
        o
    }
    fn to_img(&self, buf: &[f32]) -> Img {
        let (w, h) = (self.width as usize, self.height as usize);
        (0..h).map(|j| (0..w).map(|i| { let k = (j * w + i) * 3; Rgb([buf[k], buf[k + 1], buf[k + 2]]) }).collect()).collect()
    }

    /// async_render / Viewport::render over every device of `scene`: sample ranges (one peer-memory combine on the
    /// first device) or round-robin row bands (no exchange between devices).
    pub fn render_gpu_multi(&self, integrator: Integrator, mode: Mode, scene: &GpuSceneMulti, split: Split, seed: u64) -> Result<Img, String> {
        let cam = self.gpu_camera();
        let o = self.gpu_opts(integrator, mode, seed);
        let mut buf = vec![0f32; self.width as usize * self.height as usize * 3];
        check(unsafe { rtw_render_multi(scene.h, &cam, &o, split as u32, 0, buf.as_mut_ptr()) })?;
        Ok(self.to_img(&buf))
    }

    /// render_multi (viewport.rs:249-269) with frame k on device k mod n: one call, no exchange between devices.
    pub fn render_multi_gpu_frames(&self, integrator: Integrator, scene: &GpuSceneMulti, seed: u64) -> Result<Vec<Img>, String> {
        let cam = self.gpu_camera();
        let o = self.gpu_opts(integrator, Mode::Random, seed);
        let n = self.width as usize * self.height as usize * 3;
        let mut buf = vec![0f32; n * self.number_of_frames];
        check(unsafe { rtw_render_frames_multi(scene.h, &cam, &o, self.start_frame as u32, self.number_of_frames as u32, buf.as_mut_ptr()) })?;
        Ok(buf.chunks(n).map(|f| self.to_img(f)).collect())
    }

    fn gpu_camera(&self) -> rtw_camera {
        // the private derived fields of Viewport (viewport.rs:56-67), copied verbatim: pixel00 is
        // `upper_left_corner` (already offset by half a pixel, viewport.rs:382-387)
        rtw_camera { width: self.width as u32, height: self.height as u32, origin: f3(self.raw_origin()),
                     pixel00: f3(self.raw_upper_left_corner()), delta_u: f3(self.raw_p_delta_u()),
                     delta_v: f3(self.raw_p_delta_v()), u: f3(self.raw_u()), v: f3(self.raw_v()),
                     lens_radius: self.raw_lens_radius() }
    }

    /// Same inputs and output as Viewport::render(&self, ray_color, &Scene) -> Img, with the closure
    /// replaced by an Integrator and the loop flavour by a Mode.
    pub fn render_gpu(&self, integrator: Integrator, mode: Mode, scene: &GpuScene, seed: u64) -> Result<Img, String> {
        let cam = self.gpu_camera();
        let mut o: rtw_render_opts = unsafe { std::mem::zeroed() };
        unsafe { rtw_render_opts_default(&mut o) };
        o.samples = self.samples as u32; o.depth = self.depth as u32; o.gamma = self.gamma;
        o.integrator = integrator as u32; o.mode = mode as u32;
        o.shutter_speed = self.shutter_speed; o.fps = self.fps; o.frame = self.frame as u32; o.seed = seed;
        let (w, h) = (self.width as usize, self.height as usize);
        let mut buf = vec![0f32; w * h * 3];
        check(unsafe { rtw_render(scene.h, &cam, &o, buf.as_mut_ptr()) })?;
        Ok((0..h).map(|j| (0..w).map(|i| { let k = (j * w + i) * 3; Rgb([buf[k], buf[k + 1], buf[k + 2]]) }).collect()).collect())
    }

    /// render_multi (viewport.rs:249-269): frames start_frame..start_frame+number_of_frames.
    pub fn render_multi_gpu(&self, integrator: Integrator, scene: &GpuScene, seed: u64) -> Result<Vec<Img>, String> {
        let mut out = Vec::with_capacity(self.number_of_frames);
        let mut vp = self.clone();
        for k in self.start_frame..self.start_frame + self.number_of_frames {
            vp.frame = k;
            out.push(vp.render_gpu(integrator, Mode::Random, scene, seed)?);
        }
        Ok(out)
    }
}
