//! Safe wrapper over `libraysnail_b200.so` for the raysnail crate.
//!
//! `SceneBuilder` is what the additive `Hittable::flatten` / `Material::flatten` /
//! `Texture::flatten` methods push into; `Context::render` is the body of
//! `TakePhotoSettings::shot_to_target_gpu`.  The tested twin of this file is
//! `raysnail_b200/csrc/host/scene.cpp` (`FlatScene::from_world`) + `raysnail_b200/api.py`.
use raysnail_b200_sys as sys;
use std::ffi::CStr;
use std::ptr;

#[derive(Debug)]
pub struct Error {
    pub status: i32,
    pub message: String,
}

/// A `Hittable` the backend cannot express (a user-defined implementor outside the reference's
/// own set — every shipped one, `RayMarcher` included, has a node type): the caller falls back to
/// `shot_to_target`.
#[derive(Debug)]
pub struct Unsupported(pub &'static str);

/// The flattened `World` (include/raysnail_b200.h: `rs_scene_desc`).
#[derive(Default)]
pub struct SceneBuilder {
    pub nodes: Vec<sys::rs_node>,
    pub roots: Vec<i32>,
    pub root_bbox: Vec<f64>,
    pub lights: Vec<i32>,
    pub xforms: Vec<sys::rs_xform>,
    pub tri_verts: Vec<f64>,
    pub tri_normals: Vec<f64>,
    pub tri_material: Vec<i32>,
    pub materials: Vec<sys::rs_material>,
    pub textures: Vec<sys::rs_texture>,
    pub texels: Vec<u8>,
    pub perlin_perm: Vec<i32>,
    pub perlin_values: Vec<f64>,
    pub default_material: i32,
    pub bg0: [f32; 4],
    pub bg1: [f32; 4],
}

impl SceneBuilder {
    pub fn push_node(&mut self, n: sys::rs_node) -> i32 {
        self.nodes.push(n);
        (self.nodes.len() - 1) as i32
    }
    pub fn push_material(&mut self, m: sys::rs_material) -> i32 {
        self.materials.push(m);
        (self.materials.len() - 1) as i32
    }
    pub fn push_texture(&mut self, t: sys::rs_texture) -> i32 {
        self.textures.push(t);
        (self.textures.len() - 1) as i32
    }
    /// `TfFacade`: one entry per `TransformStack` element, in stack order (transform.rs:133-157).
    pub fn push_xforms(&mut self, stack: &[sys::rs_xform]) -> (i32, i32) {
        let first = self.xforms.len() as i32;
        self.xforms.extend_from_slice(stack);
        (first, stack.len() as i32)
    }
    /// One `Triangle` (triangle_mesh.rs:17-60): p0 p1 p2, their normals, material (-1 = None).
    /// Returns the root reference (`!index`) to pass to `push_root`.
    pub fn push_triangle(&mut self, verts: [f64; 9], normals: [f64; 9], material: i32) -> i32 {
        self.tri_verts.extend_from_slice(&verts);
        self.tri_normals.extend_from_slice(&normals);
        self.tri_material.push(material);
        !((self.tri_material.len() - 1) as i32)
    }
    /// An object of the world's `HittableList`, in list order; `bbox` = its `Hittable::bbox`
    /// (min xyz, max xyz).  The position in this list is the object id of the parity probes.
    pub fn push_root(&mut self, node_or_triangle: i32, bbox: [f64; 6]) {
        self.roots.push(node_or_triangle);
        self.root_bbox.extend_from_slice(&bbox);
    }
    pub fn push_light(&mut self, node: i32) {
        self.lights.push(node);
    }
    /// The descriptor borrows `self`: keep the builder alive until `upload_scene` returns.
    pub fn desc(&self) -> sys::rs_scene_desc {
        sys::rs_scene_desc {
            nodes: self.nodes.as_ptr(),
            n_nodes: self.nodes.len() as i32,
            roots: self.roots.as_ptr(),
            root_bbox: self.root_bbox.as_ptr(),
            n_roots: self.roots.len() as i32,
            lights: self.lights.as_ptr(),
            n_lights: self.lights.len() as i32,
            xforms: self.xforms.as_ptr(),
            n_xforms: self.xforms.len() as i32,
            tri_verts: self.tri_verts.as_ptr(),
            tri_normals: self.tri_normals.as_ptr(),
            tri_material: self.tri_material.as_ptr(),
            n_tris: self.tri_material.len() as i32,
            materials: self.materials.as_ptr(),
            n_materials: self.materials.len() as i32,
            textures: self.textures.as_ptr(),
            n_textures: self.textures.len() as i32,
            texels: self.texels.as_ptr(),
            n_texel_bytes: self.texels.len() as i64,
            perlin_perm: self.perlin_perm.as_ptr(),
            n_perlin_perm: self.perlin_perm.len() as i32,
            perlin_values: self.perlin_values.as_ptr(),
            n_perlin_values: self.perlin_values.len() as i32,
            default_material: self.default_material,
            bg0: self.bg0,
            bg1: self.bg1,
        }
    }
}

/// One CUDA device.  Not `Sync`: one host thread at a time (include/raysnail_b200.h).
pub struct Context {
    ctx: *mut sys::rs_ctx,
    /// Base Philox key.  Render call k uses `rs_pass_seed(seed, k)`, so that consecutive
    /// `shot_to_target` calls (the passes loop, raysnail.rs:379-427) see fresh random numbers the way
    /// the reference's freshly seeded FastRng does — with a constant key every pass would return
    /// bit-identical pixels and averaging them would never reduce the noise.
    pub seed: u64,
    /// number of `render` calls so far (the pass index mixed into the key)
    pub pass: i32,
}

impl Context {
    /// Fails without an sm_100-class GPU: the backend has no CPU fallback.
    pub fn new(device: i32) -> Result<Context, Error> {
        sys::check_abi().map_err(|message| Error { status: sys::RS_ERR_INVALID_ARGUMENT, message })?;
        let mut ctx: *mut sys::rs_ctx = ptr::null_mut();
        let rc = unsafe { sys::rs_create(&mut ctx, device) };
        if rc != sys::RS_OK {
            return Err(Error { status: rc, message: last_error(ptr::null()) });
        }
        Ok(Context { ctx, seed: 0x5EED_0007, pass: 0 })
    }
    fn check(&self, rc: i32) -> Result<(), Error> {
        if rc == sys::RS_OK {
            Ok(())
        } else {
            Err(Error { status: rc, message: last_error(self.ctx) })
        }
    }
    /// Replaces `World::new` + `BVH::new` (world.rs:40-53, bvh.rs:47-112).
    pub fn upload_scene(&mut self, scene: &SceneBuilder) -> Result<(), Error> {
        let d = scene.desc();
        self.check(unsafe { sys::rs_upload_scene(self.ctx, &d) })
    }
    /// Replaces the body of `shot_to_target` (camera.rs:261-287): returns the row-major
    /// `Vec<[f32; 4]>` the painter returns.  `samples` is the requested count (`--samples`):
    /// `Painter::samples` floors it to a square (painter.rs:110-118).  `mask[y * w + x] == 0`
    /// is a pixel `PixelController::calculate_pixel` declined (painter.rs:204-210).
    pub fn render(&mut self, cam: &sys::rs_camera, samples: usize, depth: usize, gamma: bool, mask: Option<&[u8]>) -> Result<Vec<[f32; 4]>, Error> {
        let sqrt_spp = (samples as f64).sqrt().floor() as i32;
        let params = sys::rs_render_params {
            sqrt_spp,
            depth: depth as i32,
            gamma: gamma as i32,
            stratum_row_begin: 0,
            stratum_row_end: sqrt_spp,
            max_wave_paths: 0,
            seed: unsafe { sys::rs_pass_seed(self.seed, self.pass) },
            flags: 0,
            reserved: 0,
        };
        let n = cam.width as usize * cam.height as usize;
        if let Some(m) = mask {
            if m.len() != n {
                return Err(Error { status: sys::RS_ERR_INVALID_ARGUMENT, message: "pixel mask size".into() });
            }
        }
        let mut out = vec![[0f32; 4]; n];
        let rc = unsafe { sys::rs_render(self.ctx, cam, &params, mask.map_or(ptr::null(), |m| m.as_ptr()), out.as_mut_ptr() as *mut f32) };
        self.check(rc)?;
        self.pass += 1;
        Ok(out)
    }
    /// The whole passes loop (raysnail.rs:379-427) on the device; `redo_mode`: `sys::RS_REDO_*`.
    pub fn render_passes(&mut self, cam: &sys::rs_camera, samples: usize, depth: usize, passes: usize, redo_mode: i32, threshold: f32) -> Result<(Vec<[f32; 4]>, Vec<i64>), Error> {
        let sqrt_spp = (samples as f64).sqrt().floor() as i32;
        let params = sys::rs_render_params { sqrt_spp, depth: depth as i32, gamma: 1, stratum_row_begin: 0, stratum_row_end: sqrt_spp, max_wave_paths: 0, seed: self.seed, flags: 0, reserved: 0 };
        let n = cam.width as usize * cam.height as usize;
        let mut out = vec![[0f32; 4]; n];
        let mut counts = vec![0i64; passes.max(1)];
        let rc = unsafe { sys::rs_render_passes(self.ctx, cam, &params, passes as i32, redo_mode, threshold, out.as_mut_ptr() as *mut f32, ptr::null_mut(), counts.as_mut_ptr()) };
        self.check(rc)?;
        counts.truncate(passes);
        Ok((out, counts))
    }
    pub fn stats(&self) -> sys::rs_stats {
        let mut s = sys::rs_stats::default();
        unsafe { sys::rs_get_stats(self.ctx, &mut s) };
        s
    }
    pub fn raw(&self) -> *mut sys::rs_ctx {
        self.ctx
    }
}

impl Drop for Context {
    fn drop(&mut self) {
        unsafe { sys::rs_destroy(self.ctx) }
    }
}

fn last_error(ctx: *const sys::rs_ctx) -> String {
    let p = unsafe { sys::rs_last_error(ctx) };
    if p.is_null() {
        String::new()
    } else {
        unsafe { CStr::from_ptr(p) }.to_string_lossy().into_owned()
    }
}


/// Philox4x32-10 exactly as the kernels evaluate it (`rs_philox4x32_10`), in plain Rust: a reference
/// harness that swaps `FastRng` (src/prelude/random.rs:109-145) for this draws the very numbers a
/// GPU path sees.  Stream layout: key = (seed lo, seed hi), counter = (pixel, sample, depth,
/// purpose | call << 8); see the RS_RNG_* purposes in include/raysnail_b200.h.
pub fn philox4x32_10(counter: [u32; 4], key: [u32; 2]) -> [u32; 4] {
    let (mut c, mut k) = (counter, key);
    for _ in 0..10 {
        let p0 = 0xD251_1F53u64 * c[0] as u64;
        let p1 = 0xCD9E_8D57u64 * c[2] as u64;
        c = [(p1 >> 32) as u32 ^ c[1] ^ k[0], p1 as u32, (p0 >> 32) as u32 ^ c[3] ^ k[1], p0 as u32];
        k = [k[0].wrapping_add(0x9E37_79B9), k[1].wrapping_add(0xBB67_AE85)];
    }
    c
}

/// The four uniforms of one draw site: `(word >> 8) * 2^-24`, exact in f32 and f64.
pub fn philox_uniform(seed: u64, pixel: u32, sample: u32, depth: u32, purpose: u32, call: u32) -> [f64; 4] {
    let w = philox4x32_10([pixel, sample, depth, purpose | (call << 8)], [seed as u32, (seed >> 32) as u32]);
    [w[0], w[1], w[2], w[3]].map(|x| (x >> 8) as f64 * (1.0 / 16_777_216.0))
}
