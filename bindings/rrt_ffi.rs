// UNBUILT here: this image has no rustc/cargo.  Kept in sync with include/rrt.h by
// tests/test_abi.py::test_rust_binding_declares_every_entry_point (names, field order, field types).
// `mod rrt_ffi;` in the reference's src/main.rs; replaces src/cuda/mod.rs and src/gpu/mod.rs.
use std::os::raw::{c_char, c_int, c_void};

#[repr(C)] #[derive(Clone, Copy, Default)]
pub struct RrtHittable { pub kind: i32, pub material: i32, pub first_child: i32, pub n_children: i32,
                         pub a: [f64; 3], pub b: [f64; 3], pub c: [f64; 3] }
#[repr(C)] #[derive(Clone, Copy, Default)]
pub struct RrtMaterial { pub kind: i32, pub texture: i32, pub albedo: [f64; 3], pub fuzz: f64, pub refraction_index: f64 }
#[repr(C)] #[derive(Clone, Copy, Default)]
pub struct RrtTexture { pub kind: i32, pub even: i32, pub odd: i32, pub image: i32, pub perlin: i32, pub _pad: i32,
                        pub color: [f64; 3], pub scale: f64 }
#[repr(C)] pub struct RrtImage { pub width: i32, pub height: i32, pub rgb: *const u8 }
#[repr(C)] pub struct RrtPerlin { pub randvec: [[f64; 3]; 256], pub perm_x: [i32; 256], pub perm_y: [i32; 256], pub perm_z: [i32; 256] }
#[repr(C)] pub struct RrtSceneDesc {
    pub hittables: *const RrtHittable, pub n_hittables: i32,
    pub children: *const i32,          pub n_children: i32,
    pub materials: *const RrtMaterial, pub n_materials: i32,
    pub textures: *const RrtTexture,   pub n_textures: i32,
    pub images: *const RrtImage,       pub n_images: i32,
    pub perlins: *const RrtPerlin,     pub n_perlins: i32,
    pub world: i32, pub lights: i32 }
#[repr(C)] #[derive(Clone, Copy)]
pub struct RrtCamera { pub aspect_ratio: f64, pub image_width: i32, pub samples_per_pixel: i32, pub max_depth: i32, pub _pad: i32,
                       pub background: [f64; 3], pub vfov: f64, pub lookfrom: [f64; 3], pub lookat: [f64; 3], pub vup: [f64; 3],
                       pub defocus_angle: f64, pub focus_dist: f64 }
#[repr(C)] #[derive(Clone, Copy)]
pub struct RrtRenderParams { pub integrator: i32, pub shard_mode: i32, pub rank: i32, pub world: i32, pub seed: u64 }
#[repr(C)] #[derive(Clone, Copy, Default)]
pub struct RrtStats { pub width: i32, pub height: i32, pub spp_effective: i32, pub kernel_launches: i32,
                      pub paths: f64, pub kernel_ms: f64, pub total_ms: f64, pub h2d_bytes: u64, pub d2h_bytes: u64,
                      pub n_prims: i32, pub n_bvh_nodes: i32, pub n_media: i32, pub n_lights: i32, pub sm_count: i32, pub kernel_features: i32,
                      pub n_devices: i32, pub n_grid_cells: i32, pub exchange_ms: f64 }
#[repr(C)] pub struct RrtCtx { _p: [u8; 0] }
#[repr(C)] pub struct RrtScene { _p: [u8; 0] }

extern "C" {
    pub fn rrt_init(device: c_int, out: *mut *mut RrtCtx) -> c_int;
    pub fn rrt_shutdown(ctx: *mut RrtCtx);
    pub fn rrt_init_multi(devices: *const c_int, n_devices: c_int, out: *mut *mut RrtCtx) -> c_int;
    pub fn rrt_comm_unique_id(id128: *mut u8) -> c_int;
    pub fn rrt_comm_init(ctx: *mut RrtCtx, id128: *const u8, rank: c_int, world: c_int) -> c_int;
    pub fn rrt_render_collective(ctx: *mut RrtCtx, scene: *const RrtScene, cam: *const RrtCamera, p: *const RrtRenderParams,
                                 accum_dev_root: *mut f32, accum_host_root: *mut f32) -> c_int;
    pub fn rrt_render_frame_collective(ctx: *mut RrtCtx, desc: *const RrtSceneDesc, cam: *const RrtCamera, p: *const RrtRenderParams, accum_host_root: *mut f32) -> c_int;
    pub fn rrt_tile_scatter_host(cam: *const RrtCamera, world: c_int, shares: *const *const f32, frame: *mut f32) -> c_int;
    pub fn rrt_set_stream(ctx: *mut RrtCtx, cuda_stream: *mut c_void) -> c_int;
    pub fn rrt_scene_create(ctx: *mut RrtCtx, desc: *const RrtSceneDesc, out: *mut *mut RrtScene) -> c_int;
    pub fn rrt_scene_destroy(scene: *mut RrtScene);
    pub fn rrt_image_height(cam: *const RrtCamera) -> c_int;
    pub fn rrt_render(ctx: *mut RrtCtx, scene: *const RrtScene, cam: *const RrtCamera, p: *const RrtRenderParams, accum_host: *mut f32) -> c_int;
    pub fn rrt_render_frame(ctx: *mut RrtCtx, desc: *const RrtSceneDesc, cam: *const RrtCamera, p: *const RrtRenderParams, accum_host: *mut f32) -> c_int;
    pub fn rrt_probe(ctx: *mut RrtCtx, scene: *const RrtScene, cam: *const RrtCamera, precision: c_int, prim_id: *mut i32, t: *mut f64) -> c_int;
    pub fn rrt_probe_static(ctx: *mut RrtCtx, scene: *const RrtScene, cam: *const RrtCamera, precision: c_int, prim_id: *mut i32, t: *mut f64) -> c_int;
    pub fn rrt_resolve_rgb8(ctx: *mut RrtCtx, accum_host: *const f32, w: c_int, h: c_int, spp: c_int, rgb: *mut u8) -> c_int;
    pub fn rrt_render_device(ctx: *mut RrtCtx, scene: *const RrtScene, cam: *const RrtCamera, p: *const RrtRenderParams, accum_dev: *mut f32) -> c_int;
    pub fn rrt_resolve_rgb8_device(ctx: *mut RrtCtx, accum_dev: *const f32, w: c_int, h: c_int, spp: c_int, rgb_dev: *mut u8) -> c_int;
    pub fn rrt_get_stats(ctx: *const RrtCtx, out: *mut RrtStats) -> c_int;
    pub fn rrt_shard_plan(cam: *const RrtCamera, p: *const RrtRenderParams, sample_begin: *mut i32, sample_end: *mut i32,
                          n_super_tiles: *mut i32, n_pixels: *mut f64, owner_mask: *mut u8) -> c_int;
    pub fn rrt_flatten_check(desc: *const RrtSceneDesc, n_prims: *mut i32, n_nodes: *mut i32, bvh_depth: *mut i32,
                             n_media: *mut i32, n_lights: *mut i32) -> c_int;
    pub fn rrt_scene_features(desc: *const RrtSceneDesc, features: *mut i32) -> c_int;
    pub fn rrt_flatten_rects(desc: *const RrtSceneDesc, n3: *mut i32) -> c_int;
    pub fn rrt_flatten_grid(desc: *const RrtSceneDesc, kind: c_int, res3: *mut i32, n_items: *mut i32, n_refs: *mut i32, max_per_cell: *mut i32) -> c_int;
    pub fn rrt_philox_kat(c0: u32, c1: u32, c2: u32, c3: u32, k0: u32, k1: u32, out4: *mut u32);
    pub fn rrt_philox_kat_rounds(rounds: c_int, c0: u32, c1: u32, c2: u32, c3: u32, k0: u32, k1: u32, out4: *mut u32) -> c_int;
    pub fn rrt_philox_rounds() -> c_int;
    pub fn rrt_kat_light(ctx: *mut RrtCtx, scene: *const RrtScene, n: c_int, origin: *const f64, dir: *const f64, u: *const f64,
                         pdf_each: *mut f64, pdf_list: *mut f64, random_dir: *mut f64) -> c_int;
    pub fn rrt_kat_rng(ctx: *mut RrtCtx, seed: u64, pixel: u32, sample0: u32, bounce: u32, block: u32, n: c_int,
                       raw4: *mut u32, uniform5: *mut f32) -> c_int;
    pub fn rrt_version() -> *const c_char;
    pub fn rrt_last_error() -> *const c_char;
}

pub fn last_error() -> String {
    unsafe { std::ffi::CStr::from_ptr(rrt_last_error()).to_string_lossy().into_owned() }
}
