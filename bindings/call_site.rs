// UNBUILT here: this image has no rustc/cargo.  Kept in sync with include/rrt.h by
// tests/test_abi.py.
//
// src/rrt_render.rs — the one call that replaces both GPU back ends of the reference:
//   src/main.rs:58-66   cuda::render_in_one_weekend()  -> render_with_rrt(...)
//   src/main.rs:67-85   gpu::render_in_one_weekend() and its CPU fallback -> removed (no fallback)
// and, called from each book's `run` in place of `cam.render(&world[, lights])`
// (in_one_weekend/mod.rs:125, the_next_week/mod.rs:623, the_rest_of_your_life/mod.rs:160), renders every
// book scene on the B200 library instead of the rayon loop of Camera::render.
use crate::render_io;
use crate::rrt_ffi::*;
use crate::rrt_flatten::SceneFlattener;

use crate::books::the_rest_of_your_life::camera::Camera;
use crate::books::the_rest_of_your_life::hittable::HittableRef;

/// Which book's `ray_color` / `get_ray` to run (include/rrt.h rrt_integrator).
#[derive(Clone, Copy)]
pub enum Integrator {
    InOneWeekend = 1,
    TheNextWeek = 2,
    TheRestOfYourLife = 3,
}

/// `devices`: CUDA device indices.  One entry renders on that GPU (rrt_init); several render one frame on
/// all of them — sample ranges reduced over NCCL, or interleaved tiles gathered, inside the library
/// (rrt_init_multi).  `tile_sharding` selects RRT_SHARD_TILE for the multi-GPU case.
pub fn render_with_rrt(
    cam: &Camera,
    world: &HittableRef,
    lights: Option<&HittableRef>,
    integrator: Integrator,
    devices: &[i32],
    tile_sharding: bool,
    seed: u64,
) -> Result<(), String> {
    // 1. the scene graph as flat arrays; `_keep` owns everything `desc` points into until the call returns
    let (desc, _keep) = SceneFlattener::new().finish(world, lights);
    let rcam = RrtCamera {
        aspect_ratio: cam.aspect_ratio,
        image_width: cam.image_width,
        samples_per_pixel: cam.samples_per_pixel,
        max_depth: cam.max_depth,
        _pad: 0,
        background: cam.background.e,
        vfov: cam.vfov,
        lookfrom: cam.lookfrom.e,
        lookat: cam.lookat.e,
        vup: cam.vup.e,
        defocus_angle: cam.defocus_angle,
        focus_dist: cam.focus_dist,
    };
    let params = RrtRenderParams {
        integrator: integrator as i32,
        shard_mode: if devices.len() > 1 && tile_sharding { 1 } else if devices.len() > 1 { 2 } else { 0 },
        rank: 0,
        world: 1,
        seed,
    };

    // 2. CudaDevice::new(0) (src/cuda/mod.rs:343) — or a group over several GPUs
    let mut ctx: *mut RrtCtx = std::ptr::null_mut();
    let rc = unsafe {
        if devices.len() <= 1 {
            rrt_init(devices.first().copied().unwrap_or(0), &mut ctx)
        } else {
            rrt_init_multi(devices.as_ptr(), devices.len() as i32, &mut ctx)
        }
    };
    if rc != 0 {
        return Err(format!("cuda init failed: {}", last_error())); // src/cuda/mod.rs:343 wording
    }

    // 3. upload + render + read back, one blocking call (the pass loop of src/cuda/mod.rs:403-436)
    let height = unsafe { rrt_image_height(&rcam) };
    let mut accum = vec![0f32; cam.image_width as usize * height as usize * 4];
    let rc = unsafe { rrt_render_frame(ctx, &desc, &rcam, &params, accum.as_mut_ptr()) };
    let mut stats = RrtStats::default();
    unsafe { rrt_get_stats(ctx, &mut stats) };
    unsafe { rrt_shutdown(ctx) };
    if rc != 0 {
        return Err(format!("kernel launch failed: {}", last_error())); // src/cuda/mod.rs:415-423 wording
    }
    eprintln!(
        "rrt: {:.0} paths in {:.1} ms on {} GPU(s) ({:.1} Mpaths/s), exchange {:.2} ms",
        stats.paths,
        stats.kernel_ms,
        stats.n_devices,
        stats.paths / stats.kernel_ms / 1e3,
        stats.exchange_ms
    );

    // 4. unchanged output path: accum is row-major RGBA f32, .w = samples (src/render_io.rs:3-11)
    render_io::write_ppm_from_accum(cam.image_width as u32, height as u32, &accum, stats.spp_effective as u32);
    Ok(())
}

/// src/main.rs:58-66 after the change:
///
/// ```ignore
/// if backend == "cuda" || backend == "gpu" {
///     let (cam, world, lights, integrator) = books::build(&book, scene);   // the builders of books/*/mod.rs, minus cam.render
///     if let Err(err) = rrt_render::render_with_rrt(&cam, &world, lights.as_ref(), integrator, &[0], false, 1) {
///         eprintln!("CUDA render failed: {err}");
///     }
///     return;
/// }
/// ```
pub fn main_rs_call_site_doc() {}
