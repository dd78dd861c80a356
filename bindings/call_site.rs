// UNBUILT here: this image has no rustc/cargo.  Kept in sync with include/rrt.h by
// tests/test_abi.py::test_rust_binding_declares_every_entry_point.  See INTEGRATION.md.
// Replaces src/main.rs:58-85.
let (desc, _keep) = SceneFlattener::new().finish(&world, lights.as_ref());  // _keep owns the Vecs desc points into
let cam = RrtCamera { aspect_ratio: cam.aspect_ratio, image_width: cam.image_width, samples_per_pixel: cam.samples_per_pixel,
                      max_depth: cam.max_depth, _pad: 0, background: cam.background.e, vfov: cam.vfov, lookfrom: cam.lookfrom.e,
                      lookat: cam.lookat.e, vup: cam.vup.e, defocus_angle: cam.defocus_angle, focus_dist: cam.focus_dist };
let params = RrtRenderParams { integrator: 3 /* book */, shard_mode: 0, rank: 0, world: 1, seed: 1 };
let mut ctx = std::ptr::null_mut();
if unsafe { rrt_init(0, &mut ctx) } != 0 { return Err(format!("cuda init failed: {}", last_error())); }
let h = unsafe { rrt_image_height(&cam) };
let mut accum = vec![0f32; (cam.image_width * h * 4) as usize];
let rc = unsafe { rrt_render_frame(ctx, &desc, &cam, &params, accum.as_mut_ptr()) };
unsafe { rrt_shutdown(ctx) };
if rc != 0 { return Err(format!("kernel launch failed: {}", last_error())); }
let spp = if params.integrator == 3 { let r = (cam.samples_per_pixel as f64).sqrt() as u32; r * r } else { cam.samples_per_pixel as u32 };
render_io::write_ppm_from_accum(cam.image_width as u32, h as u32, &accum, spp);   // unchanged
