// main.cpp — the `rrt` command line: same shape as the reference's src/main.rs:14-99
//     rrt [--backend cpu|gpu|cuda | --cpu | --gpu | --cuda] <book> [scene]  > image.ppm
// with the same book aliases (main.rs:7-12, 87-92) and the_next_week scene numbers.  Every backend
// name selects the one implementation this build has — the B200 library behind include/rrt.h —
// because it replaces src/cuda and src/gpu and there is no CPU renderer to fall back to
// (main.rs:77-84's fallback is deliberately gone).  Where the reference takes its render settings
// from the compile-time const OVERRIDES (src/config.rs:50-62), this CLI takes the same eleven
// fields as flags; with no flags the shipped override values apply, as in the reference.
// Output: text "P3" PPM on stdout (render_io.rs:3-30), progress/timing on stderr.
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include <unistd.h>

#include "../../include/rrt.h"
#include "../../include/rrt_host.h"

static bool parse3(const char* s, double* v) { return std::sscanf(s, "%lf,%lf,%lf", v, v + 1, v + 2) == 3; }

static void usage() {
  std::fprintf(stderr, "Usage: rrt [--backend cpu|gpu|cuda] <book> [scene] [render flags] > image.ppm\n");
  std::fprintf(stderr, "books: in_one_weekend, the_next_week, the_rest_of_your_life\n");
  std::fprintf(stderr, "example: rrt the_next_week 3\n");
  std::fprintf(stderr, "render flags (RenderOverrides, src/config.rs:1-14): --width N --spp N --depth N --aspect R --vfov D\n"
                       "  --lookfrom x,y,z --lookat x,y,z --vup x,y,z --defocus-angle D --focus-dist F --background r,g,b\n"
                       "  --no-overrides (scene defaults)  --seed N (scene RNG)  --sample-seed N  --device N  -o FILE  --probe\n"
                       "  --p6             binary PPM (P6; gamma and quantisation on the GPU, rrt_resolve_rgb8) instead of the text P3\n"
                       "  --devices 0,1,..  render on several GPUs of this box (sample ranges + NCCL reduce; --shard tile: interleaved\n"
                       "                    32x32 tiles + gather)\n");
}

int main(int argc, char** argv) {
  std::string backend = "cuda";
  std::vector<std::string> positional;
  rrth_overrides ov;
  rrth_overrides_shipped(&ov);
  uint64_t scene_seed = 0x5EED1234ull, sample_seed = 1;
  int device = 0;
  std::vector<int> devices;   // --devices: a multi-GPU group (rrt_init_multi)
  int shard_mode = RRT_SHARD_NONE;
  const char* out_path = nullptr;
  bool probe = false, p6 = false;
  for (int i = 1; i < argc; ++i) {
    std::string a = argv[i];
    auto next = [&](const char* what) -> const char* {
      if (i + 1 >= argc) { std::fprintf(stderr, "%s expects a value\n", what); std::exit(2); }
      return argv[++i];
    };
    if (a == "--gpu" || a == "--cuda" || a == "--cpu") { backend = a.substr(2); continue; }
    if (a == "--backend") { backend = next("--backend"); continue; }
    if (a.rfind("--backend=", 0) == 0) { backend = a.substr(10); continue; }
    if (a == "--no-overrides") { rrth_overrides_none(&ov); continue; }
    if (a == "--width") { ov.image_width = std::atoi(next("--width")); ov.has |= RRTH_O_IMAGE_WIDTH; continue; }
    if (a == "--spp") { ov.samples_per_pixel = std::atoi(next("--spp")); ov.has |= RRTH_O_SAMPLES_PER_PIXEL; continue; }
    if (a == "--depth") { ov.max_depth = std::atoi(next("--depth")); ov.has |= RRTH_O_MAX_DEPTH; continue; }
    if (a == "--aspect") { ov.aspect_ratio = std::atof(next("--aspect")); ov.has |= RRTH_O_ASPECT_RATIO; continue; }
    if (a == "--vfov") { ov.vfov = std::atof(next("--vfov")); ov.has |= RRTH_O_VFOV; continue; }
    if (a == "--defocus-angle") { ov.defocus_angle = std::atof(next("--defocus-angle")); ov.has |= RRTH_O_DEFOCUS_ANGLE; continue; }
    if (a == "--focus-dist") { ov.focus_dist = std::atof(next("--focus-dist")); ov.has |= RRTH_O_FOCUS_DIST; continue; }
    if (a == "--lookfrom") { if (!parse3(next("--lookfrom"), ov.lookfrom)) { usage(); return 2; } ov.has |= RRTH_O_LOOKFROM; continue; }
    if (a == "--lookat") { if (!parse3(next("--lookat"), ov.lookat)) { usage(); return 2; } ov.has |= RRTH_O_LOOKAT; continue; }
    if (a == "--vup") { if (!parse3(next("--vup"), ov.vup)) { usage(); return 2; } ov.has |= RRTH_O_VUP; continue; }
    if (a == "--background") { if (!parse3(next("--background"), ov.background)) { usage(); return 2; } ov.has |= RRTH_O_BACKGROUND; continue; }
    if (a == "--seed") { scene_seed = std::strtoull(next("--seed"), nullptr, 0); continue; }
    if (a == "--sample-seed") { sample_seed = std::strtoull(next("--sample-seed"), nullptr, 0); continue; }
    if (a == "--device") { device = std::atoi(next("--device")); continue; }
    if (a == "--devices") {
      const char* v = next("--devices");
      devices.clear();
      for (const char* q = v; *q;) {
        char* end = nullptr;
        long d = std::strtol(q, &end, 10);
        if (end == q) { std::fprintf(stderr, "--devices expects a comma-separated list of device indices\n"); return 2; }
        devices.push_back((int)d);
        q = *end == ',' ? end + 1 : end;
      }
      continue;
    }
    if (a == "--shard") {
      std::string v = next("--shard");
      if (v == "tile") shard_mode = RRT_SHARD_TILE;
      else if (v == "sample") shard_mode = RRT_SHARD_SAMPLE;
      else { std::fprintf(stderr, "--shard expects tile or sample\n"); return 2; }
      continue;
    }
    if (a == "-o") { out_path = next("-o"); continue; }
    if (a == "--probe") { probe = true; continue; }
    if (a == "--p6") { p6 = true; continue; }
    if (a == "-h" || a == "--help") { usage(); return 0; }
    positional.push_back(a);
  }
  for (auto& c : backend) c = (char)std::tolower((unsigned char)c);
  if (backend != "cpu" && backend != "gpu" && backend != "cuda") {
    std::fprintf(stderr, "--backend expects a value: cpu, gpu, or cuda\n");
    return 2;
  }
  if (backend == "cpu")
    std::fprintf(stderr, "note: this build has no CPU renderer; rendering on the B200 library (there is no fallback)\n");
  const std::string book_arg = positional.empty() ? "in_one_weekend" : positional[0];
  int scene_no = -1;
  if (positional.size() > 1) {
    char* end = nullptr;
    long v = std::strtol(positional[1].c_str(), &end, 10);
    if (end && *end == 0 && end != positional[1].c_str()) scene_no = (int)v;  // arg.parse::<i32>().ok()
  }
  const int book = rrth_normalize_book(book_arg.c_str());
  if (book == RRTH_BOOK_UNKNOWN) { usage(); return 1; }

  rrth_scene* hs = nullptr;
  if (rrth_scene_build(book, scene_no, scene_seed, &ov, nullptr, &hs) != 0) {
    std::fprintf(stderr, "scene build failed: %s\n", rrth_last_error());
    return 1;
  }
  const rrt_camera* cam = rrth_scene_camera(hs);
  const int W = cam->image_width, H = rrt_image_height(cam);
  // stdout is the image (render_io.rs:3-30 prints the PPM there).  Libraries under us may print too —
  // NCCL's version banner goes to stdout whenever NCCL_DEBUG is set in the environment — so everything
  // between here and the PPM is sent to stderr and the real stdout is put back for the image alone.
  std::fflush(stdout);
  const int saved_stdout = dup(1);
  dup2(2, 1);
  auto restore_stdout = [&]() { std::fflush(stdout); if (saved_stdout >= 0) dup2(saved_stdout, 1); };
  rrt_ctx* ctx = nullptr;
  const int init_rc = devices.empty() ? rrt_init(device, &ctx) : rrt_init_multi(devices.data(), (int)devices.size(), &ctx);
  if (init_rc != RRT_OK) {
    std::fprintf(stderr, "CUDA render failed: %s\n", rrt_last_error());  // main.rs:62-65 wording
    rrth_scene_free(hs);
    return 1;
  }
  int rc = 0;
  if (probe) {  // deterministic probe: one "id t" line per pixel
    rrt_scene* sc = nullptr;
    std::vector<int32_t> ids((size_t)W * H);
    std::vector<double> t((size_t)W * H);
    if (rrt_scene_create(ctx, rrth_scene_desc(hs), &sc) != RRT_OK || rrt_probe(ctx, sc, cam, 64, ids.data(), t.data()) != RRT_OK) {
      std::fprintf(stderr, "CUDA render failed: %s\n", rrt_last_error());
      rc = 1;
    } else {
      restore_stdout();
      FILE* f = out_path ? std::fopen(out_path, "w") : stdout;
      if (!f) { std::perror(out_path); rc = 1; }
      else {
        std::fprintf(f, "# probe %d %d\n", W, H);
        for (size_t i = 0; i < ids.size(); ++i) std::fprintf(f, "%d %.17g\n", ids[i], t[i]);
        if (out_path) std::fclose(f);
      }
    }
    rrt_scene_destroy(sc);
  } else {
    std::vector<float> accum((size_t)W * H * 4);
    rrt_render_params p;
    std::memset(&p, 0, sizeof p);
    p.integrator = rrth_scene_integrator(hs);
    p.shard_mode = devices.empty() ? RRT_SHARD_NONE : shard_mode; p.rank = 0; p.world = 1; p.seed = sample_seed;
    if (devices.empty()) std::fprintf(stderr, "Rendering %dx%d, %d spp, depth %d on device %d ...\n", W, H, cam->samples_per_pixel, cam->max_depth, device);
    else std::fprintf(stderr, "Rendering %dx%d, %d spp, depth %d on %d devices (%s) ...\n", W, H, cam->samples_per_pixel, cam->max_depth,
                      (int)devices.size(), shard_mode == RRT_SHARD_TILE ? "interleaved tiles" : "sample ranges");
    if (rrt_render_frame(ctx, rrth_scene_desc(hs), cam, &p, accum.data()) != RRT_OK) {
      std::fprintf(stderr, "CUDA render failed: %s\n", rrt_last_error());
      rc = 1;
    } else {
      rrt_stats st;
      rrt_get_stats(ctx, &st);
      std::fprintf(stderr, "Done: %.0f paths in %.1f ms (%.1f Mpaths/s), %d prims, %d BVH nodes, %d GPU(s), exchange %.2f ms\n", st.paths,
                   st.kernel_ms, st.paths / st.kernel_ms / 1e3, st.n_prims, st.n_bvh_nodes, st.n_devices, st.exchange_ms);
      std::vector<uint8_t> rgb;
      if (p6) {   // write_color's arithmetic (color.rs:19-36) on the device, three bytes a pixel
        rgb.resize((size_t)W * H * 3);
        if (rrt_resolve_rgb8(ctx, accum.data(), W, H, st.spp_effective, rgb.data()) != RRT_OK) {
          std::fprintf(stderr, "CUDA resolve failed: %s\n", rrt_last_error());
          rc = 1;
        }
      }
      restore_stdout();
      if (p6) {
        FILE* f = rc ? nullptr : (out_path ? std::fopen(out_path, "wb") : stdout);
        if (!f) { if (!rc) std::perror(out_path); rc = 1; }
        else {
          std::fprintf(f, "P6\n%d %d\n255\n", W, H);
          if (std::fwrite(rgb.data(), 1, rgb.size(), f) != rgb.size()) rc = 1;
          if (out_path) std::fclose(f); else std::fflush(f);
        }
      } else if (rrth_write_ppm_from_accum(out_path, W, H, accum.data(), (uint32_t)st.spp_effective) != 0) {
        std::fprintf(stderr, "could not write the image: %s\n", rrth_last_error());
        rc = 1;
      }
    }
  }
  rrt_shutdown(ctx);
  rrth_scene_free(hs);
  return rc;
}
