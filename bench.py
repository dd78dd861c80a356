#!/usr/bin/env python
"""bench.py — samples/s (Mpaths/s) of the path-tracing hot path on N B200s.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--config C2] [--impl b200|reference]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \
        --master-port P bench.py --gpus N --steps K --warmup W

A step = one full render of the workload frame (every pixel, every sample, full depth) followed by
the gamma resolve to RGB8, through the C ABI of librrt.so (include/rrt.h).

  value  device-resident: scene already uploaded, accum/RGB8 stay in HBM; CUDA events on the stream
         the kernels are launched on; max over ranks.
  e2e    the one-shot reference-facing call rrt_render_frame (twin of cuda::imp::render,
         src/cuda/mod.rs:342-441) with HOST buffers: host scene description in, flatten + BVH build
         + H2D + render + D2H of the RGBA-sum frame into pinned host memory, wall clock.

N = 1 renders BASELINE.json configs[1] (C2) as the headline and, after it, every other config at its
full size and spp (`configs`: C1, C3a, C3b, C3c, C4, C5).  N > 1 is STRONG scaling: the same C2 frame
(1200x675, 500 spp) split into N sample ranges, rendered and exchanged inside librrt.so
(rrt_render_collective: ncclReduce of the RGBA-sum frames to rank 0, inside the timed region), plus
`strong_scaling`: the two configs north_star shards — C4 (800x800, 10 000 spp) by sample ranges and C5
(3840x3840, 4096 spp) by interleaved tiles with a gather of owned tiles.  The N = 1 line carries C4 and
C5 in `configs`, so efficiency at N is strong_scaling[key].value / (N x configs[key].value at N = 1).

--impl reference times the CPU path (the ORACLE restatement of the reference's Rust renderer —
the Rust itself cannot be built in this image) on all host threads, on a bounded row sample of the
same workload.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
for _p in (ROOT, os.path.join(ROOT, "tests")):
    if _p not in sys.path:
        sys.path.insert(0, _p)

METRIC = "samples/sec (Mpaths/s)"
UNIT = "Mpaths/s"
SM_LANES = 128  # FP32 lanes per SM (4 SMSPs x 32)


def workload_name(key, n_gpus):
    from rustraytrace_b200 import configs
    return configs.DESCRIPTION[key]


def env_int(name, default):
    try:
        return int(os.environ.get(name, default))
    except ValueError:
        return default


class ClockSampler(threading.Thread):
    """nvidia-smi clocks/throttle reasons during the timed region (B200_PROFILING.md recipe)."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
    NAMES = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.rows, self.proc = index, [], None

    def run(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            for line in self.proc.stdout:
                self.rows.append([x.strip() for x in line.split(",")])
        except Exception:
            pass

    def stop(self):
        if self.proc is not None:
            self.proc.terminate()
        self.join(timeout=2)
        sm, mx, reasons, pw = [], [], set(), []
        for r in self.rows:
            try:
                sm.append(float(r[0])); mx.append(float(r[1])); pw.append(float(r[2]))
            except (ValueError, IndexError):
                continue
            for name, v in zip(self.NAMES, r[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2], "sm_max_mhz": max(mx), "reasons": sorted(reasons), "samples": len(sm),
                "power_w_max": max(pw)}


def measured_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            return json.load(f), "measured"
    return {"sm_max_mhz": 1965.0, "hbm_gbs": 6650.0}, "fallback"


def cpu_sample(key, seconds, threads):
    """The oracle on a bounded row sample of the workload (full width, full spp per sampled pixel).
    Returns (Mpaths/s, sample description, flops/path from the oracle's counters)."""
    import oracle_ffi as O
    from rustraytrace_b200 import configs
    s = configs.build(key)
    o = O.OracleScene(s.desc)
    H = s.height
    # calibrate on a thin slice at reduced spp, then size the real sample for ~`seconds`
    cal = configs.build(key, samples_per_pixel=4)
    t0 = time.perf_counter()
    _, c = O.OracleScene(cal.desc).render(cal.camera, cal.integrator, seed=3, nthreads=threads, rows=(0, H, max(1, H // 32)))
    rate = c["paths"] / max(1e-6, time.perf_counter() - t0)
    paths_per_row = s.width * (s.camera.samples_per_pixel if s.integrator != 3 else int(s.camera.samples_per_pixel ** 0.5) ** 2)
    n_rows = int(max(threads, min(H, rate * seconds / paths_per_row)))
    step = max(1, H // n_rows)
    t0 = time.perf_counter()
    _, c = o.render(s.camera, s.integrator, seed=1, nthreads=threads, rows=(0, H, step))
    dt = time.perf_counter() - t0
    rows = len(range(0, H, step))
    return (c["paths"] / dt / 1e6, f"{rows} of {H} rows (every {step}th) at full width and spp, {int(c['paths'])} paths, {dt:.1f} s",
            O.flops_per_path(c), dt, c["paths"])


def run_reference(args, rank, world):
    """--impl reference: the CPU path on the host cores (rank 0 only)."""
    if rank != 0:
        return 0
    from rustraytrace_b200 import configs
    threads = os.cpu_count() or 1
    vals, dts, sample = [], [], ""
    per_step = max(2.0, min(20.0, 90.0 / max(1, args.steps + args.warmup)))
    for i in range(args.warmup + args.steps):
        v, sample, _, dt, _ = cpu_sample(args.config, per_step, threads)
        if i >= args.warmup:
            vals.append(v); dts.append(dt)
    value = sum(vals) / len(vals)
    line = {"impl": "reference", "metric": METRIC, "value": round(value, 4), "unit": UNIT, "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": round(1e3 * sum(dts) / len(dts), 2),
            "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": workload_name(args.config, args.gpus), "key": args.config,
                       "note": "CPU restatement (oracle port) of the reference's Rust renderer; rustc/cargo absent. "
                               "The scene description it renders is built by librrt_host.so (host/books.cpp), the same "
                               "builder the GPU arm uses; the render itself runs in oracle/liboracle.so only"},
            "cpu_baseline": {"value": round(value, 4), "unit": UNIT, "cores": threads, "kind": "port", "sample": sample},
            "e2e": {"value": round(value, 4), "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)
    return 0


def measure(ctx, key, mode, rank, world, steps, warmup, stream, flush, e2e_steps=0, spp_override=0, sampler=None):
    """One config, strong scaling over `world` ranks: device-resident throughput (CUDA events on the
    launching stream, L2 flushed between steps) and, when e2e_steps > 0, the host-buffer end-to-end call.
    Collective: every rank calls it with the same arguments.  Returns a dict (complete on rank 0)."""
    import torch
    import torch.distributed as dist
    import rustraytrace_b200 as R
    from rustraytrace_b200 import configs
    F = R._ffi
    ov = {"samples_per_pixel": spp_override} if spp_override > 0 else {}
    scene = configs.build(key, **ov)
    W, H = scene.width, scene.height
    gs = ctx.scene(scene.desc)
    shard = {"sample": F.SHARD_SAMPLE, "tile": F.SHARD_TILE}[mode]
    accum = torch.zeros((H, W, 4), dtype=torch.float32, device="cuda") if rank == 0 else None
    rgb8 = torch.zeros((H, W, 3), dtype=torch.uint8, device="cuda") if rank == 0 else None

    def step():
        if world > 1:
            ctx.render_collective(gs, scene.camera, scene.integrator, seed=1, shard_mode=shard,
                                  accum_ptr=accum.data_ptr() if rank == 0 else None)
        else:
            ctx.render_device(gs, scene.camera, scene.integrator, accum.data_ptr(), seed=1)
        st = ctx.stats()
        if rank == 0:
            ctx.resolve_device(accum.data_ptr(), W, H, st["spp_effective"], rgb8.data_ptr())
        return st

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(warmup):
        flush.fill_(1)
        st = step()
    barrier()
    if sampler:
        sampler.start()
        time.sleep(0.3)
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
    kernel_ms = exchange_ms = my_paths = 0.0
    launches = 0
    barrier()
    t_wall0 = time.perf_counter()
    for a, b in ev:
        flush.fill_(1)          # L2 flush between timed iterations (outside the event pair)
        a.record(stream)
        st = step()
        b.record(stream)
        kernel_ms += st["kernel_ms"]; exchange_ms += st["exchange_ms"]
        launches += st["kernel_launches"] + (1 if rank == 0 else 0)   # render (+ tile scatter) + resolve on rank 0
        my_paths += st["paths"]
    barrier()
    wall_s = time.perf_counter() - t_wall0
    dev_ms = sum(a.elapsed_time(b) for a, b in ev)
    clocks = sampler.stop() if sampler else None
    tot = torch.tensor([dev_ms, kernel_ms], dtype=torch.float64, device="cuda")
    paths_t = torch.tensor([my_paths], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(tot, op=dist.ReduceOp.MAX)
        dist.all_reduce(paths_t, op=dist.ReduceOp.SUM)
    dev_ms_max, kernel_ms_max = tot.tolist()
    total_paths = paths_t.item()
    out = {"key": key, "workload": configs.DESCRIPTION[key], "width": W, "height": H, "spp": st["spp_effective"],
           "max_depth": scene.camera.max_depth, "steps": steps, "warmup": warmup,
           "value": round(total_paths / (dev_ms_max * 1e-3) / 1e6, 2), "unit": UNIT,
           "ms_per_step": round(dev_ms_max / steps, 3), "kernel_ms": round(kernel_ms_max / steps, 3),
           "exchange_ms_rank0": round(exchange_ms / steps, 3), "kernel_share_of_step": round(kernel_ms_max / dev_ms_max, 4),
           "paths_per_step": total_paths / steps, "my_paths_per_step": my_paths / steps, "gpu_launches": launches, "wall_s": round(wall_s, 3),
           "prims": st["n_prims"], "bvh_nodes": st["n_bvh_nodes"], "sm_count": st["sm_count"], "kernel_features": st["kernel_features"],
           "sharding": "none" if world == 1 else (("sample ranges + ncclReduce(SUM) to rank 0" if mode == "sample" else
                                                   "interleaved 32x32 tiles + gather of owned tiles to rank 0") + f", {world} ranks, inside librrt.so"),
           "clocks": clocks, "integrator": scene.integrator}
    f_path = configs.F_PATH.get(key)
    if f_path:
        peaks, _ = measured_peaks()
        peak = st["sm_count"] * SM_LANES * 2 * float(peaks.get("sm_max_mhz", 1965.0)) * 1e6 / 1e12
        out["flops_per_path"] = f_path
        # per GPU: this rank's paths through this rank's kernel
        out["roofline_frac"] = round((my_paths / steps) * f_path / (kernel_ms / steps * 1e-3) / 1e12 / peak, 4) if kernel_ms > 0 else None
    if rank == 0:
        a = accum.double().sum(dim=(0, 1)).tolist()
        out["frame_mean_radiance"] = [round(a[i] / a[3], 5) for i in range(3)]
        assert abs(a[3] - float(W) * H * st["spp_effective"]) < 0.5, "frame is missing samples"

    # ---- e2e: the host-buffer call, every rank its share, exchange inside the library, D2H on rank 0
    if e2e_steps > 0:
        host = torch.empty((H, W, 4), dtype=torch.float32, pin_memory=True) if rank == 0 else None
        host_np = host.numpy() if rank == 0 else None
        h2d = d2h = 0
        e2e_s = 0.0
        for i in range(1 + e2e_steps):
            barrier()
            t0 = time.perf_counter()
            if world == 1:
                ctx.render_frame(scene.desc, scene.camera, scene.integrator, seed=1, out=host_np)
                s2 = ctx.stats()
                h2d, d2h = s2["h2d_bytes"], s2["d2h_bytes"]
            else:
                # flatten + H2D on every rank, share, exchange, D2H on rank 0, free: the collective twin of rrt_render_frame
                ctx.render_frame_collective(scene.desc, scene.camera, scene.integrator, seed=1, shard_mode=shard, out=host_np)
                s2 = ctx.stats()
                h2d, d2h = s2["h2d_bytes"], s2["d2h_bytes"]
            barrier()
            if i > 0:
                e2e_s += time.perf_counter() - t0
        t = torch.tensor([e2e_s], dtype=torch.float64, device="cuda")
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        out["e2e"] = {"value": round(total_paths / steps * e2e_steps / t.item() / 1e6, 2), "unit": UNIT, "h2d_bytes_per_step": int(h2d),
                      "d2h_bytes_per_step": int(d2h), "steps": e2e_steps, "ms_per_step": round(1e3 * t.item() / e2e_steps, 3),
                      "call": "rrt_render_frame (host scene description in, pinned RGBA-sum frame out)" if world == 1 else
                              "rrt_render_frame_collective (host scene description in on every rank, share + NCCL exchange in librrt.so, pinned frame out on rank 0)"}
    gs.close()
    return out


# the other BASELINE.json configs in the driver's line: (key, timed steps, warm-up steps); every one at its full size and spp
OTHER_CONFIGS = [("C1", 3, 2), ("C3a", 3, 1), ("C3b", 3, 1), ("C3c", 3, 1), ("C4", 2, 1), ("C5", 2, 1)]
STRONG_CONFIGS = [("C4", 2, 1), ("C5", 2, 1)]   # the configs north_star shards: sample ranges / tiles


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--config", default="C2")
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--spp", type=int, default=0, help="override samples per pixel (diagnostics only; marks the line)")
    ap.add_argument("--cpu-seconds", type=float, default=15.0)
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg (profiling runs)")
    ap.add_argument("--no-e2e", action="store_true", help="skip the e2e leg (profiling runs)")
    ap.add_argument("--no-configs", action="store_true", help="skip the other configs (profiling runs)")
    args = ap.parse_args()
    rank, local_rank, world = env_int("RANK", 0), env_int("LOCAL_RANK", 0), env_int("WORLD_SIZE", 1)
    if args.impl != "reference" and args.gpus != world:
        print(f"bench.py: --gpus {args.gpus} but WORLD_SIZE={world}; launch with torch.distributed.run for N>1 (using {world})", file=sys.stderr)
    if args.impl == "reference":
        return run_reference(args, rank, world)
    args.warmup = max(3, args.warmup) if args.spp == 0 else args.warmup

    import torch
    import torch.distributed as dist
    import rustraytrace_b200 as R
    from rustraytrace_b200 import configs, distributed
    F = R._ffi

    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the render path has no CPU fallback (use --impl reference for the CPU arm)")
    torch.cuda.set_device(local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    n_gpus = world

    ctx = R.Context(local_rank)
    stream = torch.cuda.Stream()
    torch.cuda.set_stream(stream)
    ctx.set_stream(stream.cuda_stream)
    if world > 1:
        distributed.init_comm_from_torch(ctx)   # torch carries the NCCL id; the exchange itself is librrt's
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")  # > 126 MB L2
    diag = args.spp > 0 or args.no_configs

    mode = configs.CONFIGS[args.config][3]
    sampler = ClockSampler(local_rank) if rank == 0 else None
    head = measure(ctx, args.config, mode, rank, world, args.steps, args.warmup, stream, flush,
                   e2e_steps=0 if args.no_e2e else args.steps, spp_override=args.spp, sampler=sampler)
    others, strong = {}, {}
    if not diag:
        if n_gpus == 1:
            for key, k, w in OTHER_CONFIGS:
                if key != args.config:
                    others[key] = measure(ctx, key, configs.CONFIGS[key][3], rank, world, k, w, stream, flush, e2e_steps=1 if key in ("C4", "C5") else 0)
        else:
            for key, k, w in STRONG_CONFIGS:
                strong[key] = measure(ctx, key, configs.CONFIGS[key][3], rank, world, k, w, stream, flush, e2e_steps=1)

    if rank == 0:
        peaks, peak_src = measured_peaks()
        sm_count = head["sm_count"]
        W, H = head["width"], head["height"]
        f_path = configs.F_PATH.get(args.config)
        cpu = None
        if n_gpus == 1 and not args.no_cpu:
            v, sample, f_live, _, _ = cpu_sample(args.config, args.cpu_seconds, os.cpu_count() or 1)
            cpu = {"value": round(v, 4), "unit": UNIT, "cores": os.cpu_count() or 1, "kind": "port", "sample": sample,
                   "flops_per_path_live": round(f_live, 1)}
        ref_cuda = None
        if n_gpus == 1 and not args.no_cpu and head["integrator"] == F.INT_BOOK1:
            # the reference's existing src/cuda path, same scene, same box, its own launch shape
            # (oracle/_ref/libref_cuda.so, tests/ref_cuda_ffi.py); it supports book-1 scenes only
            try:
                import ref_cuda_ffi as RC
                if RC.available():
                    rs = configs.build(args.config, samples_per_pixel=min(128, head["spp"]))
                    RC.render(configs.build(args.config, samples_per_pixel=4))   # warm-up
                    racc, rms, rpasses = RC.render(rs)
                    rpaths = float(W) * H * rs.camera.samples_per_pixel
                    rm = racc[..., :3].sum(axis=(0, 1)) / racc[..., 3].sum()
                    ref_cuda = {"value": round(rpaths / rms / 1e3, 2), "unit": UNIT, "kernel_ms": round(rms, 3),
                                "sample": f"full frame at {rs.camera.samples_per_pixel} spp, {rpasses} pass(es) of <=256 spp, 8x8 blocks, sync per pass",
                                "what": "reference src/cuda/mod.rs kernel (brute-force spheres, f32, xorshift32), compiled for sm_100a",
                                "frame_mean_radiance": [round(float(x), 5) for x in rm]}
            except Exception as e:  # the comparison arm must never take the bench down
                ref_cuda = {"unavailable": repr(e)}
        sm_max = float(peaks.get("sm_max_mhz", 1965.0))
        peak_tf = sm_count * SM_LANES * 2 * sm_max * 1e6 / 1e12
        # dominant kernel = render_kernel: algorithmic flops per launch / its mean launch duration (rank 0's launch)
        k_ms = head["kernel_ms"]
        achieved = (head["my_paths_per_step"] * f_path / (k_ms * 1e-3) / 1e12) if f_path else None
        traffic, traffic_src = None, None
        import glob
        # ncu_r<round><letter>_render_kernel_c2_spp64.csv only: captures of experiments carry a tag (ncu_r02h_pool_..., _bvh4_...)
        caps = sorted(p for p in glob.glob(os.path.join(ROOT, "profiles", "ncu_r*_render_kernel_c2_spp64.csv"))
                      if os.path.basename(p).split("_")[2] == "render")
        ncu_csv = caps[-1] if caps else ""   # the newest capture (names sort by round)
        if args.config == "C2" and ncu_csv:   # the committed `ncu --set full` capture of this kernel
            tot = 0.0
            for ln in open(ncu_csv):
                k, u, v = (ln.strip().split(",") + ["", ""])[:3]
                if k in ("dram__bytes_read.sum", "dram__bytes_write.sum"):
                    tot += float(v) * {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}.get(u, 1)
            traffic, traffic_src = int(tot), "profiles/" + os.path.basename(ncu_csv) + " (64-spp launch; DRAM traffic is the frame, independent of spp)"
        clocks = head["clocks"]
        roof = {"bound": "fp32", "achieved": round(achieved, 3) if achieved else None, "peak": round(peak_tf, 2),
                "unit": "TFLOP/s", "frac": round(achieved / peak_tf, 4) if achieved else None, "traffic": traffic, "traffic_source": traffic_src,
                "algorithmic_dram_bytes": W * H * 16,
                "kernel": "render_kernel", "kernel_ms": round(k_ms, 3), "flops_per_path": f_path,
                "peak_source": f"{sm_count} SMs x 128 lanes x 2 x sm_max_mhz {sm_max:.0f} ({peak_src} MEASURED_PEAKS.json); "
                               "path tracing is FP32-pipe/latency bound, neither HBM nor tensor",
                "kernel_share_of_step": head["kernel_share_of_step"]}
        if clocks and clocks.get("sm_mhz") and achieved:
            roof["frac_at_observed_clock"] = round(achieved / (sm_count * SM_LANES * 2 * clocks["sm_mhz"] * 1e6 / 1e12), 4)

        def brief(d):
            keep = ("workload", "width", "height", "spp", "max_depth", "steps", "warmup", "value", "unit", "ms_per_step", "kernel_ms",
                    "exchange_ms_rank0", "flops_per_path", "roofline_frac", "sharding", "frame_mean_radiance", "e2e", "kernel_features")
            return {k: d[k] for k in keep if k in d}

        line = {"metric": METRIC, "value": head["value"], "unit": UNIT, "n_gpus": n_gpus, "steps": args.steps,
                "warmup": args.warmup, "ms_per_step": head["ms_per_step"], "higher_is_better": True,
                "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
                "config": {"workload": head["workload"] + (f", one frame over {n_gpus} GPUs (strong scaling)" if n_gpus > 1 else ""),
                           "key": args.config, "width": W, "height": H, "spp": head["spp"],
                           "max_depth": head["max_depth"], "scene_seed": hex(configs.SCENE_SEED),
                           "sharding": head["sharding"],
                           "l2": "flushed between steps (256 MiB fill)", "prims": head["prims"], "bvh_nodes": head["bvh_nodes"]},
                "e2e": head.get("e2e"), "gpu_launches": head["gpu_launches"], "roofline": roof, "cpu_baseline": cpu, "ref_cuda_kernel": ref_cuda,
                "clocks": clocks, "exchange_ms_rank0": head["exchange_ms_rank0"],
                "wall_s": head["wall_s"], "frame_mean_radiance": head.get("frame_mean_radiance")}
        if others:
            line["configs"] = {k: brief(v) for k, v in others.items()}
        if strong:
            line["strong_scaling"] = {k: brief(v) for k, v in strong.items()}
        if args.spp > 0:
            line["config"]["diagnostic_spp_override"] = args.spp
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    ctx.close()
    return 0


if __name__ == "__main__":
    sys.exit(main())
