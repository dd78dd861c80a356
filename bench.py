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

N = 1 renders BASELINE.json configs[1] (C2).  N > 1 is weak scaling on the sample axis: the frame
keeps its size, samples-per-pixel become 500*N, rank r renders samples [500r, 500(r+1)) of every
pixel (RRT_SHARD_SAMPLE) and the RGBA-sum frames are SUM-reduced to rank 0 over NCCL inside the
timed region (C5 shards by interleaved tiles instead).

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
    return configs.DESCRIPTION[key] + (f", spp x{n_gpus} (weak scaling on the sample axis)" if n_gpus > 1 else "")


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
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": workload_name(args.config, args.gpus), "key": args.config,
                       "note": "CPU restatement (oracle port) of the reference's Rust renderer; rustc/cargo absent"},
            "cpu_baseline": {"value": round(value, 4), "unit": UNIT, "cores": threads, "kind": "port", "sample": sample},
            "e2e": {"value": round(value, 4), "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)
    return 0


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
    args = ap.parse_args()
    rank, local_rank, world = env_int("RANK", 0), env_int("LOCAL_RANK", 0), env_int("WORLD_SIZE", 1)
    if args.impl != "reference" and args.gpus != world:
        print(f"bench.py: --gpus {args.gpus} but WORLD_SIZE={world}; launch with torch.distributed.run for N>1 (using {world})", file=sys.stderr)
    if args.impl == "reference":
        return run_reference(args, rank, world)
    args.warmup = max(3, args.warmup) if args.spp == 0 else args.warmup

    import numpy as np
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

    book, scene_no, ov, mode = configs.CONFIGS[args.config]
    base_spp = args.spp if args.spp > 0 else ov["samples_per_pixel"]
    # weak scaling on the sample axis (tile mode: on the frame's spp as well, shares are tiles)
    spp = base_spp * n_gpus
    scene = configs.build(args.config, samples_per_pixel=spp)
    W, H = scene.width, scene.height

    ctx = R.Context(local_rank)
    stream = torch.cuda.Stream()
    torch.cuda.set_stream(stream)
    ctx.set_stream(stream.cuda_stream)
    gs = ctx.scene(scene.desc)
    accum = torch.zeros((H, W, 4), dtype=torch.float32, device="cuda")
    rgb8 = torch.zeros((H, W, 3), dtype=torch.uint8, device="cuda")
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")  # > 126 MB L2

    def step():
        distributed.render_sharded(ctx, gs, scene.camera, scene.integrator, accum, seed=1, mode=mode, rank=rank, world=n_gpus)
        st = ctx.stats()
        if rank == 0:
            ctx.resolve_device(accum.data_ptr(), W, H, st["spp_effective"], rgb8.data_ptr())
        return st

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(args.warmup):
        flush.fill_(1)
        st = step()
    barrier()

    sampler = ClockSampler(local_rank) if rank == 0 else None
    if sampler:
        sampler.start()
        time.sleep(0.3)
    # ---- device-resident timed region: K steps, each bracketed by events on the launching stream
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
    kernel_ms, launches, my_paths = 0.0, 0, 0.0
    barrier()
    t_wall0 = time.perf_counter()
    for a, b in ev:
        flush.fill_(1)          # L2 flush between timed iterations (outside the event pair)
        a.record(stream)
        st = step()
        b.record(stream)
        kernel_ms += st["kernel_ms"]
        launches += 1 + (1 if rank == 0 else 0)  # render kernel (+ resolve kernel on rank 0)
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
    value = total_paths / (dev_ms_max * 1e-3) / 1e6
    frame_mean = None
    if rank == 0:
        a = accum.double().sum(dim=(0, 1)).tolist()
        frame_mean = [round(a[i] / a[3], 5) for i in range(3)]
        assert abs(a[3] - float(W) * H * st["spp_effective"]) < 0.5, "frame is missing samples"

    # ---- e2e: the one-shot host-buffer call, every rank its share, reduce on device, D2H on rank 0
    e2e = None
    if not args.no_e2e:
        host = torch.empty((H, W, 4), dtype=torch.float32, pin_memory=True)
        host_np = host.numpy()
        n_e2e = max(1, min(args.steps, 3))
        h2d = d2h = 0
        e2e_s = 0.0
        for i in range(1 + n_e2e):
            barrier()
            t0 = time.perf_counter()
            if world == 1:
                ctx.render_frame(scene.desc, scene.camera, scene.integrator, seed=1, out=host_np)
                s2 = ctx.stats()
                h2d, d2h = s2["h2d_bytes"], s2["d2h_bytes"]
            else:
                g2 = ctx.scene(scene.desc)          # flatten + BVH build + H2D
                h2d = ctx.stats()["h2d_bytes"]
                distributed.render_sharded(ctx, g2, scene.camera, scene.integrator, accum, seed=1, mode=mode, rank=rank, world=n_gpus)
                if rank == 0:
                    host.copy_(accum, non_blocking=True)
                    d2h = host.numel() * 4
                g2.close()
            barrier()
            if i > 0:
                e2e_s += time.perf_counter() - t0
        t = torch.tensor([e2e_s], dtype=torch.float64, device="cuda")
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        e2e_paths = total_paths / args.steps * n_e2e
        e2e = {"value": round(e2e_paths / t.item() / 1e6, 2), "unit": UNIT, "h2d_bytes_per_step": int(h2d),
               "d2h_bytes_per_step": int(d2h), "steps": n_e2e, "ms_per_step": round(1e3 * t.item() / n_e2e, 3),
               "call": "rrt_render_frame (host scene description in, pinned RGBA-sum frame out)" if world == 1 else
                       "rrt_scene_create + rrt_render_device + NCCL reduce + D2H on rank 0"}

    if rank == 0:
        peaks, peak_src = measured_peaks()
        sm_count = st["sm_count"]
        f_path = configs.F_PATH.get(args.config)
        cpu = None
        if n_gpus == 1 and not args.no_cpu:
            v, sample, f_live, _, _ = cpu_sample(args.config, args.cpu_seconds, os.cpu_count() or 1)
            cpu = {"value": round(v, 4), "unit": UNIT, "cores": os.cpu_count() or 1, "kind": "port", "sample": sample,
                   "flops_per_path_live": round(f_live, 1)}
        ref_cuda = None
        if n_gpus == 1 and not args.no_cpu and scene.integrator == F.INT_BOOK1:
            # the reference's existing src/cuda path, same scene, same box, its own launch shape
            # (oracle/_ref/libref_cuda.so, tests/ref_cuda_ffi.py); it supports book-1 scenes only
            try:
                import ref_cuda_ffi as RC
                if RC.available():
                    rs = configs.build(args.config, samples_per_pixel=min(128, st["spp_effective"]))
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
        # dominant kernel = render_kernel: algorithmic flops per launch / its mean launch duration
        paths_per_launch = my_paths / args.steps
        k_ms = kernel_ms / args.steps
        achieved = (paths_per_launch * f_path / (k_ms * 1e-3) / 1e12) if f_path else None
        traffic, traffic_src = None, None
        import glob
        caps = sorted(glob.glob(os.path.join(ROOT, "profiles", "ncu_r*_render_kernel_c2_spp64.csv")))
        ncu_csv = caps[-1] if caps else ""   # the newest capture (names sort by round)
        if args.config == "C2" and ncu_csv:   # the committed `ncu --set full` capture of this kernel
            tot = 0.0
            for ln in open(ncu_csv):
                k, u, v = (ln.strip().split(",") + ["", ""])[:3]
                if k in ("dram__bytes_read.sum", "dram__bytes_write.sum"):
                    tot += float(v) * {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}.get(u, 1)
            traffic, traffic_src = int(tot), "profiles/" + os.path.basename(ncu_csv) + " (64-spp launch; DRAM traffic is the frame, independent of spp)"
        roof = {"bound": "fp32", "achieved": round(achieved, 3) if achieved else None, "peak": round(peak_tf, 2),
                "unit": "TFLOP/s", "frac": round(achieved / peak_tf, 4) if achieved else None, "traffic": traffic, "traffic_source": traffic_src,
                "algorithmic_dram_bytes": W * H * 16,
                "kernel": "render_kernel", "kernel_ms": round(k_ms, 3), "flops_per_path": f_path,
                "peak_source": f"{sm_count} SMs x 128 lanes x 2 x sm_max_mhz {sm_max:.0f} ({peak_src} MEASURED_PEAKS.json); "
                               "path tracing is FP32-pipe/latency bound, neither HBM nor tensor",
                "kernel_share_of_step": round(kernel_ms_max / dev_ms_max, 4)}
        if clocks and clocks.get("sm_mhz") and achieved:
            roof["frac_at_observed_clock"] = round(achieved / (sm_count * SM_LANES * 2 * clocks["sm_mhz"] * 1e6 / 1e12), 4)
        line = {"metric": METRIC, "value": round(value, 2), "unit": UNIT, "n_gpus": n_gpus, "steps": args.steps,
                "warmup": args.warmup, "ms_per_step": round(dev_ms_max / args.steps, 3), "higher_is_better": True,
                "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
                "config": {"workload": workload_name(args.config, n_gpus),
                           "key": args.config, "width": W, "height": H, "spp": st["spp_effective"],
                           "max_depth": scene.camera.max_depth, "scene_seed": hex(configs.SCENE_SEED),
                           "sharding": "none" if n_gpus == 1 else ("sample ranges" if mode == "sample" else "interleaved 32x32 tiles") + f" over {n_gpus} ranks + NCCL sum-reduce to rank 0",
                           "l2": "flushed between steps (256 MiB fill)", "prims": st["n_prims"], "bvh_nodes": st["n_bvh_nodes"]},
                "e2e": e2e, "gpu_launches": launches, "roofline": roof, "cpu_baseline": cpu, "ref_cuda_kernel": ref_cuda, "clocks": clocks,
                "wall_s": round(wall_s, 3), "frame_mean_radiance": frame_mean}
        if args.spp > 0:
            line["config"]["diagnostic_spp_override"] = args.spp
        print(json.dumps(line), flush=True)
    gs.close()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    ctx.close()
    return 0


if __name__ == "__main__":
    sys.exit(main())
