#!/usr/bin/env python3
"""bench.py -- the headline measurement (BASELINE.json): closest-hit Mrays/s of the path-tracing hot path on B200, with the
roofline of the dominant kernel, hit-ID match against the reference's own intersection code, and the reference's CPU
path timed beside it.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference] [--reduce nccl|peer] [--skip c2,c3,c5,cpu]

N = 1  -- workload = BASELINE.json configs[3] (C4, the configuration the north-star target is quoted on; it fits one GPU):
          synthetic 1M-triangle random soup, one batch of 3840x2160 = 8 294 400 uniformly incoherent rays ("batch A").
          A step = one closest-hit pass over the whole batch.  `value` has rays and hits resident in HBM
          (ff_trace_closest_device); `e2e` goes through the host call a user makes (ff_trace_closest: page-locked host
          rays in, hit ids + t out, host<->device copies inside the timed region).  Beside it, untimed or separately timed:
          hit ids / t of the first rays of the batch against oracle/_ref (the reference's intersectRays compiled from its
          own source), the diffuse-bounce batch B, BVH build times, and secondary blocks for the other configs --
          c2 (configs[1], the frame round 1 reported), c3 (configs[2], wavefront vs megakernel) and c5 at one GPU (the
          base of the strong-scaling series below), c2 / c3 each with an oracle crop check of the benchmarked frame.
N > 1  -- (torchrun, one rank per GPU) workload = BASELINE.json configs[4] (C5): 20M-triangle soup (binned-SAH tree), 7680x4320, a FIXED
          total of 64 spp split over the ranks as disjoint sample ranges (strong scaling), the 531 MB accumulators
          combined by ONE framebuffer reduce per frame inside the library (ff_group_render: NCCL reduce, or the fused
          reduce-scatter + resolve kernel over NVLink peer memory with --reduce peer).  `value` = closest-hit rays of all
          ranks / frame time (max over ranks).  Rank 0 also renders the whole frame alone (`scaling_base`), so each line
          carries its own one-GPU base; the N = 1 line's `c5` block is the same measurement in a one-GPU job.
One JSON line on stdout (rank 0); everything else goes to stderr.
"""
import argparse
import json
import os
import subprocess
import sys
import tempfile
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC, UNIT = "closest_hit_rays_per_second", "Mrays/s"
C4_TRIS, C4_W, C4_H = 1000000, 3840, 2160
C4_RAYS = C4_W * C4_H
C5_TRIS, C5_W, C5_H, C5_SPP, C5_BOUNCES = 20000000, 7680, 4320, 64, 4
C2 = dict(w=1024, h=1024, spp=64, bounces=4)
C3 = dict(w=1920, h=1080, spp=256, bounces=8)
REF_RAYS_PER_STEP = {1: 1024, 2: 64}   # constant ray count of one reference-arm step: C4 (1M triangles per ray), C5 (20M)


def bench_config(world):
    """`config` of the JSON line -- identical in both arms, so the driver's same_config comparison holds."""
    if world == 1:
        return dict(workload="C4: synthetic 1M-triangle random soup, batch A = %d uniformly incoherent rays (one per pixel of %dx%d), closest hit, "
                             "bit-exact hit ids vs the reference's intersectRays" % (C4_RAYS, C4_W, C4_H),
                    triangles=C4_TRIS, rays_per_step=C4_RAYS, l2="flushed with a 256 MiB write between timed steps")
    return dict(workload="C5: synthetic 20M-triangle soup + one emissive quad, %dx%d, %d spp in total split over the GPUs as disjoint sample ranges "
                         "(strong scaling), %d bounces diffuse, one framebuffer reduce per frame" % (C5_W, C5_H, C5_SPP, C5_BOUNCES),
                triangles=C5_TRIS + 2, image=[C5_W, C5_H], total_spp=C5_SPP, max_bounces=C5_BOUNCES,
                l2="working set (path state of a pass, 9.6 GB) is far larger than L2")


_REAL_STDOUT = None


def claim_stdout():
    """The contract is ONE JSON line on stdout.  Libraries write there too (NCCL prints its version banner to fd 1), so
    fd 1 is pointed at stderr for the whole run and the JSON line goes to a private duplicate of the original stdout."""
    global _REAL_STDOUT
    if _REAL_STDOUT is None:
        sys.stdout.flush()
        _REAL_STDOUT = os.fdopen(os.dup(1), "w")
        os.dup2(2, 1)


def emit(line):
    out = _REAL_STDOUT if _REAL_STDOUT is not None else sys.stdout
    out.write(json.dumps(line) + "\n")
    out.flush()


def log(*a):
    print(*a, file=sys.stderr, flush=True)


def measured_peak():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md)"


def profile_numbers():
    """Per-launch figures that only a profiler can give (ncu, committed under profiles/): DRAM bytes, issue-slot use."""
    p = os.path.join(ROOT, "profiles", "traffic.json")
    try:
        return json.load(open(p))
    except Exception:
        return {}


class ClockSampler:
    """nvidia-smi clocks/throttle reasons DURING the timed region (B200_PROFILING.md recipe)."""
    Q = ("timestamp,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.proc, self.path = index, None, None
        self.t0 = self.t1 = None

    def mark_begin(self):
        """The timed region starts here.  nvidia-smi itself is started BEFORE the warm-up: its start-up (NVML initialisation,
        a few hundred ms of driver calls) used to land on the first timed steps and cost one of them 4.5 ms
        (profiles/r01m_bench_c2_after_reference_arm.json); rows are filtered to the timed region by their timestamps."""
        self.t0 = time.time()

    def mark_end(self):
        self.t1 = time.time()

    def start(self):
        try:
            f = tempfile.NamedTemporaryFile("w", suffix=".csv", delete=False)
            self.path = f.name
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q, "--format=csv,noheader,nounits",
                                          "-lms", "100"], stdout=f, stderr=subprocess.DEVNULL)
        except Exception as e:  # no nvidia-smi: report it, do not fail the run
            log("clock sampling unavailable:", e)
            self.proc = None

    def stop(self):
        out = dict(sm_mhz=None, sm_max_mhz=None, reasons=[], samples=0)
        if self.proc is None:
            return out
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        try:
            rows = [r.split(",") for r in open(self.path).read().strip().splitlines() if r.count(",") >= 8]
            os.unlink(self.path)
            if self.t0 is not None and self.t1 is not None:
                import datetime

                def stamp(r):
                    try:
                        return datetime.datetime.strptime(r[0].strip(), "%Y/%m/%d %H:%M:%S.%f").timestamp()
                    except Exception:
                        return None
                inside = [r for r in rows if stamp(r) is not None and self.t0 - 0.05 <= stamp(r) <= self.t1 + 0.05]
                if inside:
                    rows = inside
            sm = [float(r[1]) for r in rows]
            reasons = set()
            for r in rows:
                for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[5:9]):
                    if v.strip().lower().startswith("active"):
                        reasons.add(name)
            if sm:
                out = dict(sm_mhz=float(np.median(sm)), sm_max_mhz=float(rows[0][2]), reasons=sorted(reasons), samples=len(sm),
                           power_w_max=max(float(r[3]) for r in rows))
        except Exception as e:
            log("clock parse failed:", e)
        return out


# --------------------------------------------------------------------------------------------- CPU arms
class CpuArm:
    """The reference's own closest-hit routine (intersectRays + intersectTriangle, kernel.cu:35-176: brute force over every
    triangle for every ray) on the host cores, on a bounded sample of the workload's rays.  Uses oracle/_ref (the
    reference's text compiled from /root/reference) when present, else the oracle restatement.  Nothing here touches the
    GPU library: the soup and the rays come from the oracle's own host generators (the same Philox streams)."""

    def __init__(self, n_tris):
        from oracle import oracle as O
        if not O.Oracle.available():
            O.build()
        self.gen = O.Oracle()
        if O.Ref.available():
            self.lib, self.kind = O.Ref(), "reference"
        else:
            self.lib, self.kind = self.gen, "port"
        self.n_tris = n_tris
        self.scene = self.lib.scene()
        self.scene.add_mesh(self.gen.make_soup(n_tris))
        self.cores = int(self.lib.threads)

    def rays(self, first, n):
        return self.gen.make_rays_uniform(first, n)

    def intersect(self, rays):
        """(hits dict, seconds): the reference's object-space path (mode 0)."""
        t0 = time.perf_counter()
        r = self.scene.intersect(rays, threads=self.cores) if self.kind == "reference" else self.scene.intersect(rays, threads=self.cores, mode=0)
        return r, time.perf_counter() - t0

    def describe(self, n):
        return ("%d rays of batch A (uniform origins in the unit cube, uniform directions), each tested against all %d triangles by the "
                "reference's intersectRays; %d host threads" % (n, self.n_tris, self.cores))


def run_reference_arm(args):
    if int(os.environ.get("RANK", "0")) != 0:
        return 0
    world = args.gpus
    n_tris = C4_TRIS if world == 1 else C5_TRIS
    per_step = REF_RAYS_PER_STEP[1 if world == 1 else 2]
    arm = CpuArm(n_tris)
    for w in range(args.warmup):
        arm.intersect(arm.rays(1 << 30, max(16, per_step // 8)))
    rays, secs = 0, 0.0
    for k in range(args.steps):
        _, t = arm.intersect(arm.rays(k * per_step, per_step))  # step k = rays [k * per_step, (k + 1) * per_step) of batch A
        rays += per_step
        secs += t
    value = rays / secs / 1e6 if secs > 0 else 0.0
    what = arm.describe(per_step) + " per step"
    line = dict(metric=METRIC, value=value, unit=UNIT, impl="reference", n_gpus=args.gpus, steps=args.steps, warmup=args.warmup,
                ms_per_step=1e3 * secs / max(1, args.steps), higher_is_better=True, scaling="weak" if world == 1 else "strong", vs_baseline=None,
                dtype="f32", data="synthetic", config=bench_config(world),
                cpu_baseline=dict(value=value, unit=UNIT, cores=arm.cores, kind=arm.kind, sample=what),
                e2e=dict(value=value, unit=UNIT, h2d_bytes_per_step=0, d2h_bytes_per_step=0), gpu_launches=0)
    emit(line)
    return 0


# --------------------------------------------------------------------------------------------- GPU arm, one GPU
def philox_tile(ff, w, h, tw, th, salt):
    """A tile position drawn from Philox (reproducible, not hand-picked)."""
    r = ff.philox4x32_10([salt, w, h, 0], [0xF1EEF1E5, 7])
    return int(r[0]) % (w - tw + 1), int(r[1]) % (h - th + 1)


def frame_block(ff, S, ctx, meshes, name, desc, dims, tile, steps, skip_cpu, integrators=("wavefront",)):
    """One of the frame configs (C2 / C3): timed frames per integrator + an oracle crop check of the SAME full-size frame."""
    sc, cam = S.instantiate(desc, ctx, meshes, builder=1)
    info = sc.info()
    w, h, spp, bounces = dims["w"], dims["h"], dims["spp"], dims["bounces"]
    npix = w * h
    d_out = ctx.alloc(npix * 16)
    out = dict(triangles=info["n_triangles"], image=[w, h], spp=spp, max_bounces=bounces, builder="binned-sah", build_ms=info["build_ms"])
    frames = {}
    for nm in integrators:
        integ = ff.INTEGRATOR_WAVEFRONT if nm == "wavefront" else ff.INTEGRATOR_MEGAKERNEL
        sc.render(cam, min(spp, 8), bounces, host=False, device_out=d_out, integrator=integ)
        sc.render(cam, spp, bounces, host=False, device_out=d_out, integrator=integ)
        secs, trace_ms, rays, launches = 0.0, 0.0, 0, 0
        for _ in range(steps):
            ctx.flush_l2()
            ctx.synchronize()
            ctx.timer_begin()
            _, st = sc.render(cam, spp, bounces, want_stats=True, host=False, device_out=d_out, integrator=integ)
            secs += 1e-3 * ctx.timer_end()
            trace_ms += st["trace_ms"]
            rays += st["rays"]
            launches += st["kernel_launches"]
        b = dict(ms_per_frame=1e3 * secs / steps, mrays_s=rays / secs / 1e6, samples_per_sec_per_gpu=spp * npix * steps / secs,
                 rays_per_frame=rays // steps, launches_per_frame=launches // steps, passes=st["passes"])
        if nm == "wavefront":
            b["trace_ms"] = trace_ms / steps
            _, stc = sc.render(cam, spp, bounces, want_stats=True, host=False, device_out=d_out, integrator=integ | ff.TRACE_COUNT)
            b["nodes_per_ray"], b["tris_per_ray"] = stc["nodes_visited"] / stc["rays"], stc["tris_tested"] / stc["rays"]
            # the user-facing call with a HOST framebuffer
            sc.render(cam, spp, bounces, pinned=True)
            ctx.flush_l2()
            ctx.synchronize()
            ctx.timer_begin()
            frames[nm], _ = sc.render(cam, spp, bounces, want_stats=True, pinned=True)
            b["e2e_ms_per_frame"] = ctx.timer_end()
            frames[nm] = np.array(frames[nm])
        else:
            frames[nm] = sc.render(cam, spp, bounces, integrator=integ)
        out[nm] = b
    if not skip_cpu:
        # untimed parity of the benchmarked frame: a Philox-chosen tile of the full-size, full-spp frame, rendered by the CPU
        # restatement of the same algorithm (oracle/, as-intended tracer) and compared bit for bit with the GPU's pixels
        from oracle import oracle as O
        orc = O.Oracle()
        osc = S.instantiate_on(desc, orc.scene(), meshes)
        x0, y0 = philox_tile(ff, w, h, tile, tile, 2 if name == "c2" else 3)
        t0 = time.perf_counter()
        want, _ = osc.render_crop(desc["camera"], spp, bounces, x0, y0, tile, tile)
        par = dict(tile=[x0, y0, tile, tile], pixels=tile * tile, samples=tile * tile * spp, oracle_seconds=time.perf_counter() - t0)
        for nm, img in frames.items():
            got = img[y0:y0 + tile, x0:x0 + tile]
            par["mismatched_bits_" + nm] = int((got.view(np.uint32) != want.view(np.uint32)).sum())
        out["parity"] = par
    d_out.free()
    sc.close()
    return out


C5_BUILDER = 1   # binned SAH: 31 ms instead of 5.9 ms of (replicated, untimed) build for a frame that renders 3 % faster (profiles/r04i_c5_builders.jsonl)
BUILDER_NAMES = {0: "lbvh", 1: "binned-sah"}


def c5_scene(ff, S, ctx, n_tris, builder=C5_BUILDER):
    soup = ctx.make_soup(n_tris)
    desc = S.config5(soup, C5_W, C5_H, C5_SPP, C5_BOUNCES)
    sc, cam = S.instantiate(desc, ctx, builder=builder)
    sc.commit(builder)  # second build: scratch arena warm, the steady-state figure
    return sc, cam


def run_gpu_one(args):
    import gpupathtracer_b200 as ff
    from gpupathtracer_b200 import scenes as S
    if ff.device_count() == 0:
        raise SystemExit("bench.py: no CUDA device; the product path has no CPU fallback")
    skip = set(x for x in args.skip.split(",") if x)
    ctx = ff.Context(0)
    sampler = ClockSampler(0)
    sampler.start()
    warm = max(3, args.warmup)
    n = C4_RAYS

    # ---------------- C4: build (both builders), batch A resident in HBM
    soup = ctx.make_soup(C4_TRIS)
    builds = {}
    scenes = {}
    for bname, b in (("binned-sah", 1), ("lbvh", 0)):
        sc = ctx.scene()
        sc.add_triangles(soup)
        sc.commit(b)
        sc.commit(b)  # steady state: the builder scratch is warm
        i = sc.info()
        builds[bname] = dict(build_ms=i["build_ms"], mtris_s=C4_TRIS / i["build_ms"] / 1e3, nodes=i["n_nodes"], depth=i["depth"], sah_cost=i["sah_cost"])
        scenes[bname] = sc
    sc = scenes["lbvh"]   # headline: the LBVH tree (build + trace is the config; the SAH tree is reported beside it)
    info = sc.info()
    d_rays, d_hits = ctx.alloc(n * 32), ctx.alloc(n * 16)
    ctx.make_rays_uniform(0, n, device_buffer=d_rays, host=False)

    for _ in range(warm):
        sc.trace_closest_device(d_rays, n, d_hits)
    stc = sc.trace_closest_device(d_rays, n, d_hits, ff.TRACE_BVH8 | ff.TRACE_COUNT)
    n_node, n_tri = stc["nodes_visited"] / n, stc["tris_tested"] / n

    sampler.mark_begin()
    secs, kern_ms, launches = 0.0, 0.0, 0
    for _ in range(args.steps):
        ctx.flush_l2()
        ctx.synchronize()
        ctx.timer_begin()
        st = sc.trace_closest_device(d_rays, n, d_hits)
        dt = 1e-3 * ctx.timer_end()
        secs += dt
        kern_ms += st["gpu_ms"]
        launches += st["kernel_launches"]
    value = n * args.steps / secs / 1e6
    log("C4 batch A: %.3f ms per step, %.1f Mrays/s (kernel %.3f ms)" % (1e3 * secs / args.steps, value, kern_ms / args.steps))
    dev_hits = sc.hits_unpack(d_hits, n)  # ids / t of the timed device-resident run

    # ---------------- e2e: ff_trace_closest with page-locked host buffers, copies inside the timed region
    h_rays = ctx.pinned_array((n, 6), np.float32, key="rays")
    h_rays[:] = ctx.make_rays_uniform(0, n)
    h_geom, h_tri, h_t = ctx.pinned_array((n,), np.int32, key="geom"), ctx.pinned_array((n,), np.int32, key="tri"), ctx.pinned_array((n,), np.float32, key="t")
    for _ in range(2):
        sc.trace_closest_into(h_rays, h_geom, h_tri, h_t)
    e_secs, e_launches, e_trace_ms = 0.0, 0, 0.0
    for _ in range(args.steps):
        ctx.flush_l2()
        ctx.synchronize()
        ctx.timer_begin()
        st = sc.trace_closest_into(h_rays, h_geom, h_tri, h_t)
        e_secs += 1e-3 * ctx.timer_end()
        e_launches += st["kernel_launches"]
        e_trace_ms += st["trace_ms"]
    sampler.mark_end()
    clocks = sampler.stop()
    e2e_value = n * args.steps / e_secs / 1e6
    log("C4 e2e: %.3f ms per step, %.1f Mrays/s" % (1e3 * e_secs / args.steps, e2e_value))
    same_paths = int((dev_hits["geom"] != h_geom).sum() + (dev_hits["tri"] != h_tri).sum() + (dev_hits["t"].view(np.uint32) != h_t.view(np.uint32)).sum())

    # ---------------- batch B (diffuse bounce off the primary hits) and the coherent primary batch, same tree
    cam4 = ff.Camera.from_dict(S.config4_camera(C4_W, C4_H))
    d_prim, d_ph, d_b = ctx.alloc(n * 32), ctx.alloc(n * 16), ctx.alloc(n * 32)
    ctx.camera_rays(cam4, device_buffer=d_prim, host=False)
    sc.trace_closest_device(d_prim, n, d_ph)
    nb = sc.make_rays_bounce(d_prim, d_ph, n, 1234, d_b)

    def timed_batch(scene, rays, count, hits, reps=5):
        for _ in range(2):
            scene.trace_closest_device(rays, count, hits)
        tot = 0.0
        for _ in range(reps):
            ctx.flush_l2()
            tot += scene.trace_closest_device(rays, count, hits)["gpu_ms"]
        c = scene.trace_closest_device(rays, count, hits, ff.TRACE_BVH8 | ff.TRACE_COUNT)
        nn, nt = c["nodes_visited"] / count, c["tris_tested"] / count
        return dict(rays=count, ms=tot / reps, mrays_s=count * reps / tot / 1e3, nodes_per_ray=nn, tris_per_ray=nt, bytes_per_ray=48 + 80 * nn + 48 * nt)
    batch_b = timed_batch(sc, d_b, nb, d_hits)
    primary = timed_batch(sc, d_prim, n, d_ph)
    sah_a = timed_batch(scenes["binned-sah"], d_rays, n, d_hits)
    # property at full size: the SAH tree and the LBVH tree answer every ray of batch A identically
    sah_hits = scenes["binned-sah"].hits_unpack(d_hits, n)
    trees_agree = int((sah_hits["geom"] != dev_hits["geom"]).sum() + (sah_hits["tri"] != dev_hits["tri"]).sum() +
                      (sah_hits["t"].view(np.uint32) != dev_hits["t"].view(np.uint32)).sum())
    for b in (d_prim, d_ph, d_b):
        b.free()

    # ---------------- hit-ID match + CPU baseline: the reference's intersectRays on the first rays of the SAME batch
    cpu, mism = None, None
    if "cpu" not in skip:
        try:
            arm = CpuArm(C4_TRIS)
            _, t = arm.intersect(np.ascontiguousarray(h_rays[:256]))
            n_chk = int(min(65536, max(2048, int(256 / max(t, 1e-6) * 22.0) // 1024 * 1024)))  # ~22 s of CPU work (multiple of 1024 rays), at most 65 536
            sample = np.ascontiguousarray(h_rays[:n_chk])
            want, t = arm.intersect(sample)
            cpu = dict(value=n_chk / t / 1e6, unit=UNIT, cores=arm.cores, kind=arm.kind, sample=arm.describe(n_chk), seconds=t)
            mism = dict(checked=n_chk, against="oracle/_ref (the reference's intersectRays)" if arm.kind == "reference" else "oracle port",
                        ids=int(((want["geom"] != h_geom[:n_chk]) | (want["tri"] != h_tri[:n_chk])).sum()),
                        t_bits=int((want["t"].view(np.uint32) != h_t[:n_chk].view(np.uint32)).sum()),
                        hits=int(want["hit"].sum()),
                        device_vs_host_path_full_batch=same_paths, lbvh_vs_sah_tree_full_batch=trees_agree)
            log("hit-ID match on %d rays: %s" % (n_chk, mism))
            del arm
        except Exception as e:  # the baseline is reported, never required for the GPU numbers
            log("cpu baseline failed:", repr(e))
    for s_ in scenes.values():
        s_.close()
    d_rays.free()
    d_hits.free()

    # ---------------- roofline of the dominant kernel
    peak, peak_src = measured_peak()
    prof = profile_numbers()
    bytes_per_ray = 32 + 16 + 80 * n_node + 48 * n_tri  # ray in, hit out, 80 B per node visited, 48 B per triangle tested
    launch_ms = kern_ms / args.steps
    achieved = bytes_per_ray * n / (launch_ms * 1e-3) / 1e9
    traffic = prof.get("c4_trace_closest_dram_bytes_per_launch")
    roofline = dict(bound="hbm", kernel="trace_closest_persistent_kernel<false> (one launch = the whole batch)", achieved=achieved, peak=peak, unit="GB/s",
                    frac=achieved / peak, traffic=traffic, peak_source=peak_src, bytes_per_ray=bytes_per_ray, nodes_per_ray=n_node, tris_per_ray=n_tri,
                    launch_ms=launch_ms, share_of_step=kern_ms / (1e3 * secs), algorithmic_bytes_per_launch=bytes_per_ray * n,
                    dram_frac=(traffic / (launch_ms * 1e-3) / 1e9 / peak) if traffic else None,
                    limiter=prof.get("c4_trace_closest_limiter", "instruction issue / L1 wavefronts: the 62 MB scene is L2-resident, DRAM carries rays in and hits out"),
                    note="achieved = ALGORITHMIC bytes (SURVEY 8d: 48 B ray + hit, 80 B per node visited, 48 B per triangle tested; n-bar from the instrumented "
                         "kernel on the same batch) / CUDA-event time of the launch: a normalised traversal-traffic figure served mostly by L1/L2.  "
                         "traffic = DRAM bytes per launch measured by ncu (profiles/traffic.json); dram_frac = traffic / launch time / peak is the share of "
                         "HBM bandwidth actually used")

    line = dict(
        metric=METRIC, value=value, unit=UNIT, n_gpus=1, steps=args.steps, warmup=warm, ms_per_step=1e3 * secs / args.steps, higher_is_better=True,
        scaling="weak", vs_baseline=None, dtype="f32", data="synthetic", config=bench_config(1),
        bvh=dict(builder="lbvh", build_ms=builds["lbvh"]["build_ms"], nodes=info["n_nodes"], depth=info["depth"], builders=builds),
        hit_id_mismatches=mism, roofline=roofline, cpu_baseline=cpu,
        e2e=dict(value=e2e_value, unit=UNIT, h2d_bytes_per_step=n * 24, d2h_bytes_per_step=n * 12, ms_per_step=1e3 * e_secs / args.steps,
                 trace_ms_per_step=e_trace_ms / args.steps, call="ff_trace_closest: page-locked host rays6 in, geometry / triangle ids + t out; one persistent launch consumes the rays as the copy engine lands them and hands finished chunks to the copy-out stream"),
        batch_b=batch_b, primary_coherent=primary, sah_tree_batch_a=sah_a,
        gpu_launches=launches + e_launches, clocks=clocks)

    # ---------------- secondary blocks: the frame configs
    meshes = S.load_meshes()
    if "c2" not in skip:
        line["c2"] = frame_block(ff, S, ctx, meshes, "c2", S.config2(meshes, C2["w"], C2["h"], C2["spp"], C2["bounces"]), C2, 64, 5, "cpu" in skip)
        log("c2:", json.dumps(line["c2"]))
    if "c3" not in skip:
        line["c3"] = frame_block(ff, S, ctx, meshes, "c3", S.config3(meshes, C3["w"], C3["h"], C3["spp"], C3["bounces"]), C3, 32, 2, "cpu" in skip,
                                 integrators=("wavefront", "megakernel"))
        log("c3:", json.dumps(line["c3"]))
    if "c5" not in skip:
        try:
            import gpupathtracer_b200.multigpu as M
            sc5, cam5 = c5_scene(ff, S, ctx, C5_TRIS)
            sc5.commit(0)                      # the fast builder on the same scene, for the record
            lbvh5 = sc5.info()["build_ms"]
            sc5.commit(C5_BUILDER)
            i5 = sc5.info()
            grp = M.Group.single(ctx)
            grp.render(sc5, cam5, 2, C5_BOUNCES, to_host=False)
            secs5, rays5 = 0.0, 0
            reps5 = 2
            for _ in range(reps5):
                ctx.synchronize()
                ctx.timer_begin()
                _, st = grp.render(sc5, cam5, C5_SPP, C5_BOUNCES, to_host=False)
                secs5 += 1e-3 * ctx.timer_end()
                rays5 += st["rays"]
            line["c5"] = dict(n_gpus=1, triangles=i5["n_triangles"], bvh_nodes=i5["n_nodes"], builder=BUILDER_NAMES[C5_BUILDER], build_ms=i5["build_ms"], lbvh_build_ms=lbvh5,
                              image=[C5_W, C5_H], total_spp=C5_SPP, max_bounces=C5_BOUNCES, frame_ms=1e3 * secs5 / reps5, mrays_s=rays5 / secs5 / 1e6,
                              msamples_s=C5_SPP * C5_W * C5_H * reps5 / secs5 / 1e6,
                              note="one-GPU base of the strong-scaling series that bench.py --gpus N reports as `value`")
            log("c5:", json.dumps(line["c5"]))
            sc5.close()
        except Exception as e:
            log("c5 block failed:", repr(e))
            line["c5"] = dict(error=repr(e))
    emit(line)
    return 0


# --------------------------------------------------------------------------------------------- GPU arm, N > 1: C5 strong split
def run_gpu_multi(args):
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    import torch
    import torch.distributed as td
    torch.cuda.set_device(local_rank)
    td.init_process_group("nccl", device_id=torch.device("cuda", local_rank))  # plumbing only: rendezvous, barriers, max / sum of scalars
    import gpupathtracer_b200 as ff
    import gpupathtracer_b200.multigpu as M
    from gpupathtracer_b200 import scenes as S
    ctx = ff.Context(local_rank)
    sc, cam = c5_scene(ff, S, ctx, args.c5_tris)
    info = sc.info()
    grp = M.Group.from_torch(ctx, td)   # ncclUniqueId from rank 0 through the torch store; communicator + reduce live in the library
    npix = C5_W * C5_H

    def barrier():
        ctx.synchronize()
        torch.cuda.synchronize()
        td.barrier()

    def allred(x, op):
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        td.all_reduce(t, op=op)
        return float(t.item())

    def frame(to_host):
        barrier()
        ctx.timer_begin()
        img, st = grp.render(sc, cam, C5_SPP, C5_BOUNCES, split="samples", reduce=args.reduce, to_host=to_host, pinned=True)
        barrier()
        return 1e-3 * ctx.timer_end(), st, img

    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    for _ in range(max(3, args.warmup)):
        frame(False)
    sampler.mark_begin()
    secs, rays, launches, reduce_ms, render_ms = 0.0, 0, 0, 0.0, 0.0
    for _ in range(args.steps):
        dt, st, _ = frame(False)
        secs += allred(dt, td.ReduceOp.MAX)
        rays += int(allred(float(st["rays"]), td.ReduceOp.SUM))
        launches += st["kernel_launches"]
        reduce_ms += allred(st["reduce_ms"], td.ReduceOp.MAX)
        render_ms += allred(st["gpu_ms"], td.ReduceOp.MAX)
    frame(True)
    e_secs, e_rays = 0.0, 0
    e_steps = max(1, min(args.steps, 5))
    img = None
    for _ in range(e_steps):
        dt, st, img = frame(True)
        e_secs += allred(dt, td.ReduceOp.MAX)
        e_rays += int(allred(float(st["rays"]), td.ReduceOp.SUM))
    sampler.mark_end()
    clocks = sampler.stop() if rank == 0 else None
    # the other reduce path, timed beside the default one
    other = "peer" if args.reduce == "nccl" else "nccl"
    other_ms = None
    try:
        grp.render(sc, cam, C5_SPP, C5_BOUNCES, split="samples", reduce=other, to_host=False)
        tot = 0.0
        for _ in range(3):
            barrier()
            _, st = grp.render(sc, cam, world, C5_BOUNCES, split="samples", reduce=other, to_host=False)
            tot += allred(st["reduce_ms"], td.ReduceOp.MAX)
        other_ms = tot / 3
    except Exception as e:
        log("rank %d: %s reduce unavailable: %r" % (rank, other, e))
        other_ms = None
    # one-GPU base in the same job + parity of the split frame against it (rank 0 alone, untimed region)
    base = None
    parity = None
    barrier()
    if rank == 0:
        solo = M.Group.single(ctx)
        solo.render(sc, cam, 2, C5_BOUNCES, to_host=False)
        ctx.synchronize()
        ctx.timer_begin()
        _, st1 = solo.render(sc, cam, C5_SPP, C5_BOUNCES, to_host=False)   # like the timed steps: the frame stays on the device
        t1 = 1e-3 * ctx.timer_end()
        base = dict(n_gpus=1, frame_ms=1e3 * t1, value=st1["rays"] / t1 / 1e6, unit=UNIT, msamples_s=C5_SPP * npix / t1 / 1e6,
                    note="the same frame (all %d spp) rendered by rank 0 alone in this job, frame left on the device like the timed steps" % C5_SPP)
        ref_img, _ = solo.render(sc, cam, C5_SPP, C5_BOUNCES, to_host=True)   # untimed: the one-GPU frame for the parity figure
        parity = dict(parity_vs_1gpu_max_abs=float(np.abs(img[..., :3] - ref_img[..., :3]).max()), frame_max=float(ref_img[..., :3].max()),
                      note="sample-split sums re-associate the per-pixel fp32 sum; tolerance 1e-5 x max(1, frame max)")
    barrier()
    if rank == 0:
        steps = max(1, args.steps)
        value = rays / secs / 1e6
        line = dict(metric=METRIC, value=value, unit=UNIT, n_gpus=world, steps=args.steps, warmup=max(3, args.warmup), ms_per_step=1e3 * secs / steps,
                    higher_is_better=True, scaling="strong", vs_baseline=None, dtype="f32", data="synthetic", config=bench_config(world),
                    bvh=dict(builder=BUILDER_NAMES[C5_BUILDER], build_ms=info["build_ms"], nodes=info["n_nodes"], replicated=True),
                    parallelism="sample-split x%d behind the C ABI (ff_group_render), 1 framebuffer reduce per frame (%s)" % (world, args.reduce),
                    samples_per_sec=C5_SPP * npix * steps / secs, samples_per_sec_per_gpu=C5_SPP * npix * steps / secs / world,
                    rays_per_step=rays / steps, render_ms_per_step=render_ms / steps, reduce_ms=reduce_ms / steps, reduce_bytes_per_rank=npix * 16,
                    reduce_ms_by_path={args.reduce: reduce_ms / steps, other: other_ms},
                    scaling_base=base, parity=parity,
                    e2e=dict(value=e_rays / e_secs / 1e6, unit=UNIT, h2d_bytes_per_step=ctypes_sizeof_inputs(ff), d2h_bytes_per_step=npix * 16,
                             ms_per_step=1e3 * e_secs / e_steps, steps=e_steps),
                    gpu_launches=launches, clocks=clocks)
        emit(line)
    td.barrier()
    grp.close()
    td.destroy_process_group()
    return 0


def ctypes_sizeof_inputs(ff):
    import ctypes
    return ctypes.sizeof(ff.CCamera) + ctypes.sizeof(ff.CRenderParams)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="firefly_b200", choices=["firefly_b200", "reference"])
    ap.add_argument("--reduce", default="peer", choices=["nccl", "peer"])
    ap.add_argument("--skip", default="", help="comma list of secondary work to leave out at N = 1: c2,c3,c5,cpu")
    ap.add_argument("--c5-tris", type=int, default=C5_TRIS, help=argparse.SUPPRESS)  # smaller soups for dry runs
    args = ap.parse_args()
    claim_stdout()
    if args.impl == "reference":
        return run_reference_arm(args)
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if args.gpus != world and world == 1 and args.gpus > 1:
        raise SystemExit("bench.py --gpus %d must be launched with torchrun --nproc-per-node %d" % (args.gpus, args.gpus))
    return run_gpu_one(args) if world == 1 else run_gpu_multi(args)


if __name__ == "__main__":
    sys.exit(main())
