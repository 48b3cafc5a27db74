#!/usr/bin/env python3
"""bench.py -- the headline measurement (BASELINE.json): closest-hit Mrays/s and samples/s/GPU of the path-tracing
hot path on B200, with the roofline of the dominant kernel and the reference's CPU path timed beside it.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference] [--reduce nccl|peer]

Workload (N=1): BASELINE.json configs[1] -- rocketman.obj inside cube.obj, 1024x1024, 64 spp, 4 bounces, diffuse.
A step = one frame (all 64 spp) through the wavefront integrator.  For N>1 (torchrun, one rank per GPU) every rank
renders its own 64 global sample indices of the same frame (weak scaling: 64*N spp in total) and the framebuffers are
combined by ONE reduce per frame.  `value` is measured with the frame resident on the device; `e2e` goes through the
C-ABI call a user makes (ff_render into a HOST buffer: camera/params in, W*H RGBA fp32 framebuffer out).
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
WORKLOAD = "C2: rocketman.obj in cube.obj (6048 triangles), 1024x1024, 64 spp, 4 bounces diffuse, wavefront integrator"
W, H, SPP, BOUNCES = 1024, 1024, 64, 4


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
    """The reference's own closest-hit routine (intersectRays + intersectTriangle, kernel.cu:35-176: brute force over
    all 6048 triangles for every ray) on the host cores, on a bounded sample of the workload's rays: camera rays of
    the C2 frame taken with a stride so they cover the image (brute-force cost per ray does not depend on the ray).
    Uses oracle/_ref (the reference's text compiled from /root/reference) when present, else the oracle restatement."""

    def __init__(self):
        from oracle import oracle as O
        from gpupathtracer_b200 import scenes as S
        meshes = S.load_meshes()
        desc = S.config2(meshes, W, H, SPP, BOUNCES)
        if O.Ref.available():
            self.lib, self.kind = O.Ref(), "reference"
        else:
            if not O.Oracle.available():
                O.build()
            self.lib, self.kind = O.Oracle(), "port"
        self.scene = S.instantiate_on(desc, self.lib.scene(), meshes)
        self.rays, _ = self.lib.camera_rays(desc["camera"])
        self.cores = int(self.lib.threads)

    def run(self, n_rays):
        stride = max(1, len(self.rays) // n_rays)
        sample = np.ascontiguousarray(self.rays[::stride][:n_rays])
        t = self.scene.time_intersect(sample, threads=self.cores, reps=1, mode=0)
        what = ("%d camera rays of the C2 frame (stride %d), each tested against all 6048 triangles by the reference's "
                "intersectRays; %d host threads" % (len(sample), stride, self.cores))
        return len(sample), t, what


def run_reference_arm(args):
    if int(os.environ.get("RANK", "0")) != 0:
        return 0
    arm = CpuArm()
    n, t, _ = arm.run(4096)  # calibration
    per_step = int(max(4096, min(1 << 20, n / t * 3.0)))  # ~3 s of CPU work per step
    for _ in range(args.warmup):
        arm.run(max(4096, per_step // 8))
    rays, secs, what = 0, 0.0, ""
    for _ in range(args.steps):
        n, t, what = arm.run(per_step)
        rays += n
        secs += t
    value = rays / secs / 1e6 if secs > 0 else 0.0
    line = dict(metric=METRIC, value=value, unit=UNIT, impl="reference", n_gpus=args.gpus, steps=args.steps, warmup=args.warmup,
                ms_per_step=1e3 * secs / max(1, args.steps), higher_is_better=True, scaling="weak", vs_baseline=None, dtype="f32",
                data="synthetic", config=dict(workload=WORKLOAD, sample_per_step=what),
                cpu_baseline=dict(value=value, unit=UNIT, cores=arm.cores, kind=arm.kind, sample=what),
                e2e=dict(value=value, unit=UNIT, h2d_bytes_per_step=0, d2h_bytes_per_step=0), gpu_launches=0)
    emit(line)
    return 0


# --------------------------------------------------------------------------------------------- GPU arm
def run_gpu_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    dist = None
    if world > 1:
        import torch
        import torch.distributed as td
        torch.cuda.set_device(local_rank)
        td.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
        dist = td
    import gpupathtracer_b200 as ff
    from gpupathtracer_b200 import multigpu as M
    from gpupathtracer_b200 import scenes as S

    if ff.device_count() == 0:
        raise SystemExit("bench.py: no CUDA device; the product path has no CPU fallback")
    ctx = ff.Context(local_rank)
    meshes = S.load_meshes()
    desc = S.config2(meshes, W, H, SPP, BOUNCES)
    sc, cam = S.instantiate(desc, ctx, meshes, builder=args.builder)
    info = sc.info()
    npix = W * H
    sc_spb = max(1, (64 << 20) // npix)  # samples per wavefront pass the library picks for this image (ff_api.cpp accumulate_samples)
    first = rank * SPP  # weak scaling: every rank renders its own 64 global sample indices
    total_spp = SPP * world
    d_out = ctx.alloc(npix * 16)
    host_img = np.empty((H, W, 4), np.float32)

    def barrier():
        ctx.synchronize()
        if dist is not None:
            import torch
            torch.cuda.synchronize()
            dist.barrier()

    def step_device():
        """One frame, result left on the device; returns (seconds on the device clock: CUDA events on the library's stream
        bracketing the whole step incl. the framebuffer reduce, barrier + synchronize on both sides; stats)."""
        barrier()
        ctx.timer_begin()
        if world == 1:
            _, st = sc.render(cam, SPP, BOUNCES, want_stats=True, host=False, device_out=d_out, first_sample=first)
        else:
            _, st = M.render_distributed_weak(sc, cam, SPP, BOUNCES, dist, mode=args.reduce, to_host=False)
        barrier()
        return 1e-3 * ctx.timer_end(), st

    def step_e2e():
        """The user-facing call with HOST buffers: camera/params in (by value), framebuffer out to page-locked host memory."""
        barrier()
        ctx.timer_begin()
        if world == 1:
            _, st = sc.render(cam, SPP, BOUNCES, want_stats=True, first_sample=first, pinned=True)
        else:
            _, st = M.render_distributed_weak(sc, cam, SPP, BOUNCES, dist, mode=args.reduce, to_host=True, pinned=True)
        barrier()
        return 1e-3 * ctx.timer_end(), st

    def maxreduce(x):
        if dist is None:
            return x
        import torch
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def sumreduce(x):
        if dist is None:
            return x
        import torch
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
        return float(t.item())

    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    for _ in range(max(3, args.warmup)):
        step_device()
    # instrumented frame (untimed): mean nodes / triangles per ray for the algorithmic-bytes figure
    _, stc = sc.render(cam, SPP, BOUNCES, want_stats=True, host=False, device_out=d_out, first_sample=first,
                       integrator=ff.INTEGRATOR_WAVEFRONT | ff.TRACE_COUNT)
    n_node, n_tri = stc["nodes_visited"] / stc["rays"], stc["tris_tested"] / stc["rays"]

    sampler.mark_begin()
    secs, trace_ms, gpu_ms, rays, launches = 0.0, 0.0, 0.0, 0, 0
    for _ in range(args.steps):
        ctx.flush_l2()
        dt, st = step_device()
        log("rank %d device step: %.3f ms (device clock), library kernels %.3f ms" % (rank, 1e3 * dt, st["gpu_ms"]))
        secs += maxreduce(dt)
        trace_ms += st["trace_ms"]
        gpu_ms += st["gpu_ms"]
        rays += int(sumreduce(st["rays"]))
        launches += st["kernel_launches"]
    step_e2e()  # untimed: first use of the host-output path (pinned staging, NCCL channel set-up for this buffer)
    e_secs, e_rays = 0.0, 0
    for _ in range(args.steps):
        ctx.flush_l2()
        dt, st = step_e2e()
        log("rank %d e2e step: %.3f ms (device clock), library kernels %.3f ms" % (rank, 1e3 * dt, st["gpu_ms"]))
        e_secs += maxreduce(dt)
        e_rays += int(sumreduce(st["rays"]))
    sampler.mark_end()
    clocks = sampler.stop() if rank == 0 else None

    if rank == 0:
        steps = max(1, args.steps)
        value = rays / secs / 1e6
        peak, peak_src = measured_peak()
        bytes_per_ray = 32 + 16 + 80 * n_node + 48 * n_tri  # ray in, hit out, 80 B nodes, 48 B triangles
        rays_rank0 = rays / world
        n_extend = ((SPP + sc_spb - 1) // sc_spb) * (BOUNCES + 1)  # extend launches per frame: passes x (bounces + 1)
        achieved = bytes_per_ray * rays_rank0 / (trace_ms * 1e-3) / 1e9 if trace_ms > 0 else 0.0
        traffic = None
        tp = os.path.join(ROOT, "profiles", "traffic.json")
        if os.path.exists(tp):
            try:
                traffic = json.load(open(tp)).get("trace_closest_bytes_per_launch")
            except Exception:
                traffic = None
        cpu = None
        if world == 1 or True:
            try:
                arm = CpuArm()
                n, t, _ = arm.run(4096)
                n, t, what = arm.run(int(max(4096, min(1 << 20, n / t * 12.0))))  # ~12 s of CPU work
                cpu = dict(value=n / t / 1e6, unit=UNIT, cores=arm.cores, kind=arm.kind, sample=what)
            except Exception as e:  # the baseline is reported, never required for the GPU numbers
                log("cpu baseline failed:", e)
        line = dict(
            metric=METRIC, value=value, unit=UNIT, n_gpus=world, steps=args.steps, warmup=max(3, args.warmup),
            ms_per_step=1e3 * secs / steps, higher_is_better=True, scaling="weak", vs_baseline=None, dtype="f32", data="synthetic",
            config=dict(workload=WORKLOAD, image=[W, H], spp_per_gpu=SPP, total_spp=total_spp, max_bounces=BOUNCES, triangles=info["n_triangles"],
                        bvh_nodes=info["n_nodes"], builder=("lbvh", "binned-sah")[args.builder], build_ms=info["build_ms"], parallelism="sample-split x%d, 1 framebuffer reduce (%s)" % (world, args.reduce),
                        l2="flushed with a 256 MiB write between timed steps"),
            samples_per_sec_per_gpu=SPP * npix * steps / secs,
            rays_per_step=rays / steps, device_ms_per_step=gpu_ms / steps, trace_ms_per_step=trace_ms / steps,
            roofline=dict(bound="hbm", kernel="trace_closest_persistent_kernel (extend stage, all bounces of all passes)", achieved=achieved, peak=peak, unit="GB/s",
                          frac=achieved / peak, traffic=traffic, peak_source=peak_src, bytes_per_ray=bytes_per_ray, nodes_per_ray=n_node,
                          tris_per_ray=n_tri, share_of_step=trace_ms / gpu_ms if gpu_ms else None,
                          launches_per_step=n_extend, algorithmic_bytes_per_launch=bytes_per_ray * rays_rank0 / steps / n_extend,
                          note="achieved = algorithmic bytes (SURVEY 8d: 48 B ray+hit, 80 B per node visited, 48 B per triangle tested) / CUDA-event time "
                               "of the extend launches; traffic = measured DRAM bytes per launch (ncu, profiles/traffic.json): the scene is "
                               "L2-resident, so DRAM carries only rays in / hits out"),
            cpu_baseline=cpu,
            e2e=dict(value=e_rays / e_secs / 1e6, unit=UNIT, h2d_bytes_per_step=ctypes_sizeof_inputs(ff), d2h_bytes_per_step=npix * 16,
                     ms_per_step=1e3 * e_secs / steps),
            gpu_launches=launches, clocks=clocks)
        emit(line)
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()
    return 0


def ctypes_sizeof_inputs(ff):
    import ctypes
    return ctypes.sizeof(ff.CCamera) + ctypes.sizeof(ff.CRenderParams)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="firefly_b200", choices=["firefly_b200", "reference"])
    ap.add_argument("--reduce", default="nccl", choices=["nccl", "peer"])
    ap.add_argument("--builder", type=int, default=1, choices=[0, 1], help="0 = LBVH, 1 = binned SAH (build time is reported in config.build_ms, outside the timed region)")
    args = ap.parse_args()
    claim_stdout()
    if args.impl == "reference":
        return run_reference_arm(args)
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if args.gpus != world:
        if world == 1 and args.gpus > 1:
            raise SystemExit("bench.py --gpus %d must be launched with torchrun --nproc-per-node %d" % (args.gpus, args.gpus))
    return run_gpu_arm(args)


if __name__ == "__main__":
    sys.exit(main())
