#!/usr/bin/env python
"""bench.py -- headline benchmark of the B200 voxel ray caster (contract: see DESIGN.md "Measurement").

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference] [--workload NAME]

A "step" is one frame: rc_render_frame over one batch of W x H synthetic camera rays through the
default generated world (initTetraHexaTree + genWorld: seeds 42/64/100, 200 x 200 columns).
  N = 1 : BASELINE.json configs[2] -- 1920x1080, primary + sun shadow + 64-sample hemisphere AO
          (the configuration `metric` is quoted on).
  N > 1 : BASELINE.json configs[4] -- 7680x4320, same passes, 16-row bands dealt round-robin to the
          ranks; every rank's kernels store their pixels straight into rank 0's frame over NVLink peer memory
          (--frame-path gather: NCCL gather of the band buffers + one compose kernel instead); strong scaling.
One JSON line on stdout (rank 0). `value` = rays cast by all ranks / device time (max over ranks) with
everything resident in HBM; `e2e` = the same through the C-ABI with the frame read back into pinned
host memory every step (rc_render_frame + rc_read_image).
`--impl reference` times the reference's CPU side on all host threads: for the workload its live shader covers
in full (config 2) the shader's own source compiled as C++ (oracle/_ref/libref_shader.so); where AO is on -- a pass
the reference never implemented -- the restatement (oracle/oracle_render.cpp) on a bounded sample of the workload.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

UNIT = "Mrays/s"


def metric_name(workload, n_gpus):
    """BASELINE.json metric: "Mrays/s (primary+AO) at 1920x1080 per B200; frame ms at 1/2/4/8 GPUs vs CPU" -- the first half
    names the N = 1 line, the second the sharded ones (ms_per_step is the frame ms); the resolution is the line's own."""
    w, h = WORKLOADS[workload][0], WORKLOADS[workload][1]
    if n_gpus == 1:
        return f"Mrays/s (primary+AO) at {w}x{h} per B200"
    return f"Mrays/s (primary+AO), {w}x{h} frame sharded over {n_gpus} B200 (frame ms = ms_per_step)"


WORKLOADS = {
    # name: (width, height, shadows, ao_samples, golden case holding the oracle's algorithmic byte counts)
    "config1_720p_primary": (1280, 720, False, 0, "config1_720p_primary"),
    "config2_1080p_shadow": (1920, 1080, True, 0, "config2_1080p_shadow"),
    "config3_1080p_ao64": (1920, 1080, True, 64, "config3_1080p_ao64"),
    "config4_4k_ao64_world2048": (3840, 2160, True, 64, "config4_4k_ao64_world2048"),
    "config5_8k_ao64": (7680, 4320, True, 64, "config5_8k_ao64"),
}
HBM_FALLBACK_GBS = 6650.0  # B200_PROFILING.md, used only when MEASURED_PEAKS.json is absent


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default=None, choices=sorted(WORKLOADS))
    ap.add_argument("--no-cpu-baseline", action="store_true", help="skip the cpu_baseline leg (profiling runs)")
    ap.add_argument("--cpu-baseline-only", action="store_true",
                    help="internal: print the cpu_baseline object of --workload as JSON and exit (the default arm runs this leg in a "
                         "subprocess, so that the process that drives the GPU never maps an oracle library)")
    ap.add_argument("--completion-signal", dest="signal", choices=["flags", "allreduce"], default="flags",
                    help="N>1, direct frame: how 'frame complete' crosses the processes (flags in rank 0's memory, or the "
                         "one-element NCCL all-reduce they replaced)")
    ap.add_argument("--schedule", default=None, choices=["two_kernels", "ao_persistent", "fused", "persistent", "persistent_staged"],
                    help="A/B runs: how librc schedules a frame on the device (default: the library's default)")
    ap.add_argument("--no-e2e", action="store_true", help="skip the e2e leg (profiling runs)")
    ap.add_argument("--frame-path", choices=["direct", "gather"], default="direct",
                    help="N>1: how the bands become one frame on rank 0 (direct = peer stores from the kernels over NVLink; "
                         "gather = NCCL gather + compose kernel, the comparison baseline)")
    return ap.parse_args()


def hbm_peak():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    try:
        with open(p) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return HBM_FALLBACK_GBS, "fallback (B200_PROFILING.md)"


def golden_counters(case):
    """Algorithmic bytes of the reference algorithm for exactly this frame, counted by the oracle when
    tests/golden was generated (tools/make_golden.py): 16 B per nodes[] read + 4 B per child-index read."""
    g = golden_case(case)
    return g["counters"] if g else None


def golden_case(case):
    """The oracle's record of exactly the frame bench.py renders for a workload (same world, camera and lookingAtBlock):
    configs 1-3 tests/golden/frames_bench.json (tools/make_golden_bench.py), configs 4-5 frames_big.json."""
    for name in ("frames_bench.json", "frames_big.json"):
        try:
            with open(os.path.join(ROOT, "tests", "golden", name)) as f:
                cases = json.load(f)["cases"]
            if case in cases:
                return dict(cases[case], file=f"tests/golden/{name}")
        except Exception:
            pass
    return None


def sha256_of(array):
    import hashlib

    return hashlib.sha256(np.ascontiguousarray(array).tobytes()).hexdigest()


def kernel_sources_sha256():
    """Same digest csrc/Makefile bakes into librc.so (rc_build_id) and tools/ncu_traffic.py records with a capture."""
    import hashlib

    h = hashlib.sha256()
    for name in ("kernels.cu", "kernels.h", "free_flight.h"):
        with open(os.path.join(ROOT, "raytracing_test_b200", "csrc", name), "rb") as f:
            h.update(f.read())
    return h.hexdigest()


def capture_record(workload):
    """(per-kernel records, kernel-source digest) of the committed ncu capture of this workload, or ({}, None)."""
    try:
        with open(os.path.join(ROOT, "profiles", "traffic.json")) as f:
            j = json.load(f)
        return j.get(workload, {}), j.get("_kernel_sources_sha256", {}).get(workload)
    except Exception:
        return {}, None


class ClockSampler:
    """nvidia-smi clocks / throttle reasons while the benchmark runs (B200_PROFILING.md clocks line)."""

    FIELDS = ("index,clocks.sm,clocks.max.sm,power.draw,utilization.gpu,clocks_event_reasons.hw_slowdown,"
              "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
              "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.rows = []
        self.proc = None
        self.gpu = gpu_index

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", "-i", str(self.gpu), f"--query-gpu={self.FIELDS}", "--format=csv,noheader,nounits", "-lms", "50"],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._pump, daemon=True).start()
        except Exception:
            self.proc = None

    def _pump(self):
        for line in self.proc.stdout:
            self.rows.append((time.time(), [c.strip() for c in line.split(",")]))

    def stop(self, t0=None, t1=None):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        time.sleep(0.06)
        self.proc.terminate()
        rows = [r for (t, r) in self.rows if len(r) >= 9 and (t0 is None or t0 - 0.05 <= t <= t1 + 0.05)]
        if not rows:
            rows = [r for (_, r) in self.rows if len(r) >= 9]
        sm = [float(r[1]) for r in rows if r[1].replace(".", "").isdigit()]
        busy = [float(r[1]) for r in rows if r[4].isdigit() and int(r[4]) > 0 and r[1].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for i, n in enumerate(names) if any(r[5 + i].lower().startswith("active") for r in rows)]
        return {
            "sm_mhz": float(np.median(busy if busy else sm)) if sm else None,
            "sm_max_mhz": float(rows[0][2]) if rows and rows[0][2].replace(".", "").isdigit() else None,
            "power_w_max": max((float(r[3]) for r in rows if r[3].replace(".", "").isdigit()), default=None),
            "reasons": reasons, "samples": len(rows),
        }


# ---------------------------------------------------------------------------------------------
# CPU arm: the reference algorithm restated on the CPU (oracle/), used ONLY as a timed baseline
# ---------------------------------------------------------------------------------------------
def build_host_world(workload):
    """The synthetic world, in the reference's pool wire format, from the host stand-in."""
    from raytracing_test_b200 import voxel_host as vh

    if workload.startswith("config4"):
        return vh.World.default(width=2048, length=2048, max_depth=7), 7
    return vh.World.default(), 6


def host_threads():
    """Threads the CPU arm asks for: every core this process may run on.  Passed EXPLICITLY to the oracle (num_threads
    clause): launchers such as torchrun export OMP_NUM_THREADS=1 for nproc > 1, which a default-sized OpenMP team obeys."""
    try:
        return len(os.sched_getaffinity(0))
    except AttributeError:
        return os.cpu_count() or 1


def cpu_render_sample(workload, stride, threads=None, world_cache={}):
    """One bounded sample of the workload on the host cores: every stride-th 16-row band of the frame,
    all passes. Returns (rays, seconds, algorithmic bytes by phase); cpu_render_sample.last_threads = the OpenMP team
    the run really got."""
    from oracle import bindings as ob  # the checker, timed here as the CPU baseline only
    from raytracing_test_b200 import voxel_host as vh

    w, h, shadows, ao, _ = WORKLOADS[workload]
    if workload not in world_cache:
        host, depth = build_host_world(workload)
        nodes, arrays = host.pools()
        world_cache[workload] = (ob.make_world(nodes, arrays, depth), nodes, arrays)
        if ao:
            tab = np.zeros((ao, 3), dtype=np.float32)
            ob.oracle_lib().oracle_hemisphere(ao, tab.ctypes.data)
            ob.oracle_lib().oracle_set_hemisphere(ao, tab.ctypes.data)
    ow = world_cache[workload][0]
    pos, d, sun = vh.default_camera()
    frame = ob.make_frame(w, h, pos, d, sun, shadows=shadows, ao_samples=ao, threads=threads if threads else host_threads(),
                          bands=(16, stride, 0) if stride > 1 else None)
    t = time.perf_counter()
    r = ob.render(ow, frame, want_hits=False, want_rgba8=True)
    dt = time.perf_counter() - t
    cpu_render_sample.last_threads = int(ob.oracle_lib().oracle_last_threads())
    c = r["counters"]
    return int(sum(c.rays)), dt, c.as_dict()["algorithmic_bytes"]


def run_reference_arm(args, workload):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    w, h, shadows, ao, _ = WORKLOADS[workload]
    if shadows and not ao and reference_shader_sample(workload, reps=1) is not None:
        return run_reference_shader_arm(args, workload)  # config 2 = the live shader: time the reference's own source
    stride = max(1, (w * h) // (1920 * 1080) * 8)
    # keep the whole run within a few minutes whatever K is
    rays, dt, _ = cpu_render_sample(workload, stride)
    while dt * (args.steps + args.warmup) > 150.0 and stride < 64:
        stride *= 2
        rays, dt, _ = cpu_render_sample(workload, stride)
    for _ in range(max(0, args.warmup - 1)):
        cpu_render_sample(workload, stride)
    total_rays, total_s = 0, 0.0
    for _ in range(args.steps):
        rays, dt, _ = cpu_render_sample(workload, stride)
        total_rays += rays
        total_s += dt
    value = total_rays / total_s / 1e6
    cores = cpu_render_sample.last_threads  # the OpenMP team the timed runs really had
    sample = f"every {stride}th 16-row band of the {w}x{h} frame per step ({rays} rays/step), all passes, OpenMP x{cores}"
    line = {
        "metric": metric_name(workload, args.gpus), "value": value, "unit": UNIT, "impl": "reference", "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": total_s / args.steps * 1e3, "higher_is_better": True,
        "scaling": "strong" if args.gpus > 1 else "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": workload, "resolution": [w, h], "shadows": shadows, "ao_samples": ao,
                   "world": "default generated world (seeds 42/64/100, 200x200)", "sample": sample},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
        "note": "the reference renders frames only in GLSL (unavailable on this box); this is its traversal restated in "
                "C++ (oracle/oracle_render.cpp, IEEE fp32) on all host threads",
    }
    shader = reference_shader_sample(workload)
    if shader is not None:
        # beside the port: the reference's own shader source on the same cores, for the passes it implements
        line["cpu_baseline"]["reference_shader"] = shader
    print(json.dumps(line), flush=True)


def run_reference_shader_arm(args, workload):
    """--impl reference for the workload the reference's live shader covers in full (primary + shadow rays):
    each step is one pass of low_res.frag, compiled as C++, over the frame on all host threads."""
    w, h, shadows, ao, _ = WORKLOADS[workload]
    for _ in range(args.warmup):
        reference_shader_sample(workload, reps=1)
    total_rays, total_s, last = 0, 0.0, None
    for _ in range(args.steps):
        last = reference_shader_sample(workload, reps=1)
        total_rays += last["rays"]
        total_s += last["seconds"]
    value = total_rays / total_s / 1e6
    line = {
        "metric": metric_name(workload, args.gpus), "value": value, "unit": UNIT, "impl": "reference", "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": total_s / args.steps * 1e3, "higher_is_better": True,
        "scaling": "strong" if args.gpus > 1 else "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": workload, "resolution": [w, h], "shadows": shadows, "ao_samples": ao,
                   "world": "default generated world (seeds 42/64/100, 200x200), built by the reference's own code",
                   "sample": last["sample"]},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": last["cores"], "kind": "reference", "sample": last["sample"]},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
        "note": "the reference's own fragment shader (src/shaders/low_res.frag) compiled as C++ for the host cores "
                "(oracle/ref_shader.cpp); GLSL on a GL driver is unavailable on this box",
    }
    print(json.dumps(line), flush=True)


def reference_ray_caster_sample(workload, target_rays=40000):
    """The reference's OWN CPU traversal -- RAY_CASTER::castRayFromCam (src/ray_caster.cpp:54-87: double deltas, one
    getBlock descent from the root per step), compiled unmodified into oracle/_ref/libref_world.so -- run with the
    frame's step budget over a regular sample of this frame's primary ray directions, over the reference-built default
    world.  BASELINE.json configs[0] names it as the reference CPU path.  One thread by construction (the camera is a
    global of the reference).  None when the library was not built (needs /root/reference once, in the build container)."""
    from oracle import bindings as ob  # checker side, timed as a baseline only
    from raytracing_test_b200 import voxel_host as vh

    if workload.startswith("config4"):
        return None  # the reference's tree cannot hold the 2048^2 world (maxDepth 6)
    try:
        rw = ob.RefWorld()
    except (FileNotFoundError, OSError):
        return None
    w, h, _, _, _ = WORKLOADS[workload]
    pos, d, _ = vh.default_camera()
    f = np.float32
    step = max(1, int(np.sqrt(w * h / target_rays)))
    px, py = np.meshgrid(np.arange(step // 2, w, step), np.arange(step // 2, h, step))
    u = ((px.astype(f) + f(0.5)) * f(f(1.0) / f(w))).ravel()
    v = ((py.astype(f) + f(0.5)) * f(f(1.0) / f(h))).ravel()
    cam = d.astype(f)
    left = np.cross(cam, np.array([0, 1, 0], dtype=f)).astype(f)  # low_res.frag:279-285
    up = np.cross(cam, left).astype(f)
    ppx, ppy = f(np.tan(np.radians(45.0))), f(np.tan(np.radians(45.0)) * f(h) / w)
    dirs = cam[None, :] + left[None, :] * (-(ppx * (u - f(0.5))))[:, None] + up[None, :] * ((-v + f(0.5)) * ppy)[:, None]
    dirs = (dirs / np.sqrt((dirs * dirs).sum(axis=1, keepdims=True))).astype(f)
    dirs = np.ascontiguousarray(dirs)
    out = np.zeros((len(dirs), 7), dtype=np.int32)
    p = np.ascontiguousarray(pos, dtype=f)
    best = None
    for _ in range(3):
        t = time.perf_counter()
        rw.lib.refworld_pick_batch(p.ctypes.data, dirs.ctypes.data, len(dirs), 300, out.ctypes.data)
        dt = time.perf_counter() - t
        best = dt if best is None else min(best, dt)
    return {"value": len(dirs) / best / 1e6, "unit": "Mrays/s (primary rays only)", "cores": 1, "kind": "reference",
            "sample": f"castRayFromCam(300) for every {step}th pixel of every {step}th row of the {w}x{h} frame ({len(dirs)} rays), "
                      f"best of 3 ({best * 1e3:.0f} ms)",
            "hits": int((out[:, 6] > 0).sum())}


def reference_shader_sample(workload, threads=None, reps=3, state={}):
    """The reference's OWN renderer on the host cores: src/shaders/low_res.frag compiled as C++ (oracle/ref_shader.cpp ->
    oracle/_ref/libref_shader.so; GLSL itself cannot run on this box), one instance per pixel, OpenMP over rows, over the
    world the reference's unmodified putBlock / genWorld built.  It is the live shader as it stands: primary + sun-shadow
    rays (BASELINE configs 1-2; the reference has no AO pass to time).  Rays = pixels + shadow rays cast, the latter
    counted by the oracle on the same frame (tests/test_ref_shader_cpu.py: same traversal float for float).
    None when the library was not built (needs /root/reference once, in the build container)."""
    from oracle import bindings as ob  # checker side, timed as a baseline only
    from raytracing_test_b200 import voxel_host as vh

    if workload.startswith("config4"):
        return None  # MAX_DEPTH = 6 is a constant of the shader (low_res.frag:110)
    w, h, _, _, _ = WORKLOADS[workload]
    try:
        if "world" not in state:
            rw = ob.RefWorld()
            state["world"] = (rw, ob.RefShader(rw.nodes, rw.arrays, oracle_math=False),
                              ob.make_world(rw.nodes, rw.arrays, rw.max_depth))
    except (FileNotFoundError, OSError):
        return None
    rw, shader, oworld = state["world"]
    pos, d, sun = vh.default_camera()
    row_stride = max(1, (w * h) // (1920 * 1080) // 2)  # 8K: every 8th row (about one 1080p frame of work)
    key = (w, h, row_stride)
    if key not in state:
        # the oracle renders bands of rows; row_stride-spaced single rows = bands of 1 row with that stride
        frame = ob.make_frame(w, h, pos, d, sun, shadows=True, ao_samples=0, bands=(1, row_stride, 0) if row_stride > 1 else None)
        c = ob.render(oworld, frame, want_hits=False, want_rgba8=False)["counters"]
        rays = int(c.rays[0] + c.rays[1])
        assert int(c.rays[0]) == len(range(0, h, row_stride)) * w
        state[key] = rays
    rays = state[key]
    best = None
    for _ in range(reps):
        t = time.perf_counter()
        out = shader.render(w, h, pos, d, sun, row_stride=row_stride, threads=threads if threads else host_threads(), want_state=False)
        dt = time.perf_counter() - t
        best = dt if best is None else min(best, dt)
    n, cores = out["fragments"], out["threads"]  # the OpenMP team the run really had
    what = "whole" if row_stride == 1 else f"every {row_stride}th row of the"
    return {"value": rays / best / 1e6, "unit": "Mrays/s (primary + shadow rays: the live shader)", "cores": cores,
            "kind": "reference", "rays": rays, "seconds": best, "fragments": n,
            "sample": f"low_res.frag compiled as C++, {what} {w}x{h} frame ({n} fragments, {rays} rays), best of {reps} "
                      f"({best * 1e3:.0f} ms), OpenMP x{cores}"}


def cpu_baseline_leg(workload):
    """cpu_baseline object of the default arm: about 10-30 s of CPU work on rank 0."""
    w, h, shadows, ao, _ = WORKLOADS[workload]
    stride = max(1, (w * h) // (1920 * 1080))
    cpu_render_sample(workload, 64)  # builds the world, warms the caches
    best, spent, reps, rays = None, 0.0, 0, 0
    while reps < 5 and (spent < 10.0 or reps < 2):
        rays, dt, _ = cpu_render_sample(workload, stride)
        best = dt if best is None else min(best, dt)
        spent += dt
        reps += 1
    cores = cpu_render_sample.last_threads
    one_rays, one_dt, _ = cpu_render_sample(workload, max(stride, 4), threads=1)
    port = {
        "value": rays / best / 1e6, "unit": UNIT, "cores": cores, "kind": "port",
        "sample": f"{'whole' if stride == 1 else f'every {stride}th 16-row band of the'} {w}x{h} frame, all passes, "
                  f"best of {reps} ({best * 1e3:.0f} ms), OpenMP x{cores}",
        "frame_ms": best * 1e3 * stride, "single_thread_value": one_rays / one_dt / 1e6,
    }
    ref = reference_ray_caster_sample(workload)
    shader = reference_shader_sample(workload)
    if shader is not None and shadows and not ao:
        # config 2 (primary + shadow) is exactly the reference's live shader: its own source, compiled for the host, IS the baseline
        out = dict(shader, unit=UNIT, port=port)
        if ref is not None:
            out["reference_ray_caster"] = ref
        return out
    if ref is not None and not shadows and not ao:
        # config 1 (primary rays only) is exactly what the reference's own CPU ray caster does: it IS the baseline
        out = dict(ref, unit=UNIT, port=port)
        if shader is not None:
            out["reference_shader"] = shader
        return out
    if ref is not None:
        port["reference_ray_caster"] = ref  # the reference's own compiled CPU traversal, primary rays only, beside the port
    if shader is not None:
        port["reference_shader"] = shader   # the reference's own shader source on the host cores: primary + shadow rays (no AO pass exists)
    return port


# ---------------------------------------------------------------------------------------------
# B200 arm
# ---------------------------------------------------------------------------------------------
def cpu_baseline_subprocess(workload):
    """The cpu_baseline leg in a process of its own: the process that drives the GPU never maps an oracle library."""
    try:
        out = subprocess.run([sys.executable, os.path.abspath(__file__), "--cpu-baseline-only", "--workload", workload],
                             capture_output=True, text=True, timeout=900)
        for line in reversed(out.stdout.strip().splitlines()):
            if line.startswith("{"):
                return json.loads(line)
        return {"unavailable": (out.stderr or out.stdout)[-300:]}
    except Exception as e:  # noqa: BLE001
        return {"unavailable": repr(e)}


def run_b200_arm(args, workload):
    import torch
    import torch.distributed as dist

    from raytracing_test_b200 import frame_gather as fg
    from raytracing_test_b200 import rc, voxel_host as vh

    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world_size = int(os.environ.get("WORLD_SIZE", "1"))
    if world_size != args.gpus:
        raise SystemExit(f"--gpus {args.gpus} but WORLD_SIZE={world_size}: launch with torch.distributed.run --nproc-per-node {args.gpus}")
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; librc has no CPU path (use --impl reference for the CPU baseline)")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world_size > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        import datetime

        # a mismatched collective must not hold eight GPUs for NCCL's default ten minutes
        dist.init_process_group("nccl", device_id=dev, timeout=datetime.timedelta(seconds=120))

    w, h, shadows, ao, case = WORKLOADS[workload]
    golden = golden_case(case)
    host, depth = build_host_world(workload)
    pos, d, sun = vh.default_camera()
    look = tuple(int(v) for v in host.pick(pos, d, 30)[0])  # src/main.cpp:81: the host's picking ray feeds lookingAtBlock

    sc = fg.ShardedRayCaster(local_rank, rank, world_size, mode=args.frame_path, signal=args.signal)
    if args.schedule:
        sc.caster.set_option("schedule", args.schedule)
    info = sc.upload_world(host.describe())
    kw = dict(shadows=shadows, ao_samples=ao, looking_at=look)
    params = rc.frame_params(w, h, pos, d, sun, **kw)                 # the timed frame: what the reference's render() produces
    params_hits = rc.frame_params(w, h, pos, d, sun, hits=True, **kw)  # + parity records (free-flight rays work out where they end)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)  # > 126 MB L2
    stream = sc.stream  # librc's kernels, the NCCL gather, the flush and the timing events all go on this stream
    torch.cuda.set_stream(stream)
    n_side = min(args.steps, 20)  # frames of each side measurement (outside the timed K steps)
    checked = []  # (what, sha256) of every frame hashed against the golden

    def barrier():
        if world_size > 1:
            dist.barrier()
        torch.cuda.synchronize(dev)

    def max_over_ranks(x):
        if world_size == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def device_ms(render, n):
        """Mean device time of n frames (CUDA events on the launching stream, L2 flushed before each, outside the events)."""
        for _ in range(2):  # buffers this kind of frame needs (hit records ...) are allocated on first use: not in the timing
            render()
        torch.cuda.synchronize(dev)
        ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(n)]
        for i in range(n):
            flush.zero_()
            ev[i][0].record()
            render()
            ev[i][1].record()
        torch.cuda.synchronize(dev)
        return float(sum(a.elapsed_time(b) for a, b in ev)) / n

    # ---- N > 1: the SAME frame on ONE GPU first (rank 0, unsharded), so that the line carries a same-workload, same-box,
    # same-run strong-scaling figure; the other ranks wait.  Also the first parity check: that frame against the golden.
    strong = None
    if world_size > 1:
        t1 = t1_hits = 0.0
        if rank == 0:
            sc.caster.set_shard(0, 1)
            for _ in range(3):
                sc.caster.render_frame(params)
            t1 = device_ms(lambda: sc.caster.render_frame(params), n_side)
            checked.append(("one GPU, unsharded", sha256_of(sc.caster.read_image())))
            t1_hits = device_ms(lambda: sc.caster.render_frame(params_hits), min(n_side, 5))
            if golden:
                checked.append(("one GPU, unsharded, hit records", sha256_of(sc.caster.read_hits()), "hits_sha256"))
            sc.caster.set_shard(rank, world_size)
        barrier()
        strong = {"workload": workload, "t1_ms": t1, "t1_ms_with_hits": t1_hits,
                  "how": f"rank 0 rendered the unsharded frame alone ({n_side} frames, CUDA events, L2 flushed) before the sharded run"}

    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()

    # ---- warm-up ----
    for _ in range(max(args.warmup, 3)):
        flush.zero_()
        sc.render(params)
    barrier()
    st = sc.stats()
    local_rays = int(st.primary_rays + st.shadow_rays + st.ao_rays)
    schedule_used = int(st.schedule)
    launches_per_step = int(st.kernel_launches) + (1 if (world_size > 1 and rank == 0 and sc.mode == "gather") else 0)

    # ---- timed: K steps, device time by CUDA events on the launching stream, L2 flushed between steps ----
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
    barrier()
    t_begin = time.time()
    for i in range(args.steps):
        flush.zero_()
        ev[i][0].record()
        sc.render(params)
        ev[i][1].record()
    barrier()
    t_end = time.time()
    total_ms = max_over_ranks(float(sum(a.elapsed_time(b) for a, b in ev)))
    if world_size > 1:
        r = torch.tensor([local_rays], dtype=torch.int64, device=dev)
        dist.all_reduce(r, op=dist.ReduceOp.SUM)
        frame_rays = int(r.item())
    else:
        frame_rays = local_rays
    ms_per_step = total_ms / args.steps
    value = frame_rays / (ms_per_step * 1e-3) / 1e6

    # ---- side measurements, outside the timed steps -------------------------------------------------------------------
    # (a) the same frame WITH hit records: free-flight sky rays then work out the voxel they end in
    barrier()
    ms_hits = max_over_ranks(device_ms(lambda: sc.render(params_hits), n_side))
    if world_size == 1 and golden:
        checked.append(("hit records of the timed frame", sha256_of(sc.caster.read_hits()), "hits_sha256"))
    # (b) per-kernel device times: librc's own events around each kernel (RC_F_KERNEL_TIMES; the event between the two
    # launches is why the timed frames above do not ask for them); a fused schedule (--schedule A/B runs) is ONE kernel
    kern_ms = {}
    names = {v: k for k, v in rc.SCHEDULES.items()}
    tot = []
    for _ in range(n_side):
        flush.zero_()
        sc.render(params)
        tot.append(sc.stats().total_ms)  # synchronises
    kern_ms["frame"] = float(np.mean(tot))
    if schedule_used < rc.SCHED_FUSED:
        p_kt = rc.frame_params(w, h, pos, d, sun, kernel_times=True, **kw)
        p_kt_tail = rc.frame_params(w, h, pos, d, sun, kernel_times=True, always_tail=True, **kw)
        k_primary, k_ao, k_tail = [], [], []
        for _ in range(n_side):
            flush.zero_()
            sc.render(p_kt)
            s_ = sc.stats()
            k_primary.append(s_.primary_ms)
            k_ao.append(s_.ao_ms)
        for _ in range(n_side):
            flush.zero_()
            sc.render(p_kt_tail)
            k_tail.append(sc.stats().primary_ms)
        kern_ms["primary"] = float(np.mean(k_primary))
        kern_ms["ao"] = float(np.mean(k_ao))
        kern_ms["primary_with_free_flight_tail"] = float(np.mean(k_tail))
        sc.render(params)
    barrier()

    # ---- e2e: through the C-ABI with HOST buffers: params in, RGBA8 frame out into pinned memory, every step ----
    e2e = None
    if not args.no_e2e:
        # two pinned host frames: frame N+1 is traced while frame N travels (rc_read_image_async, N = 1)
        pinned = [torch.empty((h, w, 4), dtype=torch.uint8, pin_memory=True) for _ in range(2)] if rank == 0 else None
        host_imgs = [t.numpy() for t in pinned] if rank == 0 else None
        host_img = host_imgs[0] if rank == 0 else None

        def sync_step():
            if world_size == 1:
                sc.caster.render_frame(params)
                sc.caster.read_image(host_img)  # rc_read_image: D2H + stream sync
            else:
                sc.render(params)
                if rank == 0:
                    pinned[0].copy_(sc.frame(), non_blocking=True)
                torch.cuda.synchronize(dev)

        def timed(run):
            barrier()
            t0 = time.perf_counter()
            run()
            barrier()
            return max_over_ranks(time.perf_counter() - t0)

        def run_sync():
            for _ in range(args.steps):
                sync_step()

        def run_pipelined():
            # every step: params in (104 B), one full frame out into pinned memory; the frame of step i is
            # waited for (complete in host memory) during step i+1, the last one before the clock stops
            prev = None
            for i in range(args.steps):
                sc.caster.render_frame(params)
                t = sc.caster.read_image_async(host_imgs[i & 1])
                if prev is not None:
                    sc.caster.wait_image(prev)
                prev = t
            sc.caster.wait_image(prev)

        def run_pipelined_sharded():
            # N>1, direct frame: rank 0 copies frame i out of one frame slot while every rank traces frame i+1
            # into the other; the peers are held back by the per-frame completion signal
            prev = None
            for i in range(args.steps):
                sc.render(params)
                evt = sc.read_frame_async(pinned[i & 1]) if rank == 0 else None
                if prev is not None:
                    prev.synchronize()
                prev = evt
            if prev is not None:
                prev.synchronize()
            torch.cuda.synchronize(dev)

        sink = {"frames": None, "next": 0}
        copy_only = {}

        def run_host_direct():
            # N>1, no device-side gather at all: every rank reads ITS bands back into one shared, page-locked host
            # frame (POSIX shared memory mapped by all ranks; rc_read_image_async scatters a sharded frame's rows to
            # their place), all PCIe links at once; frame i travels while frame i+1 is traced.  Rank 0 consumes.
            sh = sink["frames"]
            prev = None

            def finish(ticket, i):
                sc.caster.wait_image(ticket)
                sh.mark_landed(i)
                if rank == 0:
                    sh.acquire(i)
                    sh.release(i)

            for _ in range(args.steps):
                i = sink["next"]
                sink["next"] += 1
                sc.caster.render_frame(params)
                sh.wait_slot_free(i)
                t = sc.caster.read_image_async(sh.frames[i % 2])
                if prev is not None:
                    finish(*prev)
                prev = (t, i)
            finish(*prev)

        for _ in range(3):
            sync_step()
        if rank == 0:
            checked.append(("rc_read_image, synchronous" if world_size == 1 else "rank 0's frame written by every rank's kernels (peer stores)"
                            if sc.mode == "direct" else "rank 0's frame after the NCCL gather + compose", sha256_of(host_img)))
        sync_s = timed(run_sync)
        rank0_s = None
        e2e_s, how = sync_s, "rc_render_frame(host params) + rc_read_image into pinned host memory, synchronous, every step"
        if world_size == 1:
            host_imgs[1][:] = 0
            run_pipelined()  # warm-up: creates the copy stream and the second device frame buffer
            e2e_s = timed(run_pipelined)
            how = ("rc_render_frame(host params) + rc_read_image_async into one of two pinned host frames + rc_wait_image, "
                   "every step; frame N+1 is traced while frame N is copied (double-buffered like the reference's swap chain)")
            checked.append(("rc_read_image_async, host frame 0", sha256_of(host_imgs[0])))
            checked.append(("rc_read_image_async, host frame 1", sha256_of(host_imgs[1])))
        elif sc.mode == "direct":
            run_pipelined_sharded()
            rank0_s = timed(run_pipelined_sharded)
            if rank == 0:
                checked.append(("peer-store frame, pipelined read-back, host frame 0", sha256_of(host_imgs[0])))
                checked.append(("peer-store frame, pipelined read-back, host frame 1", sha256_of(host_imgs[1])))
            try:
                sink["frames"] = fg.SharedHostFrames(w, h, rank, world_size, slots=2, pin=sc.caster)
            except RuntimeError as e:  # raised on every rank together (no room in /dev/shm)
                sink["frames"] = None
                e2e_s = rank0_s
                how = ("every rank: rc_render_frame(host params), kernels store into rank 0's frame over NVLink; rank 0: "
                       "asynchronous copy of the whole frame into one of two pinned host frames, every step, while all ranks "
                       f"trace the next frame (shared host frame unavailable: {e})")
            if sink["frames"] is not None:
                run_host_direct()
                e2e_s = timed(run_host_direct)
                # copy-only ceiling of this path: every rank copies ITS bands of the last frame into the shared host frame,
                # back to back, all ranks at once, with no rendering -- (a) the POSIX-shm frame registered in every process,
                # (b) a page-locked frame of each process's own (cudaHostAlloc): what the host's memory system and the PCIe
                # links can take, whatever the devices do
                reps = max(args.steps, 20)
                for key, dst in (("shared_posix_shm_frame", sink["frames"].frames[0]), ("own_cudaHostAlloc_frame", None)):
                    barrier()
                    t0c = time.perf_counter()
                    sc.caster.reread_image_ms(dst, reps)
                    barrier()
                    copy_only[key] = max_over_ranks(time.perf_counter() - t0c) / reps * 1e3
                if rank == 0:
                    for i, f in enumerate(sink["frames"].frames):
                        checked.append((f"frame assembled in shared host memory by every rank's own read-back, slot {i}", sha256_of(f)))
                sink["frames"].close()
                how = ("every rank: rc_render_frame(host params) + rc_read_image_async of ITS bands into one shared page-locked "
                       "host frame (POSIX shared memory, two frame slots), every step; rank 0 consumes each complete frame; no "
                       "device-side gather, all PCIe links in parallel")
        e2e = {"value": frame_rays * args.steps / e2e_s / 1e6, "unit": UNIT,
               "h2d_bytes_per_step": 104, "d2h_bytes_per_step": w * h * 4,
               "ms_per_step": e2e_s / args.steps * 1e3,
               "sync_ms_per_step": sync_s / args.steps * 1e3,
               **({"via_rank0_ms_per_step": rank0_s / args.steps * 1e3} if rank0_s is not None else {}),
               **({"copy_only_ms_per_frame": copy_only,
                   "copy_only_what": "every rank's band read-backs alone (no rendering), all ranks at once, back to back: the "
                                     "ceiling of the shared-host-frame path"} if copy_only else {}),
               "sync_value": frame_rays * args.steps / sync_s / 1e6,
               "what": how + ("; includes the NCCL band gather + compose on rank 0"
                              if world_size > 1 and sc.mode == "gather" else "")}
        if rank == 0:
            assert host_img.any(), "e2e frame is empty"
    elif world_size == 1:
        sc.caster.render_frame(params)
        checked.append(("rc_read_image, synchronous", sha256_of(sc.caster.read_image())))
    else:
        # every rank renders: a frame is a collective act (completion signal / gather); rank 0 alone would wait for ever
        sc.render(params)
        torch.cuda.synchronize(dev)
        barrier()
        if rank == 0:
            checked.append(("rank 0's composed frame", sha256_of(sc.frame().cpu().numpy())))

    clocks = sampler.stop(t_begin, t_end) if rank == 0 else None
    build_id = sc.caster.build_id()
    sync_timeouts = sc.caster.frame_sync_timeouts() if (rank == 0 and getattr(sc, "_sync_open", False)) else 0
    if rank != 0:
        sc.close()
        if world_size > 1:
            dist.destroy_process_group()
        return

    # ---- parity of exactly the frames that were timed / delivered, against the oracle's golden for this workload --------
    parity = None
    if golden:
        bad = [c[0] for c in checked if c[1] != golden[c[2] if len(c) > 2 else "rgba8_sha256"]]
        parity = {"golden": case, "file": golden["file"], "rgba8_sha256_match": not any(len(c) == 2 and c[0] in bad for c in checked),
                  "hits_sha256_match": not any(len(c) > 2 and c[0] in bad for c in checked) if any(len(c) > 2 for c in checked) else None,
                  "frames_checked": len(checked), "checked": [c[0] for c in checked], "mismatched": bad,
                  "what": "sha256 of each frame (and of one frame's hit records) this run rendered with the timed parameters, "
                          "against the oracle's frame for the same parameters; outside every timed region"}

    # ---- roofline of the dominant kernel ----
    peak, peak_src = hbm_peak()
    counters = golden["counters"] if golden else None
    roof, roof_all = None, {}
    records, capture_id = capture_record(workload)
    src_id = kernel_sources_sha256()
    # a capture belongs to the kernels that ran iff its digest of kernels.cu / kernels.h / free_flight.h is the one baked into
    # the librc.so this process loaded (rc_build_id)
    stale = (capture_id is None) or (capture_id != build_id)
    if counters:
        # one launch per rank renders 1/world_size of the frame's bands (round-robin: the shares are equal
        # to within a band), timed on rank 0
        ab = [b / world_size for b in counters["algorithmic_bytes"]]  # [primary, shadow, ao]
        fused = schedule_used >= rc.SCHED_FUSED
        per_kernel = {}
        if fused:
            per_kernel["primary_kernel"] = (sum(ab), kern_ms["frame"],
                                            "one fused launch: primary + shadow + AO rays (schedule %s)" % names[schedule_used])
        else:
            ao_name = "ao_window_persistent_kernel" if schedule_used == rc.SCHED_AO_PERSISTENT else "ao_window_kernel"
            per_kernel["primary_kernel"] = (ab[0] + ab[1], kern_ms["primary"], "primary + shadow rays, shading, AO seeds")
            if ao:
                per_kernel[ao_name] = (ab[2], kern_ms["ao"], "AO rays over the seed list (ao_steps = 5: dense-window kernel)")
        for name, (bytes_, ms, what) in per_kernel.items():
            if ms > 0:
                ach = bytes_ / (ms * 1e-3) / 1e9
                roof_all[name] = {"what": what, "algorithmic_bytes_per_launch": bytes_, "ms": ms, "achieved": ach, "frac": ach / peak}
        # what really bounds these kernels (profiles/*_ncu_full_summary.txt): instruction issue.  Issue-slot use of the
        # launches timed here = warp instructions of the launch (ncu capture) / (launch time x 4 schedulers x SMs x clock)
        sm_clock = (clocks or {}).get("sm_mhz")
        if sm_clock:
            slots_per_s = 4 * torch.cuda.get_device_properties(dev).multi_processor_count * sm_clock * 1e6
            for name, r in roof_all.items():
                rec = records.get(name)
                if rec and rec.get("warp_instructions_per_launch"):
                    inst = rec["warp_instructions_per_launch"] / world_size
                    r["issue_slots"] = {"warp_instructions_per_launch": inst, "peak_per_s": slots_per_s, "stale": stale,
                                        "capture": rec.get("source"), "ncu_issue_active_pct": rec.get("ncu_issue_active_pct")}
                    if not stale:
                        r["issue_slots"]["frac"] = inst / (r["ms"] * 1e-3) / slots_per_s
        dom = max(roof_all, key=lambda k: roof_all[k]["ms"]) if roof_all else None
        if dom:
            rec = records.get(dom, {})
            traffic = rec.get("dram_bytes_per_launch")
            # what the launch cannot avoid moving: the packed world once (upper bound on the part of it the rays touch), the
            # RGBA8 frame out, and -- two-kernel schedules only -- the AO seed list written and read back (48 B per hit pixel)
            seeds = 0 if fused else 48 * int(st.hit_pixels)  # written by the primary kernel, read by the AO kernel
            compulsory = int(info.packed_bytes) + seeds + (w * h * 4 // world_size if (fused or dom == "primary_kernel") else
                                                           4 * int(st.hit_pixels))
            roof = {"bound": "hbm", "kernel": dom, "schedule": names[schedule_used], "achieved": roof_all[dom]["achieved"], "peak": peak,
                    "unit": "GB/s", "frac": roof_all[dom]["frac"],
                    # the capture is of one GPU rendering the whole frame; a rank of N renders 1/N of its bands
                    "traffic": (traffic / world_size) if (traffic and not stale) else None, "traffic_stale": stale,
                    "compulsory_bytes": compulsory, "peak_source": peak_src,
                    "algorithmic_bytes_per_launch": roof_all[dom]["algorithmic_bytes_per_launch"],
                    "launch_ms": roof_all[dom]["ms"],
                    "issue_slots": roof_all[dom].get("issue_slots"),
                    "binding_resource": "instruction issue (issue_slots), not memory: the world is L1/L2 resident",
                    "note": "algorithmic bytes = what the REFERENCE algorithm fetches for this frame (16 B per nodes[] read + "
                            "4 B per child index, counted by the oracle); the kernel's region caching avoids most of them, so "
                            "frac exceeds 1 and says how far the design is from the reference's traffic, not how close to the "
                            "HBM limit; traffic = DRAM bytes of one launch from the ncu capture in profiles/ (null when that "
                            "capture is of other kernel sources than the librc.so that ran); compulsory_bytes = what this launch "
                            "cannot avoid moving: the packed world once (upper bound), the AO seed list (48 B per hit pixel, "
                            "two-kernel schedules) and the pixels it stores"}
    if strong is not None:
        strong["tN_ms"] = ms_per_step
        strong["n_gpus"] = world_size
        strong["efficiency"] = strong["t1_ms"] / (world_size * ms_per_step) if ms_per_step > 0 else None
        strong["tN_ms_with_hits"] = ms_hits
    line = {
        "metric": metric_name(workload, world_size), "value": value, "unit": UNIT, "n_gpus": world_size, "steps": args.steps,
        "warmup": max(args.warmup, 3),
        "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "strong" if world_size > 1 else "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "value_with_hits": frame_rays / (ms_hits * 1e-3) / 1e6, "ms_per_step_with_hits": ms_hits,
        "config": {"workload": workload, "resolution": [w, h], "shadows": shadows, "ao_samples": ao, "ao_steps": 5,
                   "max_steps": 300, "shadow_steps": 75,
                   "world": "default generated world (initTetraHexaTree + genWorld, seeds 42/64/100, 200x200 columns)"
                            if depth == 6 else "generated world 2048x2048 columns, tree depth 7 (extension)",
                   "world_bytes": {"reference_format": int(info.ref_bytes), "packed": int(info.packed_bytes)},
                   "camera": "default (35,50,35) dir normalize(1,0,1)", "rays_per_step": frame_rays,
                   "schedule": names[schedule_used],
                   "l2": "flushed between steps (256 MB memset, outside the timed events)",
                   "outputs": "the RGBA8 frame (what the reference's render() produces); no hit records, so sky rays in free "
                              "flight do not work out the voxel they would end in -- value_with_hits / ms_per_step_with_hits "
                              "are the same frame with the parity records written, as every parity test runs it",
                   "sharding": f"16-row bands round-robin over {world_size} GPUs, " + (
                       "kernels store into rank 0's frame over NVLink peer memory (CUDA IPC), " + sc.completion_signal
                       if sc.mode == "direct" else "NCCL gather to rank 0 + compose kernel")
                               if world_size > 1 else "none"},
        "e2e": e2e, "gpu_launches": launches_per_step * args.steps, "kernel_ms": kern_ms,
        "parity_check": parity, "roofline": roof, "roofline_kernels": roof_all, "clocks": clocks,
        "librc_build_id": build_id, "kernel_sources_sha256": src_id,
    }
    if strong is not None:
        line["strong_scaling"] = strong
    sc.close()
    if world_size > 1:
        dist.destroy_process_group()
    if world_size == 1 and not args.no_cpu_baseline:
        line["cpu_baseline"] = cpu_baseline_subprocess(workload)
    print(json.dumps(line), flush=True)
    if parity and parity["mismatched"]:
        raise SystemExit(f"bench.py: parity_check failed: {parity['mismatched']}")
    if sync_timeouts:
        raise SystemExit(f"bench.py: {sync_timeouts} frame-completion waits gave up (rc_frame_sync_timeouts): a rank never signalled")


def main():
    args = parse_args()
    workload = args.workload or ("config3_1080p_ao64" if args.gpus == 1 else "config5_8k_ao64")
    if args.cpu_baseline_only:
        print(json.dumps(cpu_baseline_leg(workload)), flush=True)
        return
    if args.impl == "reference":
        run_reference_arm(args, workload)
    else:
        run_b200_arm(args, workload)


if __name__ == "__main__":
    main()
