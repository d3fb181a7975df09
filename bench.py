#!/usr/bin/env python
"""bench.py -- cluster-light tests/s and light-culling ms/frame of the B200 path, against its roofline.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload NAME]

ONE headline workload at every N: C16M = BASELINE.json configs[3], 64x36x96 clusters x 16,777,216 lights at 3840x2160
(the only config the metric's "at 1/2/4/8 B200" applies to; it fits one GPU), "scaling": "strong".
One "step" = one frame of the hot path over the whole scene: AABB generation if dirty + light transform + culling +
offsets + scatter, the full light grid (lightCells + lightIndices in the reference layouts) left on EVERY GPU.
  N = 1   hvcf_cull on one context, all lights resident on the GPU.
  N > 1   hvcf_cull_routed: clusters sharded by z-slab AND lights owned in shards; every light is transformed every frame by
          its owner, tested against the slab boxes and routed into the owning ranks' inboxes over NVLink peer memory; totals,
          cells and index segments exchanged by kernels over peer memory.  No collective-library call inside the frame.
          Every rank hashes its full grid on the device and compares it with the UNSHARDED frame of rank 0
          ("sharded_bytes_equal"); a mismatch fails the run.

Keys of the JSON line (rank 0 prints exactly one line on stdout; it ENDS with a compact "summary"):
  value          nominal cluster x light tests per second, whole job, inputs resident in HBM
  e2e            same metric through the C ABI with HOST buffers: per step every rank uploads its own share of the LightData
                 (pinned H2D; N = 1: all of it) and the UBO, runs the frame, and downloads its own slab's cells + index
                 segment, so the whole job moves every light in and the whole grid out exactly once per step.  N = 1: with two
                 frames in flight (the reference's frame ring, AppConfig::FrameCount = 2: frame f+1's lights are uploaded into a
                 second device buffer while frame f is culled and read back; the bytes of the result are compared with the
                 one-at-a-time frames); the one-frame-at-a-time figure sits beside it under "one_frame_at_a_time"
  roofline       the dominant kernel of the step: its own tests x 16 FP32-pipe lane-ops / its duration (CUDA events inside
                 libhvcf, live) against the FP32 issue rate measured on this GPU
  cpu_baseline   the reference's own shader source compiled as C++ (oracle/_ref), all host threads, on a cluster sample
  summary        N = 1: ms / frac / e2e_ms of A1k, A65k, A1M, B1M and the fly-through, boundary-disagreement counts;
                 N > 1: phase split, per-rank kernel times, NVLink bytes, the NCCL-broadcast formulation beside the routed one

--impl reference times that CPU build alone (rank 0 only under torchrun).
Nothing here reads /root/reference at run time.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import tempfile
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

# name -> (grid, screen, lights)     
WORKLOADS = {
    "A1k": ((16, 9, 24), (2560, 1600), 1000),            # configs[0]
    "A65k": ((16, 9, 24), (2560, 1600), 65536),          # configs[1]
    "A1M": ((16, 9, 24), (2560, 1600), 1 << 20),         # north_star "1M lights x 16x9x24"
    "B1M": ((32, 18, 48), (3840, 2160), 1 << 20),        # configs[2]
    "C16M": ((64, 36, 96), (3840, 2160), 1 << 24),       # configs[3]  <- headline at every N
}
LANE_OPS_PER_TEST = 16  # 6 min/max + 3 sub + 3 mul + 3 add + 1 compare
L2_FLUSH_BYTES = 256 << 20


def log(*a):
    print(*a, file=sys.stderr, flush=True)


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="C16M", choices=sorted(WORKLOADS))
    ap.add_argument("--balance-iters", type=int, default=10, help="N > 1: slab re-cuts from measured per-rank kernel times")
    ap.add_argument("--block-tiles", type=int, default=0, help="tiles per route block (0 = library default)")
    ap.add_argument("--z-cuts", default="", help="N > 1: comma-separated z cuts (N + 1 numbers) instead of balancing")
    ap.add_argument("--row-cuts", default="", help="N > 1: comma-separated cuts in cluster rows (row = y + z * Y) instead of balancing")
    ap.add_argument("--no-summary", action="store_true", help="N = 1: skip the legs on the other workloads")
    ap.add_argument("--no-compare", action="store_true", help="N > 1: skip the NCCL-broadcast formulation")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-flythrough", action="store_true", help="skip the 1,024-frame animated fly-through leg (configs[4])")
    ap.add_argument("--fly-frames", type=int, default=0, help="frames of the fly-through (0 = 1,024)")
    ap.add_argument("--no-next-rows", action="store_true", help="skip the consumer / frustum-culling legs")
    ap.add_argument("--no-pipelined-e2e", action="store_true", help="N = 1: e2e as one frame at a time only (no second light buffer, no copy stream)")
    ap.add_argument("--cpu-seconds", type=float, default=8.0, help="wall-clock budget of the cpu_baseline sample")
    return ap.parse_args()


def host_threads():
    """Threads for the CPU legs: the cores this process may run on.  torch.distributed.run exports OMP_NUM_THREADS=1 to its
    workers; the reference arm must not inherit that (it would time ONE thread and call it the box)."""
    try:
        return max(1, len(os.sched_getaffinity(0)))
    except AttributeError:
        return max(1, os.cpu_count() or 1)


def pin_to_gpu_node(local):
    """N > 1: keep this rank's host thread (and with it the pinned buffers it first touches) on the CPU cores NVML reports as
    local to its GPU, so that eight ranks' H2D / D2H copies do not all cross the socket link.  Returns the number of cores kept,
    or None when NVML has nothing to say (then the affinity stays as it was)."""
    try:
        import pynvml
        pynvml.nvmlInit()
        h = pynvml.nvmlDeviceGetHandleByIndex(local)
        words = pynvml.nvmlDeviceGetCpuAffinity(h, ((os.cpu_count() or 64) + 63) // 64)
        cpus = {64 * i + b for i, w in enumerate(words) for b in range(64) if (int(w) >> b) & 1}
        want = cpus & os.sched_getaffinity(0)
        if want:
            os.sched_setaffinity(0, want)
            return len(want)
    except Exception:
        pass
    return None


# ------------------------------------------------------------------------------------------------
# clocks: nvidia-smi sampled in the background while the GPU loops run (B200_PROFILING.md)
# ------------------------------------------------------------------------------------------------
class ClockSampler:
    QUERY = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
             "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
             "clocks_event_reasons.sw_power_cap")
    NAMES = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]

    def __init__(self, device_index):
        self.samples = []  # (t, sm, smmax, power, [reasons])
        self.proc = None
        self.windows = []
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.QUERY}", "--format=csv,noheader,nounits",
                                          "-lms", "50", "-i", str(device_index)], stdout=subprocess.PIPE,
                                         stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            parts = [p.strip() for p in line.split(",")]
            if len(parts) < 7:
                continue
            try:
                sm, smmax, power = float(parts[0]), float(parts[1]), float(parts[2])
            except ValueError:
                continue
            reasons = [n for n, v in zip(self.NAMES, parts[3:7]) if v.lower().startswith("active")]
            self.samples.append((time.time(), sm, smmax, power, reasons))

    def window(self, t0, t1):
        self.windows.append((t0, t1))

    def stop(self):
        if self.proc:
            self.proc.terminate()
            try:
                self.proc.wait(timeout=2)
            except subprocess.TimeoutExpired:
                self.proc.kill()

    def summary(self):
        if not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0, "window": "none"}
        inwin = [s for s in self.samples if any(a <= s[0] <= b for a, b in self.windows)]
        where = "timed"
        if len(inwin) < 3:  # region shorter than the sampling period: use the loaded samples of the whole run
            pmax = max(s[3] for s in self.samples)
            inwin = [s for s in self.samples if s[3] >= 0.6 * pmax] or self.samples
            where = "run"
        reasons = sorted({r for s in inwin for r in s[4]})
        return {"sm_mhz": float(np.median([s[1] for s in inwin])), "sm_max_mhz": float(max(s[2] for s in inwin)),
                "power_w": float(np.median([s[3] for s in inwin])), "reasons": reasons, "samples": len(inwin), "window": where}


# ------------------------------------------------------------------------------------------------
# CPU side: the reference's own shader sources compiled as C++ (oracle/_ref) -- or, if that .so did
# not travel, the oracle port.  The only places bench.py touches oracle/.
# ------------------------------------------------------------------------------------------------
def cpu_checker():
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import cpu_checkers as cc
    if cc.Reference.available():
        ref = cc.Reference()
        return "reference", ref, host_threads()
    orc = cc.Oracle()
    return "port", orc, host_threads()


def cpu_frame_fn(kind, impl, grid, screen, n):
    """Returns (fn, tests_per_call, sample_description, threads): fn() culls one bounded sample on the host cores."""
    ubo = impl.camera_ubo(grid, screen)
    aabbs = impl.generate_aabbs(ubo)
    lights = impl.init_lights(n)
    C = len(aabbs)
    threads = host_threads()   # explicit: never OMP_NUM_THREADS
    full = C * n
    # bound one call to ~1e9 nominal tests (about a second on 16 cores): the whole frame for A1k / A65k, an evenly strided
    # cluster subset (every z-slice region represented) for the larger configs -- the figure is then an extrapolation
    stride = max(1, int(np.ceil(full / 1.0e9)))
    if stride == 1:
        sub, desc = aabbs, f"whole frame: {C} clusters x {n} lights"
    else:
        sub = np.ascontiguousarray(aabbs[::stride])
        desc = f"every {stride}th cluster ({len(sub)} of {C}) x all {n} lights; whole-frame figure extrapolated from the sample"
    if kind == "reference":
        # hvref_cull dispatches one workgroup per cluster of the grid in the UBO (gl_NumWorkGroups = the UBO's slice counts,
        # tile index = x + y X + z X Y, LightCulling.comp:61-64); a subset is run as a (len(sub), 1, 1) dispatch over the
        # gathered boxes -- the shader uses the counts only to form the tile index
        u = ubo.copy()
        if stride > 1:
            u.view(np.uint32)[38:41] = (len(sub), 1, 1)

        def fn():
            impl.cull(u, sub, lights, threads=threads)
    else:
        def fn():
            impl.cull(ubo, sub, lights, threads=threads)
    return fn, len(sub) * n, desc, threads


def time_cpu(kind, impl, workload, budget_s, min_calls=2, max_calls=64):
    grid, screen, n = WORKLOADS[workload]
    fn, tests, desc, threads = cpu_frame_fn(kind, impl, grid, screen, n)
    fn()  # warm the pages
    times = []
    t_begin = time.perf_counter()
    while len(times) < min_calls or (time.perf_counter() - t_begin < budget_s and len(times) < max_calls):
        t0 = time.perf_counter()
        fn()
        times.append(time.perf_counter() - t0)
    med = float(np.median(times))
    return {"value": tests / med, "unit": "tests/s", "cores": threads, "kind": kind,
            "sample": f"{desc}; {len(times)} calls, median {med * 1e3:.1f} ms/call", "ms_per_call": med * 1e3,
            "tests_per_call": tests}


def run_reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    kind, impl, threads = cpu_checker()
    grid, screen, n = WORKLOADS[args.workload]
    fn, tests, desc, threads = cpu_frame_fn(kind, impl, grid, screen, n)
    for _ in range(max(1, min(args.warmup, 3))):
        fn()
    # bound the whole run to a few minutes whatever K the driver passes
    steps, t_total, t_begin = 0, 0.0, time.perf_counter()
    while steps < args.steps and (steps < 3 or time.perf_counter() - t_begin < 150.0):
        t0 = time.perf_counter()
        fn()
        t_total += time.perf_counter() - t0
        steps += 1
    ms = t_total / steps * 1e3
    value = tests / (ms * 1e-3)
    full_frame_tests = grid[0] * grid[1] * grid[2] * n
    line = {
        "impl": "reference", "metric": "cluster-light tests/s", "value": value, "unit": "tests/s", "n_gpus": args.gpus,
        "steps": steps, "warmup": args.warmup, "ms_per_step": ms, "ms_per_frame_equiv": full_frame_tests / value * 1e3,
        "higher_is_better": True, "scaling": "strong",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": args.workload, "grid": list(grid), "screen": list(screen), "lights": n,
                   "light_distribution": "AppPBRClusterForward::InitLights, mt19937(12345)", "device": "host cpu"},
        "cpu_baseline": {"value": value, "unit": "tests/s", "cores": threads, "kind": kind, "sample": desc},
        "e2e": {"value": value, "unit": "tests/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
        "note": ("reference = HelloVulkan's AABBGenerator.comp/LightCulling.comp + Camera.cpp compiled as C++ with g++ "
                 "(oracle/_ref), OpenMP over workgroups; its Vulkan path is unavailable on this box"
                 if kind == "reference" else "oracle port (oracle/_ref not present)"),
    }
    print(json.dumps(line), flush=True)
    return 0


# ------------------------------------------------------------------------------------------------
# our arm
# ------------------------------------------------------------------------------------------------
def hbm_peak():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md)"


def ncu_traffic(workload):
    """dram bytes per launch of the dominant kernel from the committed ncu --set full capture, or None."""
    p = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(p):
        with open(p) as f:
            return json.load(f).get(workload)
    return None


KERNEL_NAMES = ("transform", "bin", "count", "finalize", "write")
KERNEL_SYMBOL = {"transform": "light_transform_kernel / shard_route_kernel (+ wait for the other ranks' shares)",
                 "bin": "hier_bin2_kernel / hier_bin_kernel / hier_xbin_kernel (level-0 binning)",
                 "count": "hier2_kernel<K,0> (count pass)", "finalize": "hier_finalize_kernel", "write": "hier2_kernel<K,1> (write pass)"}


def instrumented_kernels(hvcf, make_ctx, run_frame, sp, frames=6):
    """Per-kernel milliseconds of the hierarchical frame from CUDA events recorded inside libhvcf on the launching stream
    (HVCF_KERNEL_TIMES=1 at hvcf_create: an event after every kernel, which also switches their programmatic overlap off),
    and the work counters of the same frames.  make_ctx() -> context (created with the env switch on); run_frame(ctx, f).
    The caller closes the returned context."""
    os.environ["HVCF_KERNEL_TIMES"] = "1"
    try:
        ctx = make_ctx()
    finally:
        os.environ.pop("HVCF_KERNEL_TIMES", None)
    try:
        for f in range(2):
            run_frame(ctx, f & 1)
        rows = []
        for f in range(frames):
            run_frame(ctx, f & 1)
            rows.append(ctx.kernel_times(f & 1))
        st = ctx.stats((frames - 1) & 1, sp)
    except Exception:
        ctx.close()
        raise
    ms = dict(zip(KERNEL_NAMES, [float(x) for x in np.median(np.array(rows), axis=0)]))
    return ms, st, ctx


def kernel_roofline(ms, st, issue_rate, step_ms):
    """roofline object of the dominant kernel: ITS OWN tests x 16 lane-ops / ITS duration against the measured issue rate.
    bin kernel: the level-0 union tests (hvcf_stats.bin_tests); count pass: exact tests + level-1 union tests; the write
    pass: the tests it repeats (hvcf_stats.write_pass_tests); transform / finalize are HBM-side and carry no test count."""
    tests = {"bin": st["bin_tests"], "count": st["executed_tests"] + st["prefilter_tests"] - st["bin_tests"], "write": st["write_pass_tests"]}
    top = max(("bin", "count", "write"), key=lambda k: ms[k])
    total = sum(ms.values())
    ach = tests[top] * LANE_OPS_PER_TEST / (ms[top] * 1e-3)
    per_kernel = {k: {"ms": ms[k], "tests": int(tests[k]) if k in tests else None,
                      "frac": (tests[k] * LANE_OPS_PER_TEST / (ms[k] * 1e-3) / issue_rate) if k in tests and ms[k] > 0 else None}
                  for k in KERNEL_NAMES}
    return {"bound": "fp32", "kernel": KERNEL_SYMBOL[top], "achieved": ach / 1e12, "peak": issue_rate / 1e12, "unit": "TFLOP/s",
            "frac": ach / issue_rate, "traffic": None, "kernel_ms": ms[top], "kernel_share_of_step": ms[top] / total if total else None,
            "tests_per_launch": int(tests[top]), "flop_per_test": LANE_OPS_PER_TEST,
            "peak_source": "hvcf_measure_fp32_issue_rate on this GPU, this run (non-FMA FADD/FMUL issue rate; MEASURED_PEAKS.json has no "
                           "FP32 figure); nominal 148 SM x 128 lanes x 1.965 GHz = 37.2",
            "kernel_share_note": "share of the instrumented frame (an event after every kernel serialises the PDL chain); the ncu launch "
                                 "list under profiles/ must agree on the share, not on the absolute time",
            "per_kernel": per_kernel, "instrumented_step_ms": total, "step_ms": step_ms}


def timed_frames(torch, stream, run_frame, K, W, flush, barrier):
    """W warm-up frames, then K frames each bracketed by CUDA events on the launching stream (L2 flushed before each when
    `flush` is given); returns per-step ms."""
    for i in range(W):
        run_frame(i & 1)
    barrier()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(K)]
    for i in range(K):
        if flush is not None:
            flush.zero_()
        ev[i][0].record(stream)
        run_frame(i & 1)
        ev[i][1].record(stream)
    barrier()
    return np.array([a.elapsed_time(b) for a, b in ev])


def e2e_two_frames_in_flight(hvcf, torch, ctx, lib, stream, local, n, C, pinned, pinned_ubo, cells_host, idx_host, total_host, tp, Ke, We,
                             barrier, serial_step):
    """e2e with AppConfig::FrameCount = 2 frames in flight, as the reference app runs: while frame f is culled and its grid read
    back, the LightData of frame f + 1 already crosses PCIe into the OTHER of two device light buffers (hvcf_device_upload on a copy
    stream, hvcf_bind_lights per frame).  Every step still moves the same bytes in both directions inside the timed region, fill
    and drain of the pipeline included.  Checked against the one-at-a-time frames: totals every step, every byte of the last grid."""
    sp = stream.cuda_stream
    want = {}
    for f in range(2):   # the one-at-a-time result of both frame slots (two cameras)
        tot = serial_step(f, True)
        want[f] = (tot, cells_host.array[:8 * C].tobytes(), idx_host.array[:4 * tot].tobytes())
    copy_stream = torch.cuda.Stream(device=local)
    cp = copy_stream.cuda_stream
    bufs = [hvcf.DeviceBuffer(48 * n, local) for _ in range(2)]
    uploaded = [torch.cuda.Event() for _ in range(2)]
    consumed = [torch.cuda.Event() for _ in range(2)]
    try:
        def upload(i):
            copy_stream.wait_event(consumed[i & 1])      # the cull that read this buffer two frames ago is done
            hvcf._check(lib.hvcf_device_upload(bufs[i & 1].ptr, 0, pinned.ptr, 48 * n, cp))          # H2D 48*n
            uploaded[i & 1].record(copy_stream)

        def run(steps):
            bad = 0
            upload(0)
            for i in range(steps):
                f = i & 1
                if i + 1 < steps:
                    upload(i + 1)
                stream.wait_event(uploaded[f])
                ctx.bind_lights(bufs[f].ptr.value, n)
                ctx.set_ubo_ptr(f, pinned_ubo.ptr.value + f * hvcf.UBO_BYTES, sp)                     # H2D 176
                ctx.cull(f, sp)
                consumed[f].record(stream)
                hvcf._check(lib.hvcf_download_grid(ctx.handle, f, cells_host.ptr, 8 * C, idx_host.ptr, 4 * C * 150, tp, sp))
                bad += int(total_host[0]) != want[f][0]
            return bad

        run(max(We, 2))
        barrier()
        t0 = time.perf_counter()
        bad = run(Ke)
        torch.cuda.synchronize()
        ms = (time.perf_counter() - t0) / Ke * 1e3
        f = (Ke - 1) & 1
        tot = int(total_host[0])
        same = bad == 0 and cells_host.array[:8 * C].tobytes() == want[f][1] and idx_host.array[:4 * tot].tobytes() == want[f][2]
        if not same:
            raise RuntimeError(f"frames in flight differ from the one-at-a-time frames ({bad} totals of {Ke})")
        return {"ms_per_step": ms, "steps": Ke, "h2d_bytes_per_step": 48 * n + hvcf.UBO_BYTES, "d2h_bytes_per_step": 4 + 8 * C + 4 * tot,
                "bytes_equal_one_at_a_time": True,
                "mode": "two frames in flight (AppConfig::FrameCount): the LightData of frame f+1 is uploaded into the second of two device "
                        "light buffers while frame f is culled and downloaded; fill and drain inside the timed region"}
    finally:
        torch.cuda.synchronize()
        ctx.bind_lights(None, 0)
        for b in bufs:
            b.close()


def measure_single(hvcf, torch, name, K, W, local, flush_all, barrier=None, e2e_steps=50, graph=False, pipelined=False):
    """One workload on ONE GPU through the C ABI: device-timed frame, per-kernel times + roofline of the dominant kernel,
    e2e with host buffers (everything uploaded every step) and with the lights resident (what the reference app does per
    frame: UBO up, AppPBRClusterForward.cpp:120-136; here plus the grid down)."""
    import ctypes
    if barrier is None:
        barrier = torch.cuda.synchronize
    grid, screen, n = WORKLOADS[name]
    C = grid[0] * grid[1] * grid[2]
    stream = torch.cuda.current_stream()
    sp = stream.cuda_stream
    lib = hvcf.load()
    pinned = hvcf.PinnedBuffer(48 * n)
    lights = pinned.array.view(hvcf.LIGHT_DT)
    hvcf.host_init_lights(n, out=lights)
    ubos = [hvcf.host_camera_ubo(grid, screen, pos=hvcf.REF_CAMERA_POS), hvcf.host_camera_ubo(grid, screen, pos=(0.5, 1.0, 6.0))]
    pinned_ubo = hvcf.PinnedBuffer(2 * hvcf.UBO_BYTES)
    for f in range(2):
        pinned_ubo.array[f * hvcf.UBO_BYTES:(f + 1) * hvcf.UBO_BYTES] = ubos[f]
    big = 48 * n > (256 << 20)          # inputs larger than L2 (126 MB) by themselves: no flush needed
    flush = None if big else flush_all

    def make_ctx():
        c = hvcf.Context(grid=grid, max_lights=n, device=local)
        c.set_lights_ptr(pinned.ptr.value, n, False, sp)
        for f in range(2):
            c.set_ubo(f, ubos[f], sp)
        return c

    ctx = make_ctx()
    out = {"workload": name, "grid": list(grid), "lights": n}
    try:
        step_ms = timed_frames(torch, stream, lambda f: ctx.cull(f, sp), K, W, flush, barrier)
        st = ctx.stats((K - 1) & 1, sp)
        out.update({"ms": float(step_ms.mean()), "ms_min": float(step_ms.min()), "ms_median": float(np.median(step_ms)),
                    "launches_per_step": st["kernel_launches"], "total_indices": st["total_indices"],
                    "l2": "inputs larger than L2" if big else f"flushed between steps ({L2_FLUSH_BYTES >> 20} MiB memset, untimed)"})
        if graph:   # config 1 is launch-bound: the same frame captured once into a CUDA graph and replayed
            out["graph"] = measure_graph(hvcf, torch, ctx, K, W, flush)
        # ---- e2e through the C ABI with host buffers ----
        cells_host, idx_host = hvcf.PinnedBuffer(8 * C), hvcf.PinnedBuffer(4 * C * 150)
        total_host = np.zeros(1, np.uint32)
        tp = total_host.ctypes.data_as(ctypes.POINTER(ctypes.c_uint32))

        def e2e_step(f, upload_lights):
            if upload_lights:
                ctx.set_lights_ptr(pinned.ptr.value, n, False, sp)                  # H2D 48*n
            ctx.set_ubo_ptr(f, pinned_ubo.ptr.value + f * hvcf.UBO_BYTES, sp)        # H2D 176
            ctx.cull(f, sp)
            hvcf._check(lib.hvcf_download_grid(ctx.handle, f, cells_host.ptr, 8 * C, idx_host.ptr, 4 * C * 150, tp, sp))
            return int(total_host[0])

        Ke = max(3, min(K, e2e_steps))
        for upload, key in ((True, "e2e"), (False, "e2e_resident")):
            for i in range(min(W, 3)):
                e2e_step(i & 1, upload)
            barrier()
            t0 = time.perf_counter()
            for i in range(Ke):
                tot = e2e_step(i & 1, upload)
            torch.cuda.synchronize()
            ms = (time.perf_counter() - t0) / Ke * 1e3
            out[key] = {"ms_per_step": ms, "steps": Ke, "h2d_bytes_per_step": (48 * n if upload else 0) + hvcf.UBO_BYTES,
                        "d2h_bytes_per_step": 4 + 8 * C + 4 * tot}
        out["e2e"]["mode"] = "one frame at a time: upload lights + UBO, cull, hvcf_download_grid (blocking)"
        out["e2e_resident"]["mode"] = "lights resident (static lights, as ResourcesLight::AddLights once at init): upload UBO, cull, hvcf_download_grid"
        if pipelined:
            try:
                out["e2e_pipelined"] = e2e_two_frames_in_flight(hvcf, torch, ctx, lib, stream, local, n, C, pinned, pinned_ubo, cells_host,
                                                                idx_host, total_host, tp, Ke, min(W, 3), barrier, e2e_step)
            except Exception as ex:   # the one-at-a-time figure above stands
                log(f"e2e with two frames in flight failed, not reported: {ex!r}")
    finally:
        ctx.close()
    # ---- per-kernel times and the dominant kernel's roofline (a second, instrumented context) ----
    ms, st2, ictx = instrumented_kernels(hvcf, make_ctx, lambda c, f: ((flush.zero_() if flush is not None else None), c.cull(f, sp)), sp)
    ictx.close()
    issue_rate = hvcf.measure_fp32_issue_rate(local)
    out["roofline"] = kernel_roofline(ms, st2, issue_rate, out["ms"])
    if st2.get("plan_flags", 0) & 64:   # HVCF_PLAN_SINGLE_KERNEL: the frame was one launch; its time sits in the "count" slot
        out["roofline"]["kernel"] = "small_frame_kernel (hvcf_small.cu: the whole frame in one launch)"
    out["stats"] = {k: int(v) for k, v in st2.items()}
    out["nominal_tests"] = C * n
    return out


def measure_graph(hvcf, torch, ctx, K, W, flush):
    """The frame captured into a CUDA graph (torch.cuda.graphs on a side stream; libhvcf's launches are stream-ordered and
    capture-safe once the AABBs are clean) and replayed: launch-inclusive time per frame, against the kernel-only time
    (sum of the kernels' own durations would need ncu; reported is the graph replay bracketed by events)."""
    s = torch.cuda.Stream()
    g = torch.cuda.CUDAGraph()
    try:
        with torch.cuda.stream(s):
            ctx.cull(0, s.cuda_stream)     # warm: AABBs clean, kernels loaded
            s.synchronize()
            g.capture_begin()
            ctx.cull(0, s.cuda_stream)
            g.capture_end()
        for _ in range(W):
            g.replay()
        torch.cuda.synchronize()
        ms = []
        cur = torch.cuda.current_stream()
        for _ in range(K):
            if flush is not None:
                flush.zero_()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record(cur)
            g.replay()
            b.record(cur)
            b.synchronize()
            ms.append(a.elapsed_time(b))
        return {"ms": float(np.mean(ms)), "ms_min": float(min(ms)), "note": "one cudaGraphLaunch per frame, event pair around it"}
    except Exception as e:   # capture is best effort: the stream-launched figure stands
        return {"error": f"{type(e).__name__}: {e}"}


def grid_hash(torch, cells_t, idx_t):
    """64-bit position-dependent hash of the full light grid, computed on the device: lightCells (offset, count) and
    lightIndices[0, total) as int32 words.  Equal hashes on two GPUs <=> (to 2^-64) equal bytes."""
    def h(t, salt):
        x = t.reshape(-1).to(torch.int64)
        i = torch.arange(x.numel(), device=x.device, dtype=torch.int64)
        return ((x ^ salt) * (2 * i + 1)).sum()
    return int((h(cells_t, -7046029254386353131) * 31 + h(idx_t, 0x2545F4914F6CDD1D)).item())


def full_grid_hash(torch, hvcf, ctx, frame, dev, sp):
    from hellovulkan_b200.sharded import as_tensor
    b = ctx.buffers()
    total = ctx.download_total(frame, sp)
    cells = as_tensor(b.light_cells, 8 * b.cluster_count, dev).view(torch.int32)
    idx = as_tensor(b.light_indices, 4 * max(total, 1), dev).view(torch.int32)[:total]
    return grid_hash(torch, cells, idx), total


def routed_e2e_two_frames_in_flight(hvcf, torch, dist, ctx, stream, dev, local, pinned, lb, le, Cs, pinned_ubo, cells_host, idx_host, Ke,
                                    barrier, serial_step):
    """The routed frame's e2e with two frames in flight (see e2e_two_frames_in_flight): every rank uploads the NEXT frame's share of
    the LightData into the second of two device buffers of its owned lights (hvcf_bind_lights on a sharded context) while this
    frame is routed, culled and its own slab read back.  Every rank checks totals and, for the last frame, every byte against its
    one-at-a-time frames; the figure is reported only if every rank agrees (decided collectively, so no rank is left waiting)."""
    lib = hvcf.load()
    sp = stream.cuda_stream
    nown = le - lb
    res, bufs = None, []
    try:
        want = {}
        for f in range(2):
            base, tot = serial_step(f, True)
            want[f] = (base, tot, cells_host.array[:8 * Cs].tobytes(), idx_host.array[:4 * tot].tobytes())
        copy_stream = torch.cuda.Stream(device=local)
        cp = copy_stream.cuda_stream
        bufs = [hvcf.DeviceBuffer(48 * max(nown, 1), local) for _ in range(2)]
        uploaded = [torch.cuda.Event() for _ in range(2)]
        consumed = [torch.cuda.Event() for _ in range(2)]

        def upload(i):
            copy_stream.wait_event(consumed[i & 1])
            if nown:
                hvcf._check(lib.hvcf_device_upload(bufs[i & 1].ptr, 0, pinned.ptr, 48 * nown, cp))
            uploaded[i & 1].record(copy_stream)

        def run(steps):
            bad, last = 0, (0, 0)
            upload(0)
            for i in range(steps):
                f = i & 1
                if i + 1 < steps:
                    upload(i + 1)
                stream.wait_event(uploaded[f])
                if nown:
                    ctx.bind_lights(bufs[f].ptr.value, nown)
                ctx.set_ubo_ptr(f, pinned_ubo.ptr.value + f * hvcf.UBO_BYTES, sp)
                ctx.cull_routed(f, sp)
                consumed[f].record(stream)
                last = ctx.download_shard_into(cells_host.ptr, 8 * Cs, idx_host.ptr, 4 * Cs * 150, sp)
                bad += last != want[f][:2]
            return bad, last

        run(2)
        barrier()
        t0 = time.perf_counter()
        bad, (_, tot) = run(Ke)
        torch.cuda.synchronize()
        dt = time.perf_counter() - t0
        f = (Ke - 1) & 1
        if bad or cells_host.array[:8 * Cs].tobytes() != want[f][2] or idx_host.array[:4 * tot].tobytes() != want[f][3]:
            raise RuntimeError(f"frames in flight differ from the one-at-a-time frames ({bad} totals of {Ke})")
        res = (dt, float(48 * nown + hvcf.UBO_BYTES), float(8 * Cs + 4 * tot + 8))
    except Exception as ex:
        log(f"routed e2e with two frames in flight failed on this rank, not reported: {ex!r}")
    finally:
        torch.cuda.synchronize()
        try:
            ctx.bind_lights(None, 0)
        except Exception:
            pass
        for b in bufs:
            b.close()
    t = torch.tensor([1.0 if res else 0.0, -(res[0] if res else 0.0)], device=dev, dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MIN)          # everybody fine?  the slowest rank's time
    v = torch.tensor(list(res[1:]) if res else [0.0, 0.0], device=dev, dtype=torch.float64)
    dist.all_reduce(v, op=dist.ReduceOp.SUM)
    if t[0].item() < 1.0:
        return None
    return {"ms_per_step": -float(t[1].item()) / Ke * 1e3, "steps": Ke, "h2d_bytes_per_step": int(v[0].item()), "d2h_bytes_per_step": int(v[1].item()),
            "bytes_equal_one_at_a_time": True,
            "mode": "two frames in flight (AppConfig::FrameCount): every rank uploads its share of frame f+1's LightData into the second of two "
                    "bound device buffers while frame f is routed, culled and its own slab downloaded; fill and drain inside the timed region"}


def run_routed(args, hvcf, torch, dist, world, rank, local, dev, barrier, K, W, flush):
    """The headline at N > 1: hvcf_cull_routed (clusters sharded by z-slab, lights owned in shards)."""
    import ctypes
    from hellovulkan_b200.sharded import slice_touch_counts, cuts_from_cost, CutSearch
    grid, screen, n = WORKLOADS[args.workload]
    C = grid[0] * grid[1] * grid[2]
    stream = torch.cuda.current_stream()
    sp = stream.cuda_stream
    all_lights = hvcf.host_init_lights(n)
    ubos = [hvcf.host_camera_ubo(grid, screen, pos=hvcf.REF_CAMERA_POS), hvcf.host_camera_ubo(grid, screen, pos=(0.5, 1.0, 6.0))]
    pinned_ubo = hvcf.PinnedBuffer(2 * hvcf.UBO_BYTES)
    for f in range(2):
        pinned_ubo.array[f * hvcf.UBO_BYTES:(f + 1) * hvcf.UBO_BYTES] = ubos[f]

    def allgather(x):
        out = [None] * world
        dist.all_gather_object(out, x)
        return out

    state = {}

    def make(cuts, ktimes=False):
        if ktimes:
            os.environ["HVCF_KERNEL_TIMES"] = "1"
        try:
            ctx = hvcf.Context(grid=grid, max_lights=n, device=local, shard=(rank, world, n, cuts, args.block_tiles))
        finally:
            os.environ.pop("HVCF_KERNEL_TIMES", None)
        b, e = ctx.shard_light_range()
        if state.get("range") != (b, e):
            state["pinned"] = hvcf.PinnedBuffer(48 * max(e - b, 1))
            state["pinned"].array[:48 * (e - b)] = all_lights[b:e].view(np.uint8).reshape(-1)
            state["range"] = (b, e)
        ctx.set_lights_range_ptr(state["pinned"].ptr.value, b, e - b, False, sp)
        for f in range(2):
            ctx.set_ubo(f, ubos[f], sp)
        ctx.peer_attach(rank, world, allgather(ctx.peer_export()))
        barrier()
        return ctx

    def drop(ctx):
        barrier()
        ctx.peer_detach()
        barrier()
        ctx.close()

    # ---- cuts: in units of 4 cluster rows (the height of a 32-cluster block; row = y + z * Y), i.e. finer than whole z-slices.
    # A host cost proxy first (lights reaching each slice), then re-cuts from measured per-rank kernel times ----
    history = []
    X, Y, Z = grid
    unit = 4 if Y % 4 == 0 else Y                  # rows per unit
    upz = Y // unit                                # units per slice
    if args.z_cuts:
        cuts_u = [int(x) * upz for x in args.z_cuts.split(",")]
    elif args.row_cuts:
        cuts_u = [int(x) // unit for x in args.row_cuts.split(",")]
    else:
        touch = slice_touch_counts(ubos[0], all_lights, stride=16).astype(np.float64)
        cost_u = np.repeat((touch + 0.05 * touch.mean()) / upz, upz)
        cuts_u = cuts_from_cost(cost_u, world)
        search = CutSearch(Z * upz, world)
        for it in range(max(0, args.balance_iters)):
            ctx = make([c * unit for c in cuts_u], ktimes=True)
            for f in range(2):
                ctx.cull_routed(f & 1, sp)
            barrier()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(stream)
            for f in range(4):
                ctx.cull_routed(f & 1, sp)
            e1.record(stream)
            torch.cuda.synchronize()
            kt = ctx.kernel_times(1)
            err = ctx.peer_status(sp)
            drop(ctx)
            own = float(sum(kt[1:]))                       # bin + count + finalize + write: this rank's own work, no waiting
            got = allgather((own, e0.elapsed_time(e1) / 4))
            times, frame_ms = [g[0] for g in got], max(g[1] for g in got)
            history.append({"row_cuts": [c * unit for c in cuts_u], "own_kernels_ms": [round(t, 4) for t in times], "frame_ms": round(frame_ms, 4),
                            "peer_error": err})
            # the estimate of the cost over the rows comes from the per-rank kernel times; the cuts that are KEPT are the ones with the
            # shortest measured frame (ranks also wait for each other at the two phase boundaries of a frame)
            search.add(cuts_u, times, score=frame_ms)
            new_cuts = search.next_cuts()
            if new_cuts == cuts_u:
                break
            cuts_u = new_cuts
        if search.best is not None:
            cuts_u = search.best[1]
    cuts = [c * unit for c in cuts_u]              # in cluster rows
    ctx = make(cuts)
    out = {"row_cuts": list(cuts), "rows_per_slice": Y, "balance_iterations": history}

    # ---- value: lights resident on their owners, every light transformed every frame; per-step events; max over ranks ----
    step_ms = timed_frames(torch, stream, lambda f: ctx.cull_routed(f, sp), K, W, flush, barrier)
    out["peer_error"] = ctx.peer_status(sp)
    t = torch.tensor([float(step_ms.sum())], device=dev, dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    out["ms"] = float(t.item()) / K
    st = ctx.stats((K - 1) & 1, sp)
    out["launches_per_step"] = st["kernel_launches"] + 4   # + totals, push, wait, (route/wait are counted inside)
    out["inbox_records_per_rank"] = allgather(ctx.shard_inbox_records(sp))
    out["light_range_per_rank"] = allgather(list(state["range"]))
    total = ctx.download_total((K - 1) & 1, sp)
    out["total_indices"] = int(total)
    b = ctx.buffers()
    out["own_clusters"] = [b.cluster_begin, b.cluster_end]

    # ---- byte equality with the UNSHARDED frame (rank 0 culls the whole scene on one context) ----
    f_last = (K - 1) & 1
    mine, _ = full_grid_hash(torch, hvcf, ctx, f_last, dev, sp)
    ref_hash = None
    if rank == 0:
        with hvcf.Context(grid=grid, max_lights=n, device=local) as u:
            u.set_lights(all_lights, sp)
            u.set_ubo(f_last, ubos[f_last], sp)
            u.cull(f_last, sp)
            ref_hash, ref_total = full_grid_hash(torch, hvcf, u, f_last, dev, sp)
            out["unsharded_total_indices"] = int(ref_total)
    hashes = allgather(mine)
    ref_hash = allgather(ref_hash)[0]
    out["sharded_bytes_equal"] = bool(all(h == ref_hash for h in hashes))
    out["grid_hash"] = f"{ref_hash & 0xffffffffffffffff:016x}"
    if not out["sharded_bytes_equal"]:
        # failure reproduction: the scene is seeded, so the cuts, the route block size and the library's switches replay the frame
        out["repro"] = {"ranks_differing": [r for r, h in enumerate(hashes) if h != ref_hash],
                        "command": f"python -m torch.distributed.run --nnodes=1 --nproc-per-node {world} --master-addr 127.0.0.1 bench.py --gpus {world} "
                                   f"--workload {args.workload} --steps {K} --warmup {W} --row-cuts {','.join(str(c) for c in cuts)} "
                                   f"--block-tiles {args.block_tiles} --no-compare --no-flythrough --no-cpu-baseline",
                        "env": {k: v for k, v in os.environ.items() if k.startswith("HVCF_")}}
        log("SHARDED GRID DIFFERS FROM THE UNSHARDED FRAME; to replay:", json.dumps(out["repro"]))

    # ---- e2e: HOST buffers; every rank uploads its own lights + UBO and downloads its own slab's cells + index segment ----
    lb, le = state["range"]
    Cs = b.cluster_end - b.cluster_begin
    cells_host, idx_host = hvcf.PinnedBuffer(8 * max(Cs, 1)), hvcf.PinnedBuffer(4 * max(Cs, 1) * 150)

    def e2e_step(f, upload):
        if upload:
            ctx.set_lights_range_ptr(state["pinned"].ptr.value, lb, le - lb, False, sp)
        ctx.set_ubo_ptr(f, pinned_ubo.ptr.value + f * hvcf.UBO_BYTES, sp)
        ctx.cull_routed(f, sp)
        return ctx.download_shard_into(cells_host.ptr, 8 * Cs, idx_host.ptr, 4 * Cs * 150, sp)

    Ke = max(3, min(K, 20))
    for upload, key in ((True, "e2e"), (False, "e2e_resident")):
        for i in range(2):
            e2e_step(i & 1, upload)
        barrier()
        t0 = time.perf_counter()
        for i in range(Ke):
            _, tot = e2e_step(i & 1, upload)
        torch.cuda.synchronize()
        t = torch.tensor([time.perf_counter() - t0], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        h2d = torch.tensor([float((48 * (le - lb) if upload else 0) + hvcf.UBO_BYTES), float(8 * Cs + 4 * tot + 8)], device=dev, dtype=torch.float64)
        dist.all_reduce(h2d, op=dist.ReduceOp.SUM)
        out[key] = {"ms_per_step": float(t.item()) / Ke * 1e3, "steps": Ke, "h2d_bytes_per_step": int(h2d[0].item()),
                    "d2h_bytes_per_step": int(h2d[1].item())}
    if not args.no_pipelined_e2e:
        r = routed_e2e_two_frames_in_flight(hvcf, torch, dist, ctx, stream, dev, local, state["pinned"], lb, le, Cs, pinned_ubo, cells_host,
                                            idx_host, Ke, barrier, e2e_step)
        if r is not None:
            out["e2e_pipelined"] = r
    drop(ctx)

    # ---- per-kernel times on every rank (instrumented contexts) and the dominant kernel's roofline ----
    ictx = make(cuts, ktimes=True)
    for f in range(2):
        ictx.cull_routed(f & 1, sp)
    rows = []
    for f in range(6):
        flush.zero_()
        ictx.cull_routed(f & 1, sp)
        rows.append(ictx.kernel_times(f & 1))
    route_ms = ictx.route_ms(1)
    st2 = ictx.stats(1, sp)
    drop(ictx)
    ms = dict(zip(KERNEL_NAMES, [float(x) for x in np.median(np.array(rows), axis=0)]))
    issue_rate = hvcf.measure_fp32_issue_rate(local)
    roof = kernel_roofline(ms, st2, issue_rate, out["ms"])
    per_rank = allgather({"ms": ms, "roofline": roof, "stats": {k: int(v) for k, v in st2.items()}})
    own = [sum(p["ms"][k] for k in ("bin", "count", "finalize", "write")) for p in per_rank]
    crit = int(np.argmax(own))
    out["roofline"] = dict(per_rank[crit]["roofline"], rank=crit)
    out["own_kernels_ms_per_rank"] = [round(x, 4) for x in own]
    out["route_kernel_ms_per_rank"] = [round(x, 4) for x in allgather(route_ms)]
    out["kernel_ms_per_rank"] = [{k: round(v, 4) for k, v in p["ms"].items()} for p in per_rank]
    out["executed_tests_all_ranks"] = int(sum(p["stats"]["executed_tests"] + p["stats"]["prefilter_tests"] for p in per_rank))
    return out


def run_next_rows(args, hvcf, ctx, grid, screen, n, local, stream, sp, flush):
    """The rows next to the path, timed on the device with the culled frame of the workload resident:
    consumer = the Scene.frag light loop over a synthetic G-buffer of the workload's screen (hvcf_shade);
    frustum  = FrustumCulling.comp over 10,000 boxes (the reference README's own figure: < 25 us) and over 2^20."""
    import torch
    out = {}
    W, H = screen
    P = W * H
    g = hvcf.host_init_gbuffer(W, H, seed=7)
    dg = hvcf.DeviceGBuffer(g, device=local)
    try:
        ctx.cull(0, sp)
        for _ in range(3):
            ctx.shade(0, dg, stream=sp)
        torch.cuda.synchronize()
        reps, ms = 10, []
        for _ in range(reps):
            flush.zero_()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record(stream)
            ctx.shade(0, dg, stream=sp)
            b.record(stream)
            b.synchronize()
            ms.append(a.elapsed_time(b))
        cluster = dg.out_cluster.download(np.uint32)
        cells = ctx.download_light_cells()
        evals = int(cells["count"][cluster].astype(np.int64).sum())
        t = float(np.median(ms))
        gb = (4 + 48 + 16 + 4) * P  # depth + 3 float4 in, float4 + cluster index out
        hbm_gbs, _ = hbm_peak()
        out["consumer"] = {"kernel": "cluster_shade_kernel", "screen": [W, H], "pixels": P, "ms_per_frame": t, "ms_min": float(min(ms)),
                           "light_evaluations_per_frame": evals, "light_evaluations_per_s": evals / (t * 1e-3),
                           "mean_list_length_per_pixel": evals / P, "gbuffer_bytes_per_frame": gb,
                           "gbuffer_gbs": gb / (t * 1e-3) / 1e9, "hbm_frac": gb / (t * 1e-3) / 1e9 / hbm_gbs,
                           "note": "Scene.frag:55-78,103-134 + Radiance.glsl over a device G-buffer; ambient/emissive (textures) out of scope"}
    finally:
        dg.close()
    fr = {}
    frustum = hvcf.host_frustum_ubo(screen)
    for count in (10000, 1 << 20, 1 << 24):   # 2^24 boxes = 604 MB: larger than L2, the asymptotic bandwidth of the kernel
        boxes = hvcf.host_init_boxes(count, 60.0)
        db = hvcf.DeviceBuffer.from_array(boxes, device=local)
        dc = hvcf.DeviceBuffer(16 * count, device=local)
        try:
            for _ in range(3):
                hvcf.frustum_cull(frustum, db.ptr, count, dc.ptr, device=local, stream=sp)
            torch.cuda.synchronize()
            us = []
            for _ in range(20):
                flush.zero_()
                a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                a.record(stream)
                hvcf.frustum_cull(frustum, db.ptr, count, dc.ptr, device=local, stream=sp)
                b.record(stream)
                b.synchronize()
                us.append(a.elapsed_time(b) * 1e3)
            vis = int(dc.download(hvcf.CMD_DT, count)["instance_count"].sum())
            t = float(np.median(us))
            bytes_ = (32 + 4) * count
            fr[str(count)] = {"us": t, "us_min": float(min(us)), "visible": vis, "boxes_per_s": count / (t * 1e-6),
                              "algorithmic_bytes": bytes_, "gbs": bytes_ / (t * 1e-6) / 1e9,
                              "hbm_frac": bytes_ / (t * 1e-6) / 1e9 / hbm_peak()[0]}
        finally:
            db.close(), dc.close()
    fr["reference_published_us_10000"] = 25.0
    fr["vs_published_10000"] = 25.0 / fr["10000"]["us"]
    fr["note"] = ("FrustumCulling.comp, one event pair around a single launch (launch latency included); the reference README "
                  "quotes < 25 us for 10,000 AABBs on unnamed hardware (README.md:63)")
    out["frustum"] = fr
    return out


def run_extras(args, hvcf, local, stream, sp, flush):
    """Two figures SURVEY.md 8(d) asks for beside the headline:
    aabb_generation  AABBGenerator.comp alone, in microseconds, grids A / B / C (runs only when the projection changes);
    stress           the no-early-exit variant: grid A x 2^20 lights (north_star's target shape) with the cap raised from
                     150 to 4,096 so that executed ~ nominal tests, brute-force mode -> the cull kernel's fraction of the
                     measured FP32 issue rate when early exit cannot help."""
    import torch
    out = {"aabb_generation": {}}
    for name, (grid, screen) in {"A": ((16, 9, 24), (2560, 1600)), "B": ((32, 18, 48), (3840, 2160)), "C": ((64, 36, 96), (3840, 2160))}.items():
        with hvcf.Context(grid=grid, max_lights=16, device=local, mode=hvcf.MODE_BRUTE) as c:
            c.set_ubo(0, hvcf.host_camera_ubo(grid, screen), sp)
            us = []
            for i in range(13):
                flush.zero_()
                c.mark_aabb_dirty()
                a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                a.record(stream)
                c.generate_aabbs(0, sp)
                b.record(stream)
                b.synchronize()
                if i >= 3:
                    us.append(a.elapsed_time(b) * 1e3)
            C = grid[0] * grid[1] * grid[2]
            out["aabb_generation"][name] = {"clusters": C, "us": float(np.median(us)), "us_min": float(min(us)), "bytes_out": 32 * C,
                                            "note": "slice-plane table upload (4 B x (Z+1)) + one kernel, event pair around both"}
    grid, screen, n = WORKLOADS["A1M"]
    cap = 4096
    C = grid[0] * grid[1] * grid[2]
    with hvcf.Context(grid=grid, max_lights=n, max_lights_per_cluster=cap, device=local, mode=hvcf.MODE_BRUTE) as c:
        c.set_lights(hvcf.host_init_lights(n), sp)
        c.set_ubo(0, hvcf.host_camera_ubo(grid, screen), sp)
        for _ in range(2):
            c.cull(0, sp)
        c.set_timing(True)
        k_ms, t_ms = [], []
        for _ in range(5):
            flush.zero_()
            c.cull(0, sp)
            km, tm = c.timing(0)
            k_ms.append(km), t_ms.append(tm)
        st = c.stats(0, sp)
        c.set_timing(False)
    rate = hvcf.measure_fp32_issue_rate(local)
    km = float(np.median(k_ms))
    out["stress_no_early_exit"] = {"workload": "A1M, max_lights_per_cluster 4096, mode brute", "nominal_tests": C * n,
                                   "executed_tests": st["executed_tests"], "executed_over_nominal": st["executed_tests"] / (C * n),
                                   "total_indices": st["total_indices"], "overflow_clusters": st["overflow_clusters"],
                                   "cull_kernel_ms": km, "step_ms": float(np.median(t_ms)),
                                   "achieved_tlaneops": st["executed_tests"] * LANE_OPS_PER_TEST / (km * 1e-3) / 1e12,
                                   "peak_tlaneops": rate / 1e12, "frac": st["executed_tests"] * LANE_OPS_PER_TEST / (km * 1e-3) / rate}
    return out


def run_flythrough(args, hvcf, world, rank, local, dev, barrier):
    """BASELINE.json configs[4]: 1,024-frame camera orbit over grid B with 2^20 animated lights,
    frames partitioned in contiguous blocks across the ranks (replicas: no collective inside the loop).  End to end per
    frame: host camera -> UBO (CPU), pinned H2D of the UBO, device-side light animation fused into the transform, culling.
    The lights are uploaded once (rank 0 -> NCCL broadcast when N > 1)."""
    import torch
    import torch.distributed as dist
    from hellovulkan_b200.sharded import as_tensor

    grid, screen, n = WORKLOADS["B1M"]
    C = grid[0] * grid[1] * grid[2]
    frames = hvcf.FLY_FRAMES if args.fly_frames <= 0 else args.fly_frames
    per = (frames + world - 1) // world
    f0, f1 = min(rank * per, frames), min((rank + 1) * per, frames)
    ctx = hvcf.Context(grid=grid, max_lights=n, device=local)
    stream = torch.cuda.current_stream()
    sp = stream.cuda_stream
    b = ctx.buffers()
    lights_t = as_tensor(b.lights, 48 * n, dev)
    if rank == 0:
        host = np.zeros(n, hvcf.LIGHT_DT)
        hvcf.host_init_lights(n, out=host)
        lights_t.copy_(torch.from_numpy(host.view(np.uint8).reshape(-1)))
    if world > 1:
        dist.broadcast(lights_t, src=0)
    ctx.bind_lights(b.lights, n)  # the bytes already sit in the context's own buffer: only the count is new
    ctx.set_light_animation(hvcf.fly_table(), hvcf.FLY_AMPLITUDE, sp)
    ubos = [hvcf.host_camera_ubo(grid, screen, pos=hvcf.fly_camera_pos(f)) for f in range(f0, min(f0 + 4, f1))]
    for k, u in enumerate(ubos):  # warm-up
        ctx.set_ubo(k & 1, u, sp)
        ctx.set_animation_step(k & 1, hvcf.fly_step(f0 + k))
        ctx.cull(k & 1, sp)
    barrier()
    t0 = time.perf_counter()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    host_s = 0.0
    for f in range(f0, f1):
        slot = f & 1
        h0 = time.perf_counter()
        ubo = hvcf.host_camera_ubo(grid, screen, pos=hvcf.fly_camera_pos(f))   # CPU: Camera -> ClusterForwardUBO
        host_s += time.perf_counter() - h0
        ctx.set_ubo(slot, ubo, sp)
        ctx.set_animation_step(slot, hvcf.fly_step(f))
        ctx.cull(slot, sp)
    e1.record(stream)
    torch.cuda.synchronize()
    wall = time.perf_counter() - t0
    last_total = ctx.download_total((f1 - 1) & 1) if f1 > f0 else 0
    st = ctx.stats((f1 - 1) & 1) if f1 > f0 else {"kernel_launches": 0}
    t = torch.tensor([wall, e0.elapsed_time(e1) * 1e-3], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ctx.close()
    wall_max, dev_max = float(t[0].item()), float(t[1].item())
    return {"workload": "B1M fly-through", "grid": list(grid), "lights": n, "frames": frames, "frames_per_rank": per,
            "ms_per_frame": wall_max / frames * 1e3, "ms_per_frame_per_gpu": wall_max / max(1, f1 - f0) * 1e3,
            "device_ms_per_frame_per_gpu": dev_max / max(1, f1 - f0) * 1e3,
            "host_camera_ms_per_frame": host_s / max(1, f1 - f0) * 1e3, "value": frames * C * n / wall_max,
            "unit": "tests/s (nominal)", "scaling": "strong", "last_frame_total_indices": int(last_total),
            "kernel_launches_per_frame": st["kernel_launches"], "h2d_bytes_per_frame": hvcf.UBO_BYTES,
            "timer": "host perf_counter around the rank's frame loop, synchronised both sides, max over ranks",
            "note": "camera orbit + lights y = y0 + 0.5 sin(4 theta_f + phase_i) evaluated on the device from a 1,024-entry host sine table"}


def run_broadcast_formulation(args, hvcf, torch, dist, world, rank, local, dev, barrier, cuts, f_last, steps=3):
    """north_star's own formulation beside the routed frame, on the same z cuts: the light buffer lives on rank 0 and is
    broadcast with NCCL every frame (48 B LightData records, cut into chunks and pipelined with transform + level-0
    binning), every rank transforms and bins ALL the lights, the grid is exchanged by the peer kernels (hvcf_cull_sharded's
    steps).  Timed on the device, max over ranks; its full grid is hashed against the routed frame's."""
    from hellovulkan_b200.sharded import HvcfEngine, ShardedCuller
    grid, screen, n = WORKLOADS[args.workload]
    stream = torch.cuda.current_stream()
    sp = stream.cuda_stream
    slabs = [(cuts[r], cuts[r + 1]) for r in range(world)]          # in cluster rows
    eng = HvcfEngine(hvcf, grid, n, None, local, rows=slabs[rank])
    sc = ShardedCuller(eng, grid, world=world, rank=rank, slabs=[(0, grid[2])] * world, exchange="peer")
    ubos = [hvcf.host_camera_ubo(grid, screen, pos=hvcf.REF_CAMERA_POS), hvcf.host_camera_ubo(grid, screen, pos=(0.5, 1.0, 6.0))]
    if rank == 0:
        host = hvcf.host_init_lights(n)
        eng.lights.copy_(torch.from_numpy(host.view(np.uint8).reshape(-1)))
        del host
    for f in range(2):
        eng.set_ubo(f, ubos[f])
    res = {}
    try:
        for i in range(2):
            sc.frame_pipelined(i & 1, n, chunks=4, wire="raw")
        barrier()
        sc.enable_phase_timing()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        for i in range(steps):
            sc.frame_pipelined(i & 1, n, chunks=4, wire="raw")
        e1.record(stream)
        barrier()
        t = torch.tensor([e0.elapsed_time(e1)], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        phases = sc.phase_ms()
        pm = torch.tensor([phases.get(k, 0.0) / steps for k in ("lights+cull_begin", "totals", "cull_finish", "push_grid", "wait_grid")],
                          device=dev, dtype=torch.float64)
        dist.all_reduce(pm, op=dist.ReduceOp.MAX)
        sc._marks = None
        sc.frame_pipelined(f_last, n, chunks=4, wire="raw")   # the frame slot (camera) whose grid the routed frame hashed
        barrier()
        h, total = full_grid_hash(torch, hvcf, eng.ctx, f_last, dev, sp)
        res = {"ms": float(t.item()) / steps, "steps": steps, "wire": "ncclBroadcast of 48 B LightData from rank 0, 4 chunks pipelined",
               "phase_ms_max_over_ranks": dict(zip(("lights+cull_begin", "totals", "cull_finish", "push_grid", "wait_grid"), [round(float(x), 4) for x in pm.tolist()])),
               "grid_hash": f"{h & 0xffffffffffffffff:016x}", "total_indices": int(total), "peer_error": eng.peer_status()}
    finally:
        torch.cuda.synchronize()
        sc.close()
        eng.close()
    return res


def boundary_counts(impl_oracle, names, threads):
    """north_star: "boundary disagreements are counted and reported, as lights within 1 ulp of a cluster plane" -- the oracle's
    count of (cluster, light) pairs whose squared distance is within 1 ulp of r^2 (the strict `<` of LightCulling.comp:56),
    and of pairs whose result flips when the sum is contracted into FMAs (the SPIR-V carries no NoContraction).  Whole frame
    for A1k / A65k; every 64th cluster for B1M (the count is then scaled, and says so)."""
    out = {}
    for name in names:
        grid, screen, n = WORKLOADS[name]
        ubo = impl_oracle.camera_ubo(grid, screen)
        aabbs = impl_oracle.generate_aabbs(ubo)
        lights = impl_oracle.init_lights(n)
        stride = max(1, int(np.ceil(len(aabbs) * n / 6.0e8)))
        sub = aabbs if stride == 1 else np.ascontiguousarray(aabbs[::stride])
        near, flips = impl_oracle.boundary_report(ubo, sub, lights, threads=threads)
        out[name] = {"within_1ulp": near * stride, "flip_under_fma": flips * stride, "pairs": len(aabbs) * n,
                     "sample": "all pairs" if stride == 1 else f"every {stride}th cluster, scaled"}
    return out


def run_ours(args):
    import torch
    import torch.distributed as dist
    from hellovulkan_b200 import hvcf

    hvcf.load()  # raises if libhvcf.so is missing: there is no fallback
    if not torch.cuda.is_available():
        raise RuntimeError("bench.py --impl ours needs a CUDA device (libhvcf has no CPU path)")
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus:
        log(f"warning: --gpus {args.gpus} but WORLD_SIZE {world}; using WORLD_SIZE")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    numa_cores = pin_to_gpu_node(local) if world > 1 else None
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        opts = dist.ProcessGroupNCCL.Options()
        opts.is_high_priority_stream = True   # the comparison leg pipelines an NCCL broadcast with compute
        dist.init_process_group("nccl", device_id=dev, pg_options=opts)

    def barrier():
        if world > 1:
            dist.barrier(device_ids=[local])
        torch.cuda.synchronize()

    grid, screen, n = WORKLOADS[args.workload]
    C = grid[0] * grid[1] * grid[2]
    K, W = max(1, args.steps), max(args.warmup, 3)
    flush = torch.empty(L2_FLUSH_BYTES, dtype=torch.uint8, device=dev)
    sampler = ClockSampler(local) if rank == 0 else None
    nominal = C * n
    t_wall0 = time.time()
    summary = {}
    if world == 1:
        head = measure_single(hvcf, torch, args.workload, K, W, local, flush, barrier, e2e_steps=20, pipelined=not args.no_pipelined_e2e)
        ms_per_step = head["ms"]
        roofline = head["roofline"]
        e2e_res_ms = head["e2e_resident"]["ms_per_step"]
        # headline: two frames in flight (the reference app's frame ring) where that leg ran and reproduced the one-at-a-time bytes
        top = head.get("e2e_pipelined", head["e2e"])
        e2e_ms = top["ms_per_step"]
        e2e = {"value": nominal / (e2e_ms * 1e-3), "unit": "tests/s", "ms_per_step": e2e_ms, "steps": top["steps"],
               "h2d_bytes_per_step": top["h2d_bytes_per_step"], "d2h_bytes_per_step": top["d2h_bytes_per_step"],
               "mode": top["mode"],
               "one_frame_at_a_time": {"value": nominal / (head["e2e"]["ms_per_step"] * 1e-3), "ms_per_step": head["e2e"]["ms_per_step"],
                                       "mode": head["e2e"]["mode"]},
               "lights_resident": {"value": nominal / (e2e_res_ms * 1e-3), "ms_per_step": e2e_res_ms,
                                   "h2d_bytes_per_step": head["e2e_resident"]["h2d_bytes_per_step"],
                                   "d2h_bytes_per_step": head["e2e_resident"]["d2h_bytes_per_step"], "mode": head["e2e_resident"]["mode"]},
               "timer": "host perf_counter around the loop, synchronised both sides (the download blocks)"}
        launches_per_step, total_idx = head["launches_per_step"], head["total_indices"]
        parallelism = "single GPU: hvcf_cull, all lights resident"
        detail = {"l2": head["l2"], "step_ms_min": head["ms_min"], "step_ms_median": head["ms_median"], "stats": head["stats"]}
        summary[args.workload] = {"ms": round(ms_per_step, 4), "frac": round(roofline["frac"], 3), "e2e_ms": round(e2e_ms, 3),
                                  "e2e_res_ms": round(e2e_res_ms, 3), "top": roofline["kernel"].split(" ")[0]}
    else:
        head = run_routed(args, hvcf, torch, dist, world, rank, local, dev, barrier, K, W, flush)
        ms_per_step = head["ms"]
        roofline = head["roofline"]
        e2e_res_ms = head["e2e_resident"]["ms_per_step"]
        serial_mode = ("one frame at a time; every rank: pinned H2D of its own share of the LightData + the UBO, hvcf_cull_routed, D2H of its own "
                       "slab's cells + index segment (hvcf_download_shard); bytes are whole-job sums")
        top = head.get("e2e_pipelined", head["e2e"])
        e2e_ms = top["ms_per_step"]
        e2e = {"value": nominal / (e2e_ms * 1e-3), "unit": "tests/s", "ms_per_step": e2e_ms, "steps": top["steps"],
               "h2d_bytes_per_step": top["h2d_bytes_per_step"], "d2h_bytes_per_step": top["d2h_bytes_per_step"],
               "mode": top.get("mode", serial_mode),
               "one_frame_at_a_time": {"value": nominal / (head["e2e"]["ms_per_step"] * 1e-3), "ms_per_step": head["e2e"]["ms_per_step"],
                                       "mode": serial_mode},
               "lights_resident": {"value": nominal / (e2e_res_ms * 1e-3), "ms_per_step": e2e_res_ms,
                                   "h2d_bytes_per_step": head["e2e_resident"]["h2d_bytes_per_step"],
                                   "d2h_bytes_per_step": head["e2e_resident"]["d2h_bytes_per_step"]},
               "timer": "host perf_counter around the loop, barrier + synchronise both sides, max over ranks"}
        launches_per_step, total_idx = head["launches_per_step"], head["total_indices"]
        parallelism = (f"{world} ranks: clusters sharded into row slabs {head['row_cuts']} (row = y + z * {grid[1]}), lights owned in shards and routed over NVLink peer "
                       f"memory (hvcf_cull_routed); no collective call in the frame")
        detail = {k: head[k] for k in ("row_cuts", "rows_per_slice", "balance_iterations", "inbox_records_per_rank", "light_range_per_rank", "own_kernels_ms_per_rank",
                                       "kernel_ms_per_rank", "route_kernel_ms_per_rank", "sharded_bytes_equal", "grid_hash", "peer_error", "executed_tests_all_ranks", "repro") if k in head}
        inbox = head["inbox_records_per_rank"]
        own = head["own_kernels_ms_per_rank"]
        summary["routed"] = {"ms": round(ms_per_step, 4), "bytes_equal": head["sharded_bytes_equal"], "row_cuts": head["row_cuts"],
                             "own_ms_min_max": [min(own), max(own)], "inbox_Mrec_min_max": [round(min(inbox) / 1e6, 2), round(max(inbox) / 1e6, 2)],
                             "inbound_MB_max": round(max(inbox) * 16 / 1e6, 1), "e2e_ms": round(e2e_ms, 3), "e2e_res_ms": round(e2e_res_ms, 3),
                             "frac": round(roofline["frac"], 3), "top": roofline["kernel"].split(" ")[0],
                             "kernels_ms_crit_rank": head["kernel_ms_per_rank"][roofline["rank"]]}
        if not args.no_compare:
            try:
                cmp_ = run_broadcast_formulation(args, hvcf, torch, dist, world, rank, local, dev, barrier, head["row_cuts"], (K - 1) & 1)
                cmp_["bytes_equal"] = cmp_.get("grid_hash") == head["grid_hash"]
                detail["nccl_broadcast_formulation"] = cmp_
                summary["nccl_bcast"] = {"ms": round(cmp_["ms"], 4), "bytes_equal": cmp_["bytes_equal"], "phases": cmp_["phase_ms_max_over_ranks"]}
            except Exception as e:
                detail["nccl_broadcast_formulation"] = {"error": f"{type(e).__name__}: {e}"}
    if sampler:
        sampler.window(t_wall0, time.time())
    value = nominal / (ms_per_step * 1e-3)
    roofline["traffic"] = ncu_traffic(args.workload if world == 1 else f"{args.workload}_N{world}")
    hbm_gbs, hbm_src = hbm_peak()
    alg_bytes = 48 * n + 16 * n * 2 + 32 * C + 8 * C + 4 * total_idx   # AoS in, view records out + in, AABB, cells, indices
    roofline["nominal_frac_of_step"] = nominal * LANE_OPS_PER_TEST / (ms_per_step * 1e-3) / (roofline["peak"] * 1e12) / world
    roofline["hbm"] = {"algorithmic_bytes_per_step": alg_bytes, "achieved_gbs": alg_bytes / (ms_per_step * 1e-3) / 1e9, "peak_gbs": hbm_gbs * world,
                       "peak_source": hbm_src, "frac": alg_bytes / (ms_per_step * 1e-3) / 1e9 / (hbm_gbs * world)}

    # ---- the other configs, one GPU each (N = 1 only): ms / frac / e2e per workload, graph-launched config 1 ----
    extra = {}
    if world == 1 and not args.no_summary:
        for name in ("A1k", "A65k", "A1M", "B1M"):
            if name == args.workload:
                continue
            try:
                r = measure_single(hvcf, torch, name, min(K * 5, 100), W, local, flush, e2e_steps=40, graph=(name == "A1k"))
                extra[name] = r
                summary[name] = {"ms": round(r["ms"], 4), "frac": round(r["roofline"]["frac"], 3), "e2e_ms": round(r["e2e"]["ms_per_step"], 3),
                                 "e2e_res_ms": round(r["e2e_resident"]["ms_per_step"], 3), "top": r["roofline"]["kernel"].split(" ")[0]}
                if "graph" in r:
                    summary[name]["graph_ms"] = round(r["graph"].get("ms", float("nan")), 4)
                    summary[name]["kernels_only_ms"] = round(r["roofline"]["instrumented_step_ms"], 4)
            except Exception as e:
                summary[name] = {"error": f"{type(e).__name__}: {e}"[:120]}
    next_rows = None
    if rank == 0 and world == 1 and not args.no_next_rows:
        try:
            stream = torch.cuda.current_stream()
            g2, s2, n2 = WORKLOADS["A65k"]
            with hvcf.Context(grid=g2, max_lights=n2, device=local) as c2:
                c2.set_lights(hvcf.host_init_lights(n2), stream.cuda_stream)
                c2.set_ubo(0, hvcf.host_camera_ubo(g2, s2), stream.cuda_stream)
                next_rows = run_next_rows(args, hvcf, c2, g2, s2, n2, local, stream, stream.cuda_stream, flush)
            next_rows.update(run_extras(args, hvcf, local, stream, stream.cuda_stream, flush))
        except Exception as e:
            next_rows = {"error": f"{type(e).__name__}: {e}"}
    del flush
    fly = None
    if not args.no_flythrough:
        try:
            fly = run_flythrough(args, hvcf, world, rank, local, dev, barrier)
            summary["fly_B1M"] = {"ms_per_frame": round(fly["ms_per_frame"], 4), "frames": fly["frames"]}
        except Exception as e:
            fly = {"error": f"{type(e).__name__}: {e}"}
    if sampler:
        sampler.stop()
    clocks = sampler.summary() if sampler else None

    # ---- CPU baseline beside it (rank 0, N = 1 only), and the boundary-disagreement counts ----
    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        kind, impl, _ = cpu_checker()
        cpu = time_cpu(kind, impl, args.workload, args.cpu_seconds)
        try:
            sys.path.insert(0, os.path.join(ROOT, "tests"))
            import cpu_checkers as cc
            bc = boundary_counts(cc.Oracle(), ("A1k", "A65k", "B1M"), host_threads())
            cpu["boundary"] = bc
            summary["boundary"] = {k: [v["within_1ulp"], v["flip_under_fma"]] for k, v in bc.items()}
        except Exception as e:
            cpu["boundary"] = {"error": f"{type(e).__name__}: {e}"}
    if world > 1:
        dist.barrier(device_ids=[local])
        dist.destroy_process_group()
    if rank != 0:
        return 0
    ok = world == 1 or bool(head.get("sharded_bytes_equal"))
    line = {
        "metric": "cluster-light tests/s", "value": value, "unit": "tests/s", "n_gpus": world, "steps": K, "warmup": W,
        "ms_per_step": ms_per_step, "ms_per_frame": ms_per_step, "higher_is_better": True, "scaling": "strong",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": args.workload, "grid": list(grid), "screen": list(screen), "lights": n,
                   "max_lights_per_cluster": 150, "light_distribution": "AppPBRClusterForward::InitLights, mt19937(12345)",
                   "l2": f"flushed between steps ({L2_FLUSH_BYTES >> 20} MiB memset, untimed); the inputs are larger than L2 anyway",
                   "parallelism": parallelism, "host_cores_per_rank": numa_cores,
                   "tests": "nominal = clusters x lights; executed (early exit at 150 hits, LightCulling.comp:71) under roofline"},
        "total_indices": total_idx, "e2e": e2e, "gpu_launches": launches_per_step * K, "gpu_launches_per_step": launches_per_step,
        "roofline": roofline, "cpu_baseline": cpu, "clocks": clocks, "impl": "ours", "detail": detail,
    }
    if extra:
        line["workloads"] = extra
    if next_rows is not None:
        line["next_rows"] = next_rows
    if fly is not None:
        line["flythrough"] = fly
    line["summary"] = summary   # LAST: the driver keeps the tail of stdout
    print(json.dumps(line), flush=True)
    return 0 if ok else 3


def main():
    args = parse_args()
    if args.impl == "reference":
        return run_reference_arm(args)
    return run_ours(args)


if __name__ == "__main__":
    sys.exit(main())
