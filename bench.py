#!/usr/bin/env python
"""bench.py — headline benchmark of the forward re-rendering path (BASELINE.json).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]

Metric: warped Mpix/s on BASELINE config 2 — a synthetic RGB+depth clip of 49 frames at 384x672
warped along the 8 inference_orbits.py orbit presets = 392 (frame, pose) units = 101.2 Mpix per
step per GPU.  One "step" is one pass of the hot path over that batch.  Multi-GPU is weak scaling:
every rank warps its own clip's 392 units, no data-path collective in the timed step (units are
independent); `value` = units of all ranks / max-over-ranks device time.  The one exchange the
pipeline has — gathering the warped condition video on the denoiser's rank — is timed separately
and reported under "gather".

One JSON line on stdout (rank 0).  See the driver contract in the task statement for the keys.
"""
import argparse
import json
import os
import statistics
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

ALGO_BYTES_PER_PX = 44.0          # read RGB 12 + depth 4, write colour 12 + mask 4 + depth 4 + flow 8
N_FRAMES, H, W = 49, 384, 672
# dram__bytes_read.sum + dram__bytes_write.sum of one k_splat_pair launch from the committed ncu --set full capture
# (profiles/), for the 24-unit launch shape the roofline figure is quoted on
# dram__bytes_read.sum + dram__bytes_write.sum of one launch of the dominant kernel (ncu --set full, profiles/)
NCU_TRAFFIC_BYTES_PER_LAUNCH = {"atomic": 238_216_960,   # r1_final_kernels_full_24units_ncu.txt: k_splat_pair, 24 units
                                "window": 209_976_832}    # r2_kernels_full_24units_ncu.txt: k_splat_window, 24 units: 67.8 MB read + 142.2 MB written
# DRAM bytes of one whole step from the ncu launch list / full captures of the timed schedule (profiles/), per mode
NCU_DRAM_BYTES_PER_STEP = {"atomic": 3_160_000_000,      # r1 launch list
                           "window": 7_645_000_000}      # profiles/r2_launches_summary.txt: 1.943 GB read + 5.702 GB written (99 launches of 12 units: slabs stream through HBM)
# bytes each kernel itself must move per unit pixel (its own inputs + outputs, intermediates included)
OWN_BYTES_PER_PX = {"window": {"prepass": 16.0, "splat": 44.0, "normalise": 40.0},       # K1: depth in, flow + z' out; K2: flow, z', RGB in, >= 1 slab cell out; K3: >= 1 slab cell in, 20 B out
                    "atomic": {"prepass": 4.0, "splat": 24.0, "normalise": 40.0}}


ACCUMULATION = {"window": "fixed order: warp-private shared-memory windows + ordered gather (default, reproducible)",
                "atomic": "fp32 L2 reductions (order varies run to run)"}


def load_peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        with open(path) as fh:
            return float(json.load(fh)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


class ClockSampler:
    """Samples SM clock and throttle reasons of one GPU during the timed region (NVML, ~5 ms period)."""

    def __init__(self, index):
        self.index, self.samples, self.reasons, self.max_mhz = index, [], set(), None
        self._stop = threading.Event()
        self._thread = None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.handle = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.handle, pynvml.NVML_CLOCK_SM)
        except Exception:
            self.nv = None

    def _loop(self):
        nv = self.nv
        names = {
            "hw_slowdown": getattr(nv, "nvmlClocksThrottleReasonHwSlowdown", 0x8),
            "hw_thermal_slowdown": getattr(nv, "nvmlClocksThrottleReasonHwThermalSlowdown", 0x40),
            "sw_thermal_slowdown": getattr(nv, "nvmlClocksThrottleReasonSwThermalSlowdown", 0x20),
            "sw_power_cap": getattr(nv, "nvmlClocksThrottleReasonSwPowerCap", 0x4),
        }
        while not self._stop.is_set():
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self.handle, nv.NVML_CLOCK_SM))
                mask = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.handle)
                for k, bit in names.items():
                    if mask & bit:
                        self.reasons.add(k)
            except Exception:
                pass
            time.sleep(0.005)

    def start(self):
        if self.nv is not None:
            self._thread = threading.Thread(target=self._loop, daemon=True)
            self._thread.start()

    def stop(self):
        self._stop.set()
        if self._thread is not None:
            self._thread.join(1.0)
        if not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": self.max_mhz, "reasons": [], "samples": 0}
        return {"sm_mhz": statistics.median(self.samples), "sm_max_mhz": self.max_mhz,
                "reasons": sorted(self.reasons), "samples": len(self.samples)}


def build_workload(seed, with_poses=False):
    import torch
    from trajectorycrafter_b200 import prepare_matrices
    from trajectorycrafter_b200.workloads import orbit_units, synthetic_clip
    frames, depths = synthetic_clip(N_FRAMES, H, W, seed=seed)
    T1, T2, K, src = orbit_units(depths, H, W)
    M = prepare_matrices(T1, T2, K)
    if with_poses:
        return frames, depths, M, src.to(torch.int32), (T1, T2, K)
    return frames, depths, M, src.to(torch.int32)


def bind_to_gpu_cpus(index):
    """Restrict this process to the CPUs NVML reports as local to GPU `index` (ranks of one box otherwise all pin their host
    buffers on NUMA node 0 and share its memory controllers).  Returns the CPU list, or None if NVML cannot tell."""
    try:
        import pynvml
        pynvml.nvmlInit()
        hnd = pynvml.nvmlDeviceGetHandleByIndex(index)
        words = pynvml.nvmlDeviceGetCpuAffinity(hnd, (os.cpu_count() + 63) // 64)
        cpus = [64 * i + b for i, wd in enumerate(words) for b in range(64) if (int(wd) >> b) & 1]
        cpus = [c for c in cpus if c in os.sched_getaffinity(0)]
        if cpus:
            os.sched_setaffinity(0, cpus)
            return [cpus[0], cpus[-1], len(cpus)]
    except Exception:
        pass
    return None


def host_description():
    model = "unknown CPU"
    try:
        with open("/proc/cpuinfo") as fh:
            for ln in fh:
                if ln.startswith("model name"):
                    model = ln.split(":", 1)[1].strip()
                    break
    except OSError:
        pass
    return model


def reference_torch_throughput(frames, depths, poses, src, unit_ids, threads, deterministic):
    """Times the reference's OWN torch implementation (oracle/_ref copy of models/utils.py, or /root/reference in the
    build container) exactly as its callers loop it (demo.py:77-90): Warper(device='cpu'), one forward_warp call per
    unit with b = 1, mask1 = None, twice = False.  Returns (Mpix/s, seconds)."""
    import torch
    from oracle import ref_loader
    ref = ref_loader.load()
    T1, T2, K = poses
    prev_threads, prev_det = torch.get_num_threads(), torch.are_deterministic_algorithms_enabled()
    torch.set_num_threads(threads)
    torch.use_deterministic_algorithms(deterministic)
    try:
        warper = ref.Warper(device="cpu")
        s0 = int(src[unit_ids[0]])
        warper.forward_warp(frames[s0:s0 + 1], None, depths[s0:s0 + 1], T1[:1], T2[:1], K[:1], None, False, twice=False)     # warm-up unit
        t0 = time.perf_counter()
        for u in unit_ids:
            s_ = int(src[u])
            warper.forward_warp(frames[s_:s_ + 1], None, depths[s_:s_ + 1], T1[u:u + 1], T2[u:u + 1], K[u:u + 1], None, False, twice=False)
        dt = time.perf_counter() - t0
    finally:
        torch.set_num_threads(prev_threads)
        torch.use_deterministic_algorithms(prev_det)
    return len(unit_ids) * H * W / dt / 1e6, dt


def oracle_throughput(frames, depths, M, src, unit_ids, threads):
    """Times the CPU oracle (oracle/warp_oracle.c, the port of the reference's path) on `unit_ids`,
    one unit per task over `threads` host threads.  Returns (Mpix/s, seconds)."""
    from concurrent.futures import ThreadPoolExecutor
    import oracle
    src_np = src.numpy()

    def one(u):
        oracle.forward_warp(frames, None, depths, M[u:u + 1], src_index=src_np[u:u + 1], per_unit_max=True)

    t0 = time.perf_counter()
    with ThreadPoolExecutor(threads) as ex:
        list(ex.map(one, unit_ids))
    dt = time.perf_counter() - t0
    return len(unit_ids) * H * W / dt / 1e6, dt


def run_reference(args):
    """--impl reference: the reference's own CPU implementation of the path on the box's host cores.  When the verbatim
    run-time copy of /root/reference/models/utils.py is present (oracle/_ref/, made by oracle/build_ref.py in the build
    container; it travels to the GPU box) the UNMODIFIED torch Warper is timed (kind "reference"), looped as demo.py:77-90
    with every host core; otherwise the C port of the same arithmetic (kind "port").  Each step is a bounded sample of
    the 392-unit workload: one source frame under the 8 orbit presets, the frame advancing by 6 per step."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    import oracle
    from oracle import ref_loader
    oracle.build()
    frames, depths, M, src, poses = build_workload(1234, with_poses=True)
    threads = os.cpu_count() or 1
    use_ref = ref_loader.available()

    def one_step(k):
        f = (6 * k) % N_FRAMES
        unit_ids = [f * 8 + p for p in range(8)]
        if use_ref:
            return reference_torch_throughput(frames, depths, poses, src, unit_ids, threads, False)[1]
        return oracle_throughput(frames, depths, M, src, unit_ids, threads)[1]

    for k in range(max(1, min(args.warmup, 3))):
        one_step(k)
    t0 = time.perf_counter()
    for k in range(args.steps):
        one_step(k)
    dt = (time.perf_counter() - t0) / args.steps
    value = 8 * H * W / dt / 1e6
    kind = "reference" if use_ref else "port"
    sample = (f"8 of 392 units per step (source frame (6*step) mod 49 under the 8 orbit presets, b=1 calls as demo.py:77-90), "
              f"{threads} threads, {'unmodified torch Warper(device=cpu), default (non-deterministic) scatter' if use_ref else 'C port of the path'}; "
              f"host: {host_description()}")
    line = {
        "impl": "reference", "metric": "warped_mpix_per_s", "value": value, "unit": "Mpix/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt * 1e3,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": workload_config(),
        "cpu_baseline": {"value": value, "unit": "Mpix/s", "cores": threads, "kind": kind, "sample": sample},
        "e2e": {"value": value, "unit": "Mpix/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


def workload_config():
    return {"workload": "config2: synthetic RGB+depth 49x384x672 x 8 inference_orbits.py orbit presets "
                        "(392 frame/pose units = 101.2 Mpix per step per GPU), per-frame soft-z max, mask1=None",
            "units_per_step_per_gpu": N_FRAMES * 8, "frames": N_FRAMES, "h": H, "w": W,
            "l2": "per-step inputs+outputs are 3.0 GB >> 126 MB L2, so no L2 flush between steps",
            "sharding": "units are independent: each rank warps its own 392 units, no collective in the step"}


def run_ours(args):
    import torch
    import torch.distributed as dist
    from trajectorycrafter_b200 import Warper

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: this path has no CPU fallback")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    numa = bind_to_gpu_cpus(local) if world > 1 else None      # pinned host buffers are first-touched on the GPU's own NUMA node
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)
        # first collectives now: NCCL sets its communicator up lazily (buffers, proxy threads) and was seen to
        # disturb the first milliseconds after the first barrier -- keep that out of the timed region
        warm = torch.ones(1, dtype=torch.float64, device=dev)
        dist.all_reduce(warm, op=dist.ReduceOp.MAX)
        dist.barrier()
        torch.cuda.synchronize()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(v):
        if world == 1:
            return v
        t = torch.tensor([v], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    frames, depths, M, src = build_workload(1234 + rank)
    n_units = M.shape[0]
    px_per_step = n_units * H * W
    warper = Warper(device=f"cuda:{local}")
    frames_d, depths_d = frames.to(dev), depths.to(dev)
    M_d, src_d = M.to(dev), src.to(dev)      # inputs resident in HBM before the timed region (matrices and indices too)
    out = (torch.empty((n_units, 3, H, W), device=dev), torch.empty((n_units, 1, H, W), device=dev),
           torch.empty((n_units, 1, H, W), device=dev), torch.empty((n_units, 2, H, W), device=dev))
    flag = torch.zeros(1, dtype=torch.int32, device=dev)

    def step():
        warper.forward_warp_matrices(frames_d, depths_d, M_d, src_index=src_d, per_frame_max=True,
                                     check_range=flag, out=out, accumulation=args.accumulation)

    for _ in range(max(3, args.warmup)):
        step()
    barrier()
    # ---- timed region: K steps, device-timed on the launching (current) stream
    sampler = ClockSampler(local)
    launches0 = warper.launch_count()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    sampler.start()
    ev0.record()
    t_host = time.perf_counter()
    for _ in range(args.steps):
        step()
    host_ms = (time.perf_counter() - t_host) / args.steps * 1e3     # CPU time to enqueue one step (no sync inside)
    ev1.record()
    torch.cuda.synchronize()
    clocks = sampler.stop()
    barrier()
    ms_local = ev0.elapsed_time(ev1) / args.steps
    launches = warper.launch_count() - launches0
    assert int(flag.item()) == 0, "range check failed"
    ms = max_over_ranks(ms_local)
    value = world * px_per_step / (ms * 1e-3) / 1e6
    per_rank = None
    if world > 1:      # which rank set the max, and at what clock (diagnosis of weak-scaling loss; not part of the metric)
        mine = torch.tensor([ms_local, float(clocks.get("sm_mhz") or 0.0)], dtype=torch.float64, device=dev)
        allr = [torch.zeros_like(mine) for _ in range(world)]
        dist.all_gather(allr, mine)
        per_rank = {"ms_per_step": [round(float(t[0]), 4) for t in allr], "sm_mhz": [float(t[1]) for t in allr]}

    # ---- roofline of the dominant kernel: CUDA-event durations of every launch, events recorded on the stream each
    # kernel is launched on.  Pass A repeats the timed schedule with per-launch events (four streams run
    # concurrently, so a launch's duration includes the time it shares the SMs with the other lanes' kernels);
    # pass B times the same kernels alone: one stream, 24-unit launches (TCW_SERIAL=1, TCW_CHUNK_UNITS=24).
    # `step_frac` is the whole timed step against the same peak.
    peak, peak_src = load_peaks()

    def profile_pass(env):
        os.environ.update(env)
        w2 = Warper(device=f"cuda:{local}")
        w2._context()              # tcw_create reads the TCW_* switches
        for k in env:
            os.environ.pop(k)

        def one():
            w2.forward_warp_matrices(frames_d, depths_d, M_d, src_index=src_d, per_frame_max=True,
                                     check_range=flag, out=out, accumulation=args.accumulation)
        for _ in range(2):
            one()
        torch.cuda.synchronize()
        n = max(1, min(args.steps, 5))
        w2.profile(True)
        for _ in range(n):
            one()
        torch.cuda.synchronize()
        res = w2.profile_read()
        w2.profile(False)
        return res, n

    prof, prof_steps = profile_pass({})
    iso, iso_steps = profile_pass({"TCW_LANES": "1", "TCW_CHUNK_UNITS": "24"})
    roofline = None
    i_ms, i_n, i_units = iso["splat"]
    k_ms, k_n, k_units = prof["splat"]
    if i_n and k_n:
        # roofline.frac is SURVEY 8(d)'s figure: the WHOLE timed step (CUDA events around the batched submission) against
        # 44 B per warped source pixel.  Beside it: every kernel timed ALONE (one stream, 24-unit launches) against the
        # bytes that kernel itself must move, and the same kernels inside the timed multi-lane schedule.
        def gbps(bytes_per_px, ms_, units_):
            return bytes_per_px * units_ * H * W / (ms_ * 1e-3) / 1e9
        kname = {"window": "k_splat_window", "atomic": "k_splat_pair"}[args.accumulation]
        own = OWN_BYTES_PER_PX[args.accumulation]
        tot_i, tot_k = sum(v[0] for v in iso.values()), sum(v[0] for v in prof.values())
        step_gbps = ALGO_BYTES_PER_PX * px_per_step / (ms_local * 1e-3) / 1e9
        alone = {}
        for name, (ms_, n_, units_) in iso.items():
            if n_:
                alone[name] = {"ms_per_step": ms_ / iso_steps, "own_bytes_per_px": own.get(name), "avg_launch_us": ms_ / n_ * 1e3,
                               "achieved_GBps": gbps(own[name], ms_, units_) if own.get(name) else None,
                               "frac": gbps(own[name], ms_, units_) / peak if own.get(name) else None}
        roofline = {"bound": "hbm", "kernel": kname, "achieved": step_gbps, "peak": peak, "unit": "GB/s",
                    "frac": step_gbps / peak, "traffic": NCU_DRAM_BYTES_PER_STEP.get(args.accumulation),
                    "traffic_what": "dram__bytes_read.sum + dram__bytes_write.sum of ALL launches of one step (ncu launch list under profiles/), "
                                    "to be read against algorithmic_bytes_per_step",
                    "kernel_alone_traffic_per_launch": NCU_TRAFFIC_BYTES_PER_LAUNCH.get(args.accumulation),
                    "peak_source": peak_src,
                    "what": "frac / achieved: 44 B x pixels of one step / device time of the whole timed step (all kernels, the timed "
                            "schedule); kernel_alone: each kernel on one stream in 24-unit launches against its OWN bytes per pixel "
                            "(outputs larger than L2 there, so its intermediates stream through HBM); in_schedule: the same launches "
                            "inside the timed multi-lane schedule (they share the SMs, so the times add up to more than the step)",
                    "algorithmic_bytes_per_step": ALGO_BYTES_PER_PX * px_per_step,
                    "kernel_share_of_step": i_ms / tot_i if tot_i else None,
                    "kernel_alone": alone,
                    "kernel_alone_frac": alone.get("splat", {}).get("frac"),
                    "in_schedule": {"per_kernel_ms_per_step": {k: v[0] / prof_steps for k, v in prof.items()},
                                    "launches_per_step": sum(v[1] for v in prof.values()) / prof_steps,
                                    "units_per_launch": k_units / k_n, "kernel_share_of_step": k_ms / tot_k if tot_k else None},
                    "dram_GBps_actual": (NCU_DRAM_BYTES_PER_STEP[args.accumulation] / (ms_local * 1e-3) / 1e9
                                         if NCU_DRAM_BYTES_PER_STEP.get(args.accumulation) else None),
                    "step_frac": step_gbps / peak}

    # ---- the opt-in float-reduction mode on the same step, same timing (not the metric: its summation order is not fixed)
    other_modes = None
    if rank == 0 and world == 1 and not args.no_extras and args.accumulation == "window":
        def atomic_step():
            warper.forward_warp_matrices(frames_d, depths_d, M_d, src_index=src_d, per_frame_max=True, check_range=flag, out=out,
                                         accumulation="atomic")
        for _ in range(4):
            atomic_step()
        a0, a1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a0.record()
        for _ in range(args.steps):
            atomic_step()
        a1.record()
        torch.cuda.synchronize()
        a_ms = a0.elapsed_time(a1) / args.steps
        other_modes = {"atomic": {"ms_per_step": a_ms, "Mpix_per_s": px_per_step / (a_ms * 1e-3) / 1e6,
                                  "frac": ALGO_BYTES_PER_PX * px_per_step / (a_ms * 1e-3) / 1e9 / peak,
                                  "what": ACCUMULATION["atomic"] + " -- Warper(..., accumulation='atomic') / TCW_ATOMIC_F32"}}
        for _ in range(2):      # leave the default mode's graph and buffers as the later legs expect them
            step()
        torch.cuda.synchronize()

    # ---- the literal metric shape: ONE 49-unit trajectory (49x384x672, preset (0, 30)), inputs resident, device-timed
    clip49 = None
    if rank == 0 and not args.no_extras:
        sel = torch.arange(N_FRAMES, device=dev) * 8 + 1
        M49, src49 = M_d[sel].contiguous(), src_d[sel].contiguous()
        out49 = tuple(t[:N_FRAMES] for t in out)
        for _ in range(5):
            warper.forward_warp_matrices(frames_d, depths_d, M49, src_index=src49, per_frame_max=True, check_range=flag, out=out49,
                                         accumulation=args.accumulation)
        c0, c1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        reps = 40
        c0.record()
        for _ in range(reps):
            warper.forward_warp_matrices(frames_d, depths_d, M49, src_index=src49, per_frame_max=True, check_range=flag, out=out49,
                                         accumulation=args.accumulation)
        c1.record()
        torch.cuda.synchronize()
        c_ms = c0.elapsed_time(c1) / reps
        c_px = N_FRAMES * H * W
        clip49 = {"ms": c_ms, "Mpix_per_s": c_px / (c_ms * 1e-3) / 1e6, "frac": ALGO_BYTES_PER_PX * c_px / (c_ms * 1e-3) / 1e9 / peak,
                  "what": "one 49-frame trajectory (orbit preset (0, 30)) = 12.6 Mpix per call, 556 MB algorithmic; back-to-back calls "
                          "re-use 556 MB of buffers, more than the 126 MB L2"}

    # ---- the drop-in entries end to end (ADVICE): device tensors in, fresh outputs, host-side matrix preparation inside
    dropin = None
    if rank == 0 and world == 1 and not args.no_extras:
        from trajectorycrafter_b200.workloads import orbit_units
        T1, T2, K, _ = orbit_units(depths, H, W, presets=[(0, 30)])
        f49, d49 = frames_d, depths_d
        def batch_call():
            return warper.forward_warp_batch(f49, d49, T1, T2, K, accumulation=args.accumulation)
        def frame_loop():
            outs = []
            for i in range(N_FRAMES):      # the reference's loop, demo.py:77-90
                outs.append(warper.forward_warp(f49[i:i + 1], None, d49[i:i + 1], T1[i:i + 1], T2[i:i + 1], K[i:i + 1], None))
            return outs
        dropin = {}
        for name, fn in (("forward_warp_batch_49", batch_call), ("forward_warp_b1_loop_49", frame_loop)):
            r = fn(); r = fn(); torch.cuda.synchronize()      # two rounds: the caching allocator then holds both generations of outputs
            t0 = time.perf_counter()
            for _ in range(3):
                r = fn()
            torch.cuda.synchronize()
            dt = (time.perf_counter() - t0) / 3
            dropin[name] = {"ms": dt * 1e3, "Mpix_per_s": N_FRAMES * H * W / dt / 1e6}
            del r
            if os.environ.get("TCW_BENCH_PROFILE_DROPIN"):      # where the host time of the drop-in entries goes (stderr)
                import cProfile, pstats
                pr = cProfile.Profile()
                pr.enable(); fn(); torch.cuda.synchronize(); pr.disable()
                pstats.Stats(pr, stream=sys.stderr).sort_stats("cumulative").print_stats(18)
        dropin["what"] = ("wall clock incl. the host-side T2*T1^-1 / K^-1 preparation, output allocation and the range-flag read; "
                          "forward_warp_b1_loop_49 is the reference's own call pattern (49 calls with b = 1)")

    # ---- e2e: host buffers through the C-ABI host entry (H2D + kernels + D2H inside the timed region).  The step's result is
    # what every shipped caller keeps of the four outputs (`warped, mask, _, _`, demo.py:78-90): colour float32 + validity
    # mask, the mask as uint8; depth and flow12 stay on the device.  The all-four-outputs form is timed beside it at N = 1.
    e2e = None
    if not args.no_e2e:
        hf, hd = frames.pin_memory(), depths.pin_memory()
        Mh = M.contiguous()
        hc = (torch.empty((n_units, 3, H, W)).pin_memory(), torch.empty((n_units, 1, H, W), dtype=torch.uint8).pin_memory(), None, None)

        def host_call():
            return warper.forward_warp_host(hf, hd, Mh, src_index=src, outputs=hc, return_depth=False, return_flow=False, mask_u8=True)
        host_call()
        barrier()
        k_e2e = max(1, min(args.steps, 5))
        t0 = time.perf_counter()
        for _ in range(k_e2e):
            res = host_call()
        dt = (time.perf_counter() - t0) / k_e2e
        assert res[4] == 0
        dt = max_over_ranks(dt)
        e2e = {"value": world * px_per_step / dt / 1e6, "unit": "Mpix/s",
               "h2d_bytes_per_step": int(hf.numel() * 4 + hd.numel() * 4 + Mh.numel() * 4 + src.numel() * 4),
               "d2h_bytes_per_step": int(n_units * H * W * 13 + 4),
               "ms_per_step": dt * 1e3, "steps": k_e2e,
               "api": "Warper.forward_warp_host(return_depth=False, return_flow=False, mask_u8=True) -> tcw_forward_warp_host "
                      "(pinned host buffers in, colour f32 + mask u8 out)"}
        if rank == 0 and world == 1:
            houts = (hc[0], torch.empty((n_units, 1, H, W)).pin_memory(),
                     torch.empty((n_units, 1, H, W)).pin_memory(), torch.empty((n_units, 2, H, W)).pin_memory())
            warper.forward_warp_host(hf, hd, Mh, src_index=src, outputs=houts)
            t0 = time.perf_counter()
            for _ in range(k_e2e):
                warper.forward_warp_host(hf, hd, Mh, src_index=src, outputs=houts)
            dta = (time.perf_counter() - t0) / k_e2e
            e2e["all_four_outputs"] = {"value": px_per_step / dta / 1e6, "ms_per_step": dta * 1e3,
                                       "d2h_bytes_per_step": int(sum(t.numel() * 4 for t in houts) + 4),
                                       "what": "colour, mask, depth and flow12 as float32, the reference signature's full 4-tuple"}
            del houts
        del hc

    # ---- multi-GPU as the path is specified: ONE clip's 392 units sharded over the ranks (strong scaling), then the same
    # with the exchange inside the timed step -- colour float32 + mask uint8 (13 B per pixel) of ALL units on rank 0:
    # (a) one NCCL gather of per-rank packed buffers the warp writes directly, (b) the fused form: every rank's gather
    # kernel stores straight into rank 0's buffer through peer-mapped memory (no collective, no local copy)
    strong = with_gather = None
    if world > 1:
        from trajectorycrafter_b200.sharding import (PeerBuffer, balanced_units, gather_packed, packed_buffer, packed_views,
                                                     sources_of, unpack_gathered)
        sf, sd, sM, ssrc = build_workload(1234)                       # the same clip on every rank
        n_all = sM.shape[0]
        # a rank takes a block of frames of EVERY trajectory (the eight presets cost 0.43 - 0.78 ms per 49 units: one whole
        # preset per rank leaves the step at the slowest one), and writes each unit to its slot of the clip (dst_index)
        ids = balanced_units(n_all, N_FRAMES, world, rank)
        order = torch.cat([balanced_units(n_all, N_FRAMES, world, r) for r in range(world)])      # unit held at each gathered position
        nl = ids.numel()
        equal_shards = all(balanced_units(n_all, N_FRAMES, world, r).numel() == nl for r in range(world))      # true for N = 2, 4, 8
        needed, lsrc = sources_of(ssrc, ids)                          # a rank uploads only the source frames its units read
        sf_d, sd_d = sf[needed].contiguous().to(dev), sd[needed].contiguous().to(dev)
        sM_d, ss_d, ids_d = sM[ids].contiguous().to(dev), lsrc.to(dev), ids.to(dev)
        o_depth, o_flow = torch.empty((nl, 1, H, W), device=dev), torch.empty((nl, 2, H, W), device=dev)
        o_col, o_msk = torch.empty((nl, 3, H, W), device=dev), torch.empty((nl, 1, H, W), device=dev)
        a_depth, a_flow = torch.empty((n_all, 1, H, W), device=dev), torch.empty((n_all, 2, H, W), device=dev)   # scatter form: clip-sized
        packed = packed_buffer(nl, H, W, dev)
        p_col, p_msk = packed_views(packed, nl, H, W)
        peer = PeerBuffer(warper, n_all, H, W, dst=0)

        def warp_into(col, msk, u8, scatter=False):
            if scatter:      # every unit straight to its slot of the clip on rank 0; depth / flow12 stay local
                warper.forward_warp_matrices(sf_d, sd_d, sM_d, src_index=ss_d, per_frame_max=True, check_range=flag,
                                             out=(col, msk, a_depth, a_flow), mask_u8=u8, dst_index=ids_d)
            else:
                warper.forward_warp_matrices(sf_d, sd_d, sM_d, src_index=ss_d, per_frame_max=True, check_range=flag,
                                             out=(col, msk, o_depth, o_flow), mask_u8=u8)

        def timed(fn, steps):
            for _ in range(3):
                fn()
            barrier()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            for _ in range(steps):
                fn()
            b.record()
            torch.cuda.synchronize()
            barrier()
            return max_over_ranks(a.elapsed_time(b) / steps)

        ksteps = max(1, args.steps)
        ms_strong = timed(lambda: warp_into(o_col, o_msk, False), ksteps)
        gathered = [None]
        def nccl_step():
            warp_into(p_col, p_msk, True)
            gathered[0] = gather_packed(packed, n_all, H, W, dst=0)
        ms_nccl = timed(nccl_step, ksteps) if equal_shards else None      # the single packed gather needs equal shards
        ms_peer = timed(lambda: warp_into(peer.colour, peer.mask, True, scatter=True), ksteps)
        # the assembled clip on rank 0 must equal what one GPU computes for the whole clip, bit for bit (fixed summation order)
        same = None
        if rank == 0:
            pick = torch.tensor([0, n_all // 2 - 1, n_all // 2, n_all - 1])
            ref_c, ref_m, _, _ = warper.forward_warp_matrices(sf.to(dev), sd.to(dev), sM[pick].to(dev), src_index=ssrc[pick].to(dev),
                                                              per_frame_max=True, check_range=flag, mask_u8=True)
            same = bool(torch.equal(peer.colour[pick.to(dev)], ref_c) and torch.equal(peer.mask[pick.to(dev)], ref_m))
            if ms_nccl is not None:
                parts = unpack_gathered(gathered[0], n_all, H, W)
                where = torch.empty(n_all, dtype=torch.int64)
                where[order] = torch.arange(n_all)                   # gathered position of every unit (rank-major blocks)
                g_col = torch.cat([p_[0] for p_ in parts])[where[pick].to(dev)]
                same = same and bool(torch.equal(g_col, ref_c))
        wire = 13 * (n_all - balanced_units(n_all, N_FRAMES, world, 0).numel()) * H * W
        px_all = n_all * H * W
        strong = {"value": px_all / (ms_strong * 1e-3) / 1e6, "unit": "Mpix/s", "ms": ms_strong, "units_per_rank": nl,
                  "what": "config 2's 392 units of ONE clip sharded over the ranks (a frame block of every trajectory per rank: "
                          "sharding.balanced_units), outputs left on each rank, no collective"}
        with_gather = {"value": px_all / (ms_peer * 1e-3) / 1e6, "unit": "Mpix/s", "ms": ms_peer,
                       "ingress_GBps": wire / (ms_peer * 1e-3) / 1e9, "bytes_into_rank0": wire,
                       "nvlink_bound_ms": wire / 770e9 * 1e3,
                       "how": "fused: the gather kernel of every rank stores colour f32 + mask u8 (13 B/px) of each of its units straight "
                              "into that unit's slot of rank 0's buffer (tcw_forward_warp_scatter over a CUDA IPC peer mapping, NVLink / "
                              "NVSwitch); the step ends when all ranks' kernels have finished",
                       "nccl_single_gather": None if ms_nccl is None else {
                           "value": px_all / (ms_nccl * 1e-3) / 1e6, "ms": ms_nccl, "ingress_GBps": wire / (ms_nccl * 1e-3) / 1e9,
                           "how": "warp into a per-rank packed buffer, then ONE dist.gather of 13 B/px (rank-major blocks: "
                                  "the clip order still needs a permutation on rank 0, not timed)"},
                       "assembled_equals_single_gpu_bits": same}
        if rank != 0:
            peer.close()          # unmap before the owner frees
        barrier()
        if rank == 0:
            peer.close()
    gather = None

    # ---- next row of SURVEY.md §8f (N1), reported beside the metric, never part of it: the resize + pack kernel on
    # the reference's own warp resolution, 49 x 576x1024 -> 384x672 (demo.py:91-100,117-120)
    pack = None
    if rank == 0 and world == 1 and not args.no_extras:
        Hs, Ws = 576, 1024
        pf = torch.rand((N_FRAMES, 3, Hs, Ws), device=dev) * 2 - 1
        pm = (torch.rand((N_FRAMES, 1, Hs, Ws), device=dev) > 0.2).float()
        for _ in range(3):
            warper.pack_condition(pf, pm, (H, W))
        p0, p1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        p0.record()
        for _ in range(10):
            warper.pack_condition(pf, pm, (H, W))
        p1.record()
        torch.cuda.synchronize()
        p_ms = p0.elapsed_time(p1) / 10
        p_bytes = N_FRAMES * (Hs * Ws * 16 + H * W * 16)
        pack = {"kernel": "k_resize_pack", "ms": p_ms, "shape": f"{N_FRAMES}x{Hs}x{Ws} -> {H}x{W}", "algorithmic_bytes": p_bytes,
                "achieved_GBps": p_bytes / (p_ms * 1e-3) / 1e9, "frac_of_hbm_peak": p_bytes / (p_ms * 1e-3) / 1e9 / peak,
                "note": "includes the two torch.empty output allocations of the Python entry; 462 MB in + 202 MB out, larger than L2"}
        del pf, pm
        # the same lines fused behind the warp (tcw_forward_warp_pack) against the two calls, on the reference's own shapes:
        # 49 frames warped at 576x1024 (demo.py:545-547), condition video at 384x672
        from trajectorycrafter_b200.workloads import orbit_units as _orb, synthetic_clip as _clip
        bf, bd = _clip(N_FRAMES, Hs, Ws, seed=5)
        bT1, bT2, bK, _ = _orb(bd, Hs, Ws, presets=[(0, 30)])
        bf, bd = bf.to(dev), bd.to(dev)
        bflag = torch.zeros(1, dtype=torch.int32, device=dev)

        def two_calls():
            wf, wm, _, _ = warper.forward_warp_batch(bf, bd, bT1, bT2, bK, check_range=bflag)
            return warper.pack_condition(wf, wm, (H, W))

        def fused():
            return warper.forward_warp_pack(bf, bd, bT1, bT2, bK, None, (H, W), check_range=bflag)
        res = {}
        for name, fn in (("warp_then_pack", two_calls), ("fused", fused)):
            for _ in range(3):
                r_ = fn()
            torch.cuda.synchronize()
            p0.record()
            for _ in range(10):
                r_ = fn()
            p1.record()
            torch.cuda.synchronize()
            res[name] = p0.elapsed_time(p1) / 10
        same = bool(torch.equal(two_calls()[0], fused()[0]) and torch.equal(two_calls()[1], fused()[1]))
        px = N_FRAMES * Hs * Ws
        pack["fused_behind_warp"] = {"warp_then_pack_ms": res["warp_then_pack"], "fused_ms": res["fused"],
                                     "Mpix_per_s_fused": px / (res["fused"] * 1e-3) / 1e6, "bit_identical": same,
                                     "result_bytes_per_warped_px": {"warp_then_pack": 28 + 16 * H * W / (Hs * Ws), "fused": 16 * H * W / (Hs * Ws)},
                                     "what": f"{N_FRAMES} units warped at {Hs}x{Ws}, condition video + masks at {H}x{W}; device time incl. the "
                                             "host-side matrix preparation and output allocation of the Python entries"}
        del bf, bd, r_

    # ---- the other SURVEY 8(f) rows, reported beside the metric, never part of it
    extras = None
    if rank == 0 and world == 1 and not args.no_extras:
        extras = {}
        from trajectorycrafter_b200 import GlobalPointCloudWarper
        from trajectorycrafter_b200.workloads import orbit_units as _ou

        def dev_time(fn, reps):
            fn(); torch.cuda.synchronize()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            for _ in range(reps):
                fn()
            b.record()
            torch.cuda.synchronize()
            return a.elapsed_time(b) / reps
        # A6 / N3: forward_warp(twice=True) = five chained splats (models/utils.py:296-347), one 384x672 frame per call
        tT1, tT2, tK, _ = _ou(depths[:1], H, W, presets=[(0, 12)])
        f1, d1 = frames_d[:1], depths_d[:1]
        t_ms = dev_time(lambda: warper.forward_warp(f1, None, d1, tT1, tT2, tK, None, twice=True), 5)
        extras["twice"] = {"ms_per_frame": t_ms, "Mpix_per_s": H * W / (t_ms * 1e-3) / 1e6,
                           "what": "Warper.forward_warp(twice=True), b = 1, 384x672: forward colour + depth splat, flow splat, two backward "
                                   "splats, incl. the host-side matrix preparation of each call"}
        # the same as ONE call with b = 49 (one trajectory; the reference's call-wide normaliser): 4 splats of 49 units + a projection
        bT1, bT2, bK, _ = _ou(depths, H, W, presets=[(0, 12)])
        tb_ms = dev_time(lambda: warper.forward_warp(frames_d, None, depths_d, bT1, bT2, bK, None, twice=True), 5)
        tb_bytes = N_FRAMES * H * W * (44 + 36 + 44 + 28)      # colour+depth warp 44, flow splat (2 ch) 36, colour back-splat 44, depth back-splat 28 B/px
        extras["twice_b49"] = {"ms": tb_ms, "Mpix_per_s": N_FRAMES * H * W / (tb_ms * 1e-3) / 1e6,
                               "algorithmic_bytes": tb_bytes, "frac_of_hbm_peak": tb_bytes / (tb_ms * 1e-3) / 1e9 / peak,
                               "what": "Warper.forward_warp(twice=True) with b = 49 at 384x672 in one call: forward warp, flow splat, two backward "
                                       "splats on the window kernels (fixed summation order), plus the torch negation of the warped flow"}
        # N2: z-buffer renderer, 1 M points -> 576x1024 (point_size 2: nine footprint entries per point)
        pc = GlobalPointCloudWarper(device=f"cuda:{local}")
        g = torch.Generator().manual_seed(9)
        n_pts = 1_000_000
        P = (torch.rand((n_pts, 3), generator=g) * torch.tensor([8.0, 5.0, 6.0]) - torch.tensor([4.0, 2.5, -4.0])).to(dev)
        C = (torch.rand((n_pts, 3), generator=g) * 2 - 1).to(dev)
        Tt = torch.eye(4)[None]
        Kt = torch.tensor([[500.0, 0, 512.0], [0, 500.0, 288.0], [0, 0, 1]])[None]
        z_ms = dev_time(lambda: pc.render_pointcloud_zbuffer_vectorized_point_size(P, C, Tt, Kt, (576, 1024), 2), 20)
        z_bytes = n_pts * 24 + 576 * 1024 * (8 + 8 + 16)          # points + colours in; z-buffer cleared + resolved, frame + mask out
        extras["pointcloud_zbuffer"] = {"ms": z_ms, "Mpoints_per_s": n_pts / (z_ms * 1e-3) / 1e6, "algorithmic_bytes": z_bytes,
                                        "achieved_GBps": z_bytes / (z_ms * 1e-3) / 1e9, "frac_of_hbm_peak": z_bytes / (z_ms * 1e-3) / 1e9 / peak,
                                        "what": "render_pointcloud_zbuffer_vectorized_point_size, 1 M points, point_size 2, 576x1024, incl. the "
                                                "Python entry; 9 M 64-bit atomicMin per call, so the bound is L2 atomics, not HBM"}
        # N4: depth post-processing of one 49x576x1024 clip (the reference's native depth resolution)
        res = torch.rand((N_FRAMES, 576, 1024, 3), generator=g).to(dev)
        p_ms = dev_time(lambda: warper.depth_postprocess(res), 10)
        p_bytes = N_FRAMES * 576 * 1024 * (12 + 12 + 4)
        extras["depth_post"] = {"ms": p_ms, "algorithmic_bytes": p_bytes, "achieved_GBps": p_bytes / (p_ms * 1e-3) / 1e9,
                                "frac_of_hbm_peak": p_bytes / (p_ms * 1e-3) / 1e9 / peak,
                                "what": "tcw_depth_post on 49x576x1024x3: two reads of the raw output (min/max pass, then the post-processing), one write"}
        del P, C, res

    # ---- CPU baseline (rank 0, N=1 only), bounded samples of the same workload on the box's host cores:
    # the reference's own torch path (oracle/_ref, BASELINE.md section 5: all cores default / deterministic / one thread)
    # and the C port of the same arithmetic (all cores, one full step)
    cpu_baseline = cpu_ref = cpu_port = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        import oracle
        from oracle import ref_loader
        oracle.build()
        threads = os.cpu_count() or 1
        v, secs = oracle_throughput(frames, depths, M, src, list(range(n_units)), threads)
        cpu_port = {"value": v, "unit": "Mpix/s", "cores": threads, "kind": "port",
                    "sample": f"all 392 units of one step, once ({secs:.1f} s wall on {threads} threads)"}
        cpu_baseline = cpu_port
        if ref_loader.available():
            _, _, _, _, poses = build_workload(1234 + rank, with_poses=True)
            ids = [f * 8 + p for f in (6, 24, 42) for p in range(8)]              # 24 units: 3 source frames x 8 presets
            v_def, s_def = reference_torch_throughput(frames, depths, poses, src, ids, threads, False)
            v_det, s_det = reference_torch_throughput(frames, depths, poses, src, ids[:16], threads, True)
            v_one, s_one = reference_torch_throughput(frames, depths, poses, src, ids[:8], 1, False)
            cpu_ref = {"value": v_def, "unit": "Mpix/s", "cores": threads, "kind": "reference",
                       "sample": f"24 of 392 units (source frames 6/24/42 x 8 presets), b=1 calls as demo.py:77-90, {s_def:.1f} s wall",
                       "modes": {"all_cores_default": {"Mpix_per_s": v_def, "units": 24, "seconds": s_def},
                                 "all_cores_deterministic": {"Mpix_per_s": v_det, "units": 16, "seconds": s_det},
                                 "one_thread": {"Mpix_per_s": v_one, "units": 8, "seconds": s_one}},
                       "host": host_description(), "torch": torch.__version__, "torch_threads": threads,
                       "what": "the UNMODIFIED reference Warper(device='cpu') (verbatim run-time copy of models/utils.py in oracle/_ref/)"}
            cpu_baseline = {k: cpu_ref[k] for k in ("value", "unit", "cores", "kind", "sample")}

    if rank == 0:
        line = {
            "metric": "warped_mpix_per_s", "value": value, "unit": "Mpix/s", "n_gpus": world,
            "steps": args.steps, "warmup": max(3, args.warmup), "ms_per_step": ms, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": workload_config(), "clocks": clocks, "e2e": e2e, "gpu_launches": int(launches),
            "roofline": roofline, "cpu_baseline": cpu_baseline,
            "details": {"accumulation": ACCUMULATION[args.accumulation], "host_enqueue_ms_per_step": round(host_ms, 3)},
        }
        if cpu_ref is not None:
            line["cpu_baseline_reference"] = cpu_ref
        if cpu_port is not None:
            line["cpu_baseline_port"] = cpu_port
        if other_modes is not None:
            line["other_modes"] = other_modes
        if clip49 is not None:
            line["clip49"] = clip49
        if dropin is not None:
            line["dropin"] = dropin
        if extras is not None:
            line["next_rows"] = extras
        if strong is not None:
            line["strong"] = strong
            line["with_gather"] = with_gather
        if numa is not None:
            line["details"]["cpu_affinity_first_last_count"] = numa
        if per_rank is not None:
            line["per_rank"] = per_rank
        if pack is not None:
            line["pack_epilogue"] = pack
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--accumulation", default="window", choices=sorted(ACCUMULATION))
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-extras", action="store_true", help="skip the clip49 / drop-in / pack-epilogue legs")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
