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
NCU_TRAFFIC_BYTES_PER_LAUNCH = 238_216_960   # profiles/r1_final_kernels_full_24units_ncu.txt: 109.06 MB read + 129.16 MB written


ACCUMULATION = {"window": "fixed order: warp-private shared-memory windows + ordered gather (default, reproducible)",
                "atomic": "fp32 L2 reductions (order varies run to run)", "integer": "int64 fixed point (order-free)"}


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


def build_workload(seed):
    import torch
    from trajectorycrafter_b200 import prepare_matrices
    from trajectorycrafter_b200.workloads import orbit_units, synthetic_clip
    frames, depths = synthetic_clip(N_FRAMES, H, W, seed=seed)
    T1, T2, K, src = orbit_units(depths, H, W)
    M = prepare_matrices(T1, T2, K)
    return frames, depths, M, src.to(torch.int32)


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
    """--impl reference: the reference's own CPU implementation of the path, timed on the host cores.
    The reference is pure Python/torch and cannot travel to the GPU box, so this is the oracle port
    (kind "port"), threaded over units with every host core."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    import oracle
    oracle.build()
    frames, depths, M, src = build_workload(1234)
    threads = os.cpu_count() or 1
    # bounded sample per step: every 6th frame x 8 presets = 72 of the 392 units
    unit_ids = [f * 8 + p for f in range(0, N_FRAMES, 6) for p in range(8)]
    for _ in range(max(1, min(args.warmup, 2))):
        oracle_throughput(frames, depths, M, src, unit_ids[:threads], threads)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        oracle_throughput(frames, depths, M, src, unit_ids, threads)
    dt = (time.perf_counter() - t0) / args.steps
    value = len(unit_ids) * H * W / dt / 1e6
    sample = f"{len(unit_ids)} of 392 units per step (every 6th frame x 8 orbit presets), {threads} threads"
    line = {
        "impl": "reference", "metric": "warped_mpix_per_s", "value": value, "unit": "Mpix/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt * 1e3,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": workload_config(),
        "cpu_baseline": {"value": value, "unit": "Mpix/s", "cores": threads, "kind": "port", "sample": sample},
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
        # headline: the kernel timed ALONE (pass B) -- a roofline fraction describes one kernel owning the GPU;
        # "overlapped" repeats the figure for the launches of the timed 4-lane schedule, which share the SMs with
        # the other lanes' kernels, and "step_frac" is the whole timed step against the same peak
        def ach(ms_, units_):
            return ALGO_BYTES_PER_PX * units_ * H * W / (ms_ * 1e-3) / 1e9
        kname = {"window": "k_splat_window", "atomic": "k_splat_pair", "integer": "k_splat_pair_det<i64>"}[args.accumulation]
        tot_i, tot_k = sum(v[0] for v in iso.values()), sum(v[0] for v in prof.values())
        roofline = {"bound": "hbm", "kernel": kname, "achieved": ach(i_ms, i_units), "peak": peak, "unit": "GB/s",
                    "frac": ach(i_ms, i_units) / peak, "traffic": NCU_TRAFFIC_BYTES_PER_LAUNCH,
                    "peak_source": peak_src, "avg_launch_us": i_ms / i_n * 1e3, "launches": i_n,
                    "units_per_launch": i_units / i_n,
                    "algorithmic_bytes_per_launch": ALGO_BYTES_PER_PX * i_units * H * W / i_n,
                    "kernel_share_of_step": i_ms / tot_i if tot_i else None,
                    "per_kernel_ms_per_step": {k: v[0] / iso_steps for k, v in iso.items()},
                    "how": "CUDA events on the launching stream around every launch, inside bench.py after the timed "
                           "region; frac/achieved: the kernel alone on the GPU (one stream, 24-unit launches); "
                           "overlapped: the launches of the timed schedule (4 streams x 3-unit launches run "
                           "concurrently, so a launch's duration includes sharing the SMs and the per-kernel times add "
                           "up to more than the step); step_frac: whole timed step",
                    "overlapped": {"achieved": ach(k_ms, k_units), "frac": ach(k_ms, k_units) / peak,
                                   "avg_launch_us": k_ms / k_n * 1e3, "launches": k_n, "units_per_launch": k_units / k_n,
                                   "kernel_share_of_step": k_ms / tot_k if tot_k else None,
                                   "per_kernel_ms_per_step": {k: v[0] / prof_steps for k, v in prof.items()}},
                    "step_frac": ALGO_BYTES_PER_PX * px_per_step / (ms_local * 1e-3) / 1e9 / peak}

    # ---- e2e: host buffers through the C-ABI host entry (H2D + kernels + D2H inside the timed region)
    e2e = None
    if not args.no_e2e:
        hf, hd = frames.pin_memory(), depths.pin_memory()
        houts = (torch.empty((n_units, 3, H, W)).pin_memory(), torch.empty((n_units, 1, H, W)).pin_memory(),
                 torch.empty((n_units, 1, H, W)).pin_memory(), torch.empty((n_units, 2, H, W)).pin_memory())
        Mh = M.contiguous()
        warper.forward_warp_host(hf, hd, Mh, src_index=src, outputs=houts)
        barrier()
        k_e2e = max(1, min(args.steps, 5))
        t0 = time.perf_counter()
        for _ in range(k_e2e):
            res = warper.forward_warp_host(hf, hd, Mh, src_index=src, outputs=houts)
        dt = (time.perf_counter() - t0) / k_e2e
        assert res[4] == 0
        dt = max_over_ranks(dt)
        e2e = {"value": world * px_per_step / dt / 1e6, "unit": "Mpix/s",
               "h2d_bytes_per_step": int(hf.numel() * 4 + hd.numel() * 4 + Mh.numel() * 4 + src.numel() * 4),
               "d2h_bytes_per_step": int(sum(t.numel() * 4 for t in houts) + 4),
               "ms_per_step": dt * 1e3, "steps": k_e2e,
               "api": "Warper.forward_warp_host -> tcw_forward_warp_host (pinned host buffers)"}
        del houts

    # ---- the path's one exchange step: gather colour+mask of one 49-unit trajectory per rank on rank 0
    gather = None
    if world > 1:
        from trajectorycrafter_b200.sharding import gather_condition_video
        n_g = N_FRAMES
        col, msk = out[0][:n_g].contiguous(), out[1][:n_g].contiguous()
        for _ in range(2):
            gather_condition_video(col, msk, n_g * world, dst=0)
        barrier()
        g0, g1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        g0.record()
        for _ in range(5):
            gather_condition_video(col, msk, n_g * world, dst=0)
        g1.record()
        torch.cuda.synchronize()
        g_ms = max_over_ranks(g0.elapsed_time(g1) / 5)
        wire = n_g * H * W * 13 * (world - 1)
        gather = {"ms": g_ms, "bytes_into_rank0": wire, "ingress_GBps": wire / (g_ms * 1e-3) / 1e9,
                  "what": "NCCL gather of colour f32 + mask u8 (13 B/px) of 49 units per rank to rank 0"}

    # ---- next row of SURVEY.md §8f (N1), reported beside the metric, never part of it: the resize + pack kernel on
    # the reference's own warp resolution, 49 x 576x1024 -> 384x672 (demo.py:91-100,117-120)
    pack = None
    if rank == 0 and world == 1:
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

    # ---- CPU baseline (rank 0, N=1 only): the oracle port on the host cores, one full step once
    cpu_baseline = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        import oracle
        oracle.build()
        threads = os.cpu_count() or 1
        v, secs = oracle_throughput(frames, depths, M, src, list(range(n_units)), threads)
        cpu_baseline = {"value": v, "unit": "Mpix/s", "cores": threads, "kind": "port",
                        "sample": f"all 392 units of one step, once ({secs:.1f} s wall on {threads} threads)"}

    if rank == 0:
        line = {
            "metric": "warped_mpix_per_s", "value": value, "unit": "Mpix/s", "n_gpus": world,
            "steps": args.steps, "warmup": max(3, args.warmup), "ms_per_step": ms, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": dict(workload_config(), accumulation=ACCUMULATION[args.accumulation], host_enqueue_ms_per_step=round(host_ms, 3)), "clocks": clocks, "e2e": e2e, "gpu_launches": int(launches),
            "roofline": roofline, "cpu_baseline": cpu_baseline,
        }
        if gather is not None:
            line["gather"] = gather
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
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
