"""Where the time of one reference-style call (b = 1) goes: host matrix preparation, submission, status read."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from trajectorycrafter_b200 import Warper, prepare_matrices
from trajectorycrafter_b200.workloads import orbit_units, synthetic_clip

h, w, nf = 384, 672, 49
frames, depths = synthetic_clip(nf, h, w, seed=3)
T1, T2, K, src = orbit_units(depths, h, w, presets=[(0, 30)])
wp = Warper(device="cuda:0")
f, d = frames.cuda(), depths.cuda()
print("threads", torch.get_num_threads(), "cpus", len(os.sched_getaffinity(0)))

def timeit(name, fn, n=3):
    fn()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(n):
        fn()
    torch.cuda.synchronize()
    print(f"{name:48s} {(time.perf_counter() - t0) / n / nf * 1e6:8.1f} us per call")

timeit("prepare_matrices only", lambda: [prepare_matrices(T1[i:i+1], T2[i:i+1], K[i:i+1], None) for i in range(nf)])
Ms = [prepare_matrices(T1[i:i+1], T2[i:i+1], K[i:i+1], None) for i in range(nf)]
timeit("forward_warp_matrices, check_range=False", lambda: [wp.forward_warp_matrices(f[i:i+1], d[i:i+1], Ms[i], check_range=False) for i in range(nf)])
timeit("forward_warp_matrices, check_range=True", lambda: [wp.forward_warp_matrices(f[i:i+1], d[i:i+1], Ms[i]) for i in range(nf)])
timeit("forward_warp (reference signature)", lambda: [wp.forward_warp(f[i:i+1], None, d[i:i+1], T1[i:i+1], T2[i:i+1], K[i:i+1], None) for i in range(nf)])
Tc1, Tc2, Kc = T1.cuda(), T2.cuda(), K.cuda()
timeit("forward_warp, matrices on the GPU (demo.py)", lambda: [wp.forward_warp(f[i:i+1], None, d[i:i+1], Tc1[i:i+1], Tc2[i:i+1], Kc[i:i+1], None) for i in range(nf)])
for nt in (1, 4):
    torch.set_num_threads(nt)
    timeit(f"forward_warp, torch.set_num_threads({nt})", lambda: [wp.forward_warp(f[i:i+1], None, d[i:i+1], T1[i:i+1], T2[i:i+1], K[i:i+1], None) for i in range(nf)])
    timeit(f"prepare_matrices only, {nt} threads", lambda: [prepare_matrices(T1[i:i+1], T2[i:i+1], K[i:i+1], None) for i in range(nf)])

# the bench process: a 392-unit submission (large workspace, graph replay) before the loop, CPU affinity of the GPU
import bench
fr, dp, M, src392, _ = bench.build_workload(1234)[:5] if False else (None,) * 5
T1a, T2a, Ka, srca = orbit_units(depths, h, w)
Ma = prepare_matrices(T1a, T2a, Ka)
sa = srca.cuda()
out = wp.forward_warp_matrices(f, d, Ma, src_index=sa, check_range=False)
for _ in range(3):
    wp.forward_warp_matrices(f, d, Ma, src_index=sa, check_range=False, out=out)
torch.cuda.synchronize()
timeit("forward_warp after 392-unit submissions", lambda: [wp.forward_warp(f[i:i+1], None, d[i:i+1], T1[i:i+1], T2[i:i+1], K[i:i+1], None) for i in range(nf)])
print("bind", bench.bind_to_gpu_cpus(0), "cpus now", len(os.sched_getaffinity(0)))
timeit("forward_warp after binding to the GPU's CPUs", lambda: [wp.forward_warp(f[i:i+1], None, d[i:i+1], T1[i:i+1], T2[i:i+1], K[i:i+1], None) for i in range(nf)])
del out
timeit("forward_warp, big outputs freed", lambda: [wp.forward_warp(f[i:i+1], None, d[i:i+1], T1[i:i+1], T2[i:i+1], K[i:i+1], None) for i in range(nf)])
