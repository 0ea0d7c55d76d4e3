"""SHA-256 of the four outputs of a fixed 64-unit submission (every orbit preset): two builds of the library that claim
bit-identical sums must print the same digests.  usage: [TCW_LIB=...] python scripts/out_digest.py"""
import hashlib, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from trajectorycrafter_b200 import Warper, prepare_matrices
from trajectorycrafter_b200.workloads import orbit_units, synthetic_clip

h, w, nf = 384, 672, 49
frames, depths = synthetic_clip(nf, h, w, seed=1234)
T1, T2, K, src = orbit_units(depths, h, w)
sel = torch.arange(3, 392, 6)[:64]
M = prepare_matrices(T1[sel], T2[sel], K[sel])
wp = Warper(device="cuda:0")
out = wp.forward_warp_matrices(frames.cuda(), depths.cuda(), M, src_index=src[sel].cuda())
torch.cuda.synchronize()
print(os.environ.get("TCW_LIB", "default"), [hashlib.sha256(t.cpu().numpy().tobytes()).hexdigest()[:16] for t in out])
