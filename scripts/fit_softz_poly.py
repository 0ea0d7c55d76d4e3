import numpy as np
from mpmath import mp, mpf, log1p, exp as mexp, log as mlog
mp.prec = 120
f32 = np.float32
def fma(a,b,c):
    return (np.asarray(a,f32).astype(np.float64)*np.asarray(b,f32).astype(np.float64)+np.asarray(c,f32).astype(np.float64)).astype(f32)
def minimax_fit(fun, wfun, lo, hi, deg, n=3000, iters=60):
    x = np.cos(np.pi*(np.arange(n)+0.5)/n)*(hi-lo)/2+(hi+lo)/2
    x = x[np.abs(x)>1e-5]
    y = np.array([float(fun(mpf(float(v)))) for v in x]); s = wfun(x)
    A = np.vander(x, deg+1, increasing=True); wgt = s.copy()
    best=None
    for it in range(iters):
        c, *_ = np.linalg.lstsq(A*wgt[:,None], y*wgt, rcond=None)
        err = np.abs((A@c - y)*s); m = err.max()
        if best is None or m < best[1]: best=(c,m)
        wgt = wgt*(1+0.7*err/m)
    return best
cl, el = minimax_fit(lambda v: (log1p(v)-v+v*v/2)/v**3, lambda x: np.abs(x)**3, -1/3, 1/3, 7)
ce, ee = minimax_fit(lambda v: (mexp(v)-1-v)/v**2, lambda x: x*x, -0.3466, 0.3466, 5)
print("log err", el, "exp err", ee)
LOGC = [f32(v) for v in cl]; EXPC = [f32(v) for v in ce]
print("LOGC", [v.item().hex() for v in LOGC]); print("LOGC", [repr(float(v)) for v in LOGC])
print("EXPC", [repr(float(v)) for v in EXPC])
LN2 = f32(0.6931471805599453)
def my_log(a):
    a = np.asarray(a, f32); bits = a.view(np.int32)
    i = bits - np.int32(0x3f2aaaab); e = i >> 23
    m = (bits - (e << 23)).view(f32); f = (m - f32(1)).astype(f32)
    q = np.full_like(f, LOGC[7])
    for k in range(6,-1,-1): q = fma(q, f, np.full_like(f, LOGC[k]))
    f2 = (f*f).astype(f32); t = fma(f, q, np.full_like(f, -0.5)); r = fma(f2, t, f)
    return fma(e.astype(f32), np.full_like(f, LN2), r)
a = np.concatenate([np.random.default_rng(0).uniform(1,1001,2_000_000), np.random.default_rng(1).uniform(1,30,2_000_000), 1+np.random.default_rng(2).uniform(0,1,500000)**3]).astype(f32)
ref = np.log(a.astype(np.float64)); got = my_log(a).astype(np.float64)
ulp = np.spacing(np.abs(ref).astype(f32)).astype(np.float64)
print("log max abs err", np.abs(got-ref).max(), "max ulp err", (np.abs(got-ref)/ulp).max(), "vs correctly rounded mismatch frac", np.mean(got.astype(f32)!=ref.astype(f32)))
L2E = f32(1.4426950408889634); LN2HI = f32(0.693145751953125); LN2LO = f32(1.42860682e-6); MAGIC=f32(12582912.0)
def my_exp(x):
    x = np.asarray(x,f32)
    jf = (fma(x, np.full_like(x,L2E), np.full_like(x,MAGIC)) - MAGIC).astype(f32)
    f = fma(jf, np.full_like(x,-LN2HI), x); f = fma(jf, np.full_like(x,-LN2LO), f)
    q = np.full_like(f, EXPC[5])
    for k in range(4,-1,-1): q = fma(q, f, np.full_like(f, EXPC[k]))
    f2=(f*f).astype(f32); r = fma(f2, q, f); r = (r + f32(1)).astype(f32)
    return ((r.view(np.int32) + (jf.astype(np.int32) << 23)).view(f32))
x = -np.random.default_rng(3).uniform(0,50,3_000_000).astype(f32)
ref = np.exp(x.astype(np.float64)); got = my_exp(x).astype(np.float64)
print("exp max rel err", (np.abs(got-ref)/ref).max(), " in ulps ~", (np.abs(got-ref)/ref).max()/6e-8)
