import numpy as np
from scipy.special import erfc, erf
# gelu(x) = max(x,0) - |x| * 2^(q(|x|))  where 2^q(a) = 0.5*erfc(a/sqrt2)
A=6.0
a = np.linspace(0, A, 200001)
target = np.log2(0.5*erfc(a/np.sqrt(2)))
w = a*0.5*erfc(a/np.sqrt(2)) + 1e-4   # abs err of gelu = a * h * ln2 * dq
import sys
for deg in tuple(int(a) for a in sys.argv[1:]) or (5,6,7,8):
    # weighted LSQ in Chebyshev basis then iterate reweighting (Lawson) for minimax
    t = 2*a/A-1
    V = np.polynomial.chebyshev.chebvander(t, deg)
    wl = np.ones_like(a)
    for it in range(60):
        W = (w*wl)
        c,*_ = np.linalg.lstsq(V*W[:,None], target*W, rcond=None)
        err = np.abs((V@c-target)*w)
        wl = wl*(err/err.max()+0.05); wl/=wl.max()
    # convert to monomial in a
    p = np.polynomial.chebyshev.Chebyshev(c, domain=[0,A]).convert(kind=np.polynomial.Polynomial)
    co = p.coef.astype(np.float32)
    # evaluate in fp32 Horner
    x = np.linspace(-8,8,400001).astype(np.float32)
    ax = np.minimum(np.abs(x), np.float32(A))
    q = np.full_like(ax, co[-1])
    for k in range(deg-1,-1,-1):
        q = (q*ax + co[k]).astype(np.float32)
    h = np.exp2(q).astype(np.float32)
    g = (np.maximum(x,0) - np.abs(x)*h).astype(np.float32)
    ref = 0.5*x.astype(np.float64)*(1+erf(x.astype(np.float64)/np.sqrt(2)))
    e = np.abs(g-ref)
    rel = e/np.maximum(np.abs(ref),1e-3)
    print(deg, "max abs", e.max(), "at", x[e.argmax()], "max rel(>1e-3)", rel.max())
    print("  coefs", [f"{v:.9e}" for v in co])
