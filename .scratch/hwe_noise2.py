import sys, math, numpy as np
sys.path.insert(0,'.')
from oracle import port, ref

def cases(N, rng, n=300):
    out=[]
    for _ in range(n):
        maf = 10**rng.uniform(-4, math.log10(0.5))
        F = rng.choice([0, 0, rng.uniform(-1,1), rng.uniform(-0.2,0.2), rng.uniform(-0.02,0.02)])  # inbreeding coeff
        p=maf; q=1-p
        f0=q*q+F*p*q; f1=2*p*q*(1-F); f2=p*p+F*p*q
        f=np.clip([f0,f1,f2],0,None); f/=f.sum()
        c=rng.multinomial(N, f)
        out.append((int(c[1]),int(c[0]),int(c[2])))
    return out
rng=np.random.default_rng(3)
for N in (1000, 10000, 100003, 500000):
    ulp = np.spacing(2*N*math.log(2*N))
    rows=[]
    for (h,a,b) in cases(N, rng):
        for midp in (False,True):
            ex = port.hwe_lnp_exact(h,a,b,midp)
            rf = ref.hwe_lnp(h,a,b,midp)
            pd = port.hwe_exact(h,a,b,midp)
            rows.append((ex, rf, pd, h,a,b,midp))
    r=np.array([(x[0],x[1],x[2]) for x in rows])
    ex,rf,pd=r[:,0],r[:,1],r[:,2]
    # reference noise in ln p for tail cases (lnp < -1), and in p for p near 1
    tailm = ex < -1
    d_ln = np.abs(rf-ex)
    print(f"N={N} ulp(2N ln 2N)={ulp:.3g}  lnp range [{ex.min():.1f},{ex.max():.3g}]")
    print("   ref |dlnp| (tail) max = %.3g = %.1f ulp ; all-range |dp| max=%.3g" % (d_ln[tailm].max(), d_ln[tailm].max()/ulp, np.abs(np.exp(rf)-np.exp(ex)).max()))
    ok = ex > -700
    with np.errstate(divide='ignore'):
        dl = np.abs(np.log(pd[ok]) - ex[ok])
    dl = dl[np.isfinite(dl)]
    print("   f64 walk |dlnp| max = %.3g (where p>1e-300, n=%d); zeros at lnp>-700: %d" % (dl.max(), ok.sum(), (pd[ok]==0).sum()))
