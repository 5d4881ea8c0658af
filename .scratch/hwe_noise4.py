import sys, math, numpy as np
sys.path.insert(0,'.')
from oracle import port, ref
exec(open('.scratch/hwe_noise2.py').read().split("rng=np.random.default_rng(3)")[0])
rng=np.random.default_rng(4)
for N in (1000, 500000):
    ulp = np.spacing(2*N*math.log(2*N))
    A=[]
    for (h,a,b) in cases(N, rng, 600):
        for midp in (True,):
            ex = port.hwe_lnp_exact(h,a,b,midp); rf = ref.hwe_lnp(h,a,b,midp)
            if ex < math.log(2.0**-10): A.append((abs(rf-ex)/ulp, h,a,b,ex,rf))
    A.sort(reverse=True)
    for x in A[:8]: print(N, "%.1f ulp hets=%d hom1=%d hom2=%d exact=%.12g ref=%.12g"%x)
    print("median", np.median([x[0] for x in A]))
