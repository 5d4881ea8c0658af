import sys, math, numpy as np
sys.path.insert(0,'.')
from oracle import port, ref
exec(open('.scratch/hwe_noise2.py').read().split("rng=np.random.default_rng(3)")[0])
rng=np.random.default_rng(4)
for N in (100, 1000, 10000, 100003, 500000):
    ulp = np.spacing(2*N*math.log(2*N))
    A=[];B=[]
    for (h,a,b) in cases(N, rng, 600):
        for midp in (False,True):
            ex = port.hwe_lnp_exact(h,a,b,midp); rf = ref.hwe_lnp(h,a,b,midp)
            if ex < math.log(2.0**-10): A.append(abs(rf-ex)/ulp)
            else: B.append(abs(math.exp(rf)-math.exp(ex))/ulp)
    print(N, "ulp=%.3g"%ulp, "tail(p<2^-10): n=%d max|dlnp|=%.1f ulp"%(len(A), max(A)), " p>=2^-10: n=%d max|dp|=%.1f ulp"%(len(B), max(B)))
