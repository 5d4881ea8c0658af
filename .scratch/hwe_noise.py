import sys, math, numpy as np
sys.path.insert(0,'.')
from oracle import port, ref
import mpmath as mp
mp.mp.dps = 60

def mp_lnp(hets, h1, h2, midp=False):
    # independent: log-gamma based log-probabilities in 60 digits, direct summation
    r0, c0 = min(h1,h2), max(h1,h2)
    n = hets+h1+h2
    rare = 2*r0+hets
    if rare < 2: return math.log(0.5) if midp else 0.0
    lo_k = rare & 1
    def lp(k):
        r=(rare-k)//2; c = n - k - r
        return k*mp.log(2) - mp.loggamma(k+1)-mp.loggamma(r+1)-mp.loggamma(c+1)
    l_obs = lp(hets)
    ks = range(lo_k, rare+1, 2)
    tail = mp.mpf(0); tot=mp.mpf(0); ties=0
    for k in ks:
        d = lp(k)-l_obs
        t = mp.e**d
        tot += t
        if d <= mp.log(1+mp.mpf(2)**-38):
            tail += t
            if d >= mp.log(1-mp.mpf(2)**-38): ties+=1
    if midp: tail -= mp.mpf(ties)/2
    return float(mp.log(tail/tot))

# verify exact oracle vs mpmath on small/medium
rng = np.random.default_rng(1)
worst=0
for _ in range(40):
    n = int(rng.integers(10, 3000))
    h1 = int(rng.integers(0,n)); hets=int(rng.integers(0,n-h1+1)); h2=n-h1-hets
    for midp in (False, True):
        a = port.hwe_lnp_exact(hets,h1,h2,midp); b = mp_lnp(hets,h1,h2,midp)
        err = abs(a-b)/max(1,abs(b))
        worst=max(worst,err)
print("exact oracle vs mpmath worst rel/abs err:", worst)
