import sys, numpy as np
sys.path.insert(0,'.'); sys.path.insert(0,'tests')
import hwe_check as hc
from oracle import port
g = hc.load_golden()
for N in (1000,100003,500000):
    c = g[f"n{N}/counts"]; n_obs=c[:,:3].sum(1)
    for midp in (0,1):
        p = np.array([port.hwe_exact(int(x[1]),int(x[0]),int(x[2]),bool(midp)) for x in c])
        rf=g[f"n{N}/ref_lnp_midp{midp}"]; ex=g[f"n{N}/exact_lnp_midp{midp}"]
        # emulate: where exact lnp < ~-690 the f64 port returns 0
        v1=hc.violations_p_vs_reference(p, rf, n_obs); v2=hc.violations_p_vs_exact(p, ex, n_obs)
        print(N, midp, "vs ref:", hc.describe(v1,c,p,rf,ex), "| vs exact:", hc.describe(v2,c,p,rf,ex))
        # perturbation: 1e-9 relative at p ~ 1e-40
        i = np.argmin(np.abs(ex - np.log(1e-40)))
        pp = p.copy(); pp[i] *= (1+1e-9)
        print("   table with lnp=%.1f perturbed 1e-9: exact-check catches=%s ref-check catches=%s" % (ex[i], i in hc.violations_p_vs_exact(pp, ex, n_obs), i in hc.violations_p_vs_reference(pp, rf, n_obs)))
