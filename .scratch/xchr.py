import sys, math
sys.path.insert(0,'.')
from oracle import ref
import mpmath as mp
mp.mp.dps=50
def xchr_exact(f0,fh,f2,ma,mb,midp=False):
    nf=f0+fh+f2; nm=ma+mb; nB=2*f2+fh+mb; 
    def lp(m,h):
        fb=nB-m
        if fb<0 or (fb-h)%2 or h>fb: return None
        g2=(fb-h)//2; g0=nf-h-g2
        if g0<0 or m>nm or m<0: return None
        return h*mp.log(2)-mp.loggamma(m+1)-mp.loggamma(nm-m+1)-mp.loggamma(g0+1)-mp.loggamma(h+1)-mp.loggamma(g2+1)
    lo=lp(mb,fh)
    tail=mp.mpf(0); tot=mp.mpf(0); ties=0
    for m in range(0,nm+1):
        for h in range(0,nf+1):
            l=lp(m,h)
            if l is None: continue
            t=mp.e**(l-lo); tot+=t
            if t<=1+mp.mpf(2)**-30:
                tail+=t
                if t>=1-mp.mpf(2)**-30: ties+=1
    if midp: tail-=mp.mpf(ties)/2
    return float(tail/tot)
t=(64,8,65,110,69)
print("exact", xchr_exact(*t), "ref", math.exp(ref.hwe_xchr_lnp(t[1],t[0],t[2],t[3],t[4],False)))
