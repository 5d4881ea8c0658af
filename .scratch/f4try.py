import sys, numpy as np
sys.path.insert(0,'.'); sys.path.insert(0,'tests')
import f4, synth
from oracle import ref, port
codes, ac = f4.write_multiallelic('.scratch/ma.pgen', 300, 1000, 3)
with ref.RefPgen('.scratch/ma.pgen') as r:
    print("V,N", r.V, r.N, "vrtypes", sorted(set(r.vrtypes().tolist())))
    print("counts ok", np.array_equal(r.counts(), synth.counts_numpy(codes)))
    print("i8 ok", np.array_equal(r.decode_i8(), np.where(codes==3,-9,codes)))
g = f4.write_dphase('.scratch/dp.pgen', 90, 500, 4)
with ref.RefPgen('.scratch/dp.pgen') as r:
    print("dphase vrtypes", sorted(set(r.vrtypes().tolist())))
    print("counts ok", np.array_equal(r.counts(), synth.counts_numpy(g)))
    d = r.dosage(2); print(d[:8])
for pth in ('.scratch/ma.pgen','.scratch/dp.pgen'):
    try:
        with port.PortPgen(pth) as p: print("port counts ok", pth, np.array_equal(p.counts(), r.counts()) if False else p.counts().shape)
    except Exception as e: print("port", pth, e)
