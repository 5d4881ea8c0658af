"""CPU: the S1 fileset (BASELINE configs[0] size, tests/s1.py) through both oracles -- the reference's pgenlib
(oracle/_ref) and the plain-C restatement (oracle/pgen_oracle.c) must agree with each other and with the
generator on every variant, for the fixed-width file and the pgenlib-written one."""
import numpy as np
import pytest

import s1
import synth


def test_s1_oracles_agree(tmp_path, refmod):
    from oracle import port

    g = s1.genotypes()
    paths = s1.write(str(tmp_path), g, with_pgenlib=True)
    exp = synth.counts_numpy(g)
    miss_s = (g == 3).sum(axis=0).astype(np.uint32)
    for kind in ("fixed", "pgenlib"):
        with refmod.RefPgen(paths[kind]) as r:
            assert np.array_equal(r.counts(), exp)
            pv, ps = r.missingness()
            assert np.array_equal(ps, miss_s) and np.array_equal(pv, exp[:, 3])
        if port.available():
            with port.PortPgen(paths[kind]) as p:
                assert np.array_equal(p.counts(), exp)
                assert np.array_equal(p.missing_by_sample(), miss_s)
    chrom, pos, sex = s1.layout()
    assert set(chrom.tolist()) == {"1", "2", "X", "Y", "MT"} and set(sex.tolist()) == {0, 1, 2}
    # the sex-aware restatement on a hand-checked row (test/sql/plink_sexchr.test:45: x1 -> 0.444444 / 9)
    i8 = np.array([0, 1, 2, 0, 2, 1], dtype=np.int8)  # 3 females (0, 1, 2), 3 males (0, 2, het -> missing)
    sx = np.array([2, 2, 2, 1, 1, 1], dtype=np.uint8)
    r = s1.sex_aware_counts(i8, 1, sx)
    assert (r["obs"], r["alt"], r["missing"]) == (8, 4, 1)
