"""Known-answer values quoted from the reference's own sqllogictests (test/sql/*.test), shared by the
CPU oracle tests and the GPU parity tests.  Each entry cites the file:line it was read from."""
import numpy as np

# test/sql/plink_freq.test:63-70 (counts) and :16-23 (ALT_FREQ / OBS_CT) -- pgen_example
PGEN_EXAMPLE_COUNTS = np.array([[1, 1, 1, 1], [1, 2, 1, 0], [1, 1, 1, 1], [2, 1, 1, 0]], dtype=np.uint32)
PGEN_EXAMPLE_ALT_FREQ = [0.5, 0.5, 0.5, 0.375]
PGEN_EXAMPLE_OBS_CT = [6, 8, 6, 8]
# test/sql/plink_freq.test:213-229 -- all_missing: every call missing, ALT_FREQ NULL
ALL_MISSING_COUNTS = np.array([[0, 0, 0, 2], [0, 0, 0, 2]], dtype=np.uint32)
# test/sql/plink_hardy.test:20-52 (P_HWE), :84-91 (E_HET), :146-164 (mid-p; rs3 == rs1)
PGEN_EXAMPLE_P_HWE = [1.0, 1.0, 1.0, 0.428571]
PGEN_EXAMPLE_P_HWE_MIDP = [0.7, 0.657143, 0.7, 0.214286]
PGEN_EXAMPLE_O_HET = [0.333333, 0.5, 0.333333, 0.25]
PGEN_EXAMPLE_E_HET = [0.5, 0.5, 0.5, 0.46875]
# test/sql/plink_hardy.test:317-327 -- large_example: every variant (2,2,2 obs of 8? see counts) P = 0.480519
LARGE_EXAMPLE_P_HWE = 0.480519
# test/sql/plink_missing.test:20-23 -- MISSING_CT / OBS_CT / F_MISS per variant of pgen_example
PGEN_EXAMPLE_MISSING = [(1, 3, 0.25), (0, 4, 0.0), (1, 3, 0.25), (0, 4, 0.0)]
# test/sql/plink_missing_sample.test:25 -- SAMPLE2 and SAMPLE4 miss one of four variants
PGEN_EXAMPLE_MISSING_BY_SAMPLE = [0, 1, 0, 1]
# test/sql/plink_missing_sample.test:168 -- large_example: 750 / 2250 / 0.25 for every sample
LARGE_EXAMPLE_SAMPLE_MISSING = (750, 2250, 0.25)
# test/sql/plink_score.test:48-75 -- weights [1.0, 0.5, -0.5, 2.0], mean imputation
SCORE_W = np.array([1.0, 0.5, -0.5, 2.0])
SCORE_SUM = [-0.5, 1.0, 3.5, 6.0]
SCORE_AVG = [-0.0625, 0.125, 0.4375, 0.75]
SCORE_DOSAGE_SUM = [3.0, 3.0, 4.0, 5.0]
# test/sql/plink_score.test:203-221 -- no_mean_imputation: (ALLELE_CT, SCORE_SUM, NAMED_ALLELE_DOSAGE_SUM)
SCORE_NOIMPUTE = {1: (6, 1.5, 2.0), 3: (6, 5.0, 4.0)}
# test/sql/plink_score.test:303-320 -- center := true on rs2 alone (region 1:20000-20000), weight 1.0
SCORE_CENTER_RS2 = {2: -1.414213562373095, 3: 1.414213562373095}
# test/sql/read_pfile_sample_counts_sparse.test:13 -- rare_small totals hom_ref / het / hom_alt / missing
RARE_SMALL_SUMS = (99217, 1947, 224, 1012)
# test/sql/plink_freq_dosage.test:103-118 -- dosage_example
DOSAGE_ALT_FREQ = [0.4583333333333333, 0.4499969482421875, 0.5, 0.20833333333333334]
DOSAGE_OBS_CT = [6, 8, 8, 6]
DOSAGE_IMP_R2 = [1.2900887786091624, 0.7374153374407626, 1.0, 0.3857199844561125]
# test/sql/plink_sexchr.test:109,139 -- chrX x1: female counts 1/1/1, male ref 2 / alt 1
XCHR_X1 = dict(fhom_ref=1, fhet=1, fhom_alt=1, male_ref=2, male_alt=1, p=1.0, p_midp=0.714286)
# test/sql/plink_sexchr.test:45,54,63 -- ALT_FREQ / OBS_CT of x1, y1, mt1
SEXCHR_FREQ = {"x1": (0.444444, 9), "y1": (0.666667, 3), "mt1": (0.666667, 6)}

# test/sql/read_pgen_phased.test:45-63 -- read_pgen('test/data/phased_example.pgen', phased := true), NULL = -9,-9
PHASED_EXAMPLE_PAIRS = [
    [[0, 0], [0, 1], [1, 0], [-9, -9]],   # rs1: 0|0 0|1 1|0 ./.
    [[0, 1], [1, 0], [0, 0], [1, 1]],     # rs2: 0/1 1|0 0|0 1|1
    [[1, 1], [-9, -9], [1, 0], [0, 0]],   # rs3: 1|1 ./. 1|0 0|0
    [[0, 0], [0, 0], [0, 1], [1, 1]],     # rs4: 0|0 0|0 0|1 1|1
]

# test/sql/read_pfile_genotypes_stats.test:15-33 -- genotypes := 'stats', variant orient, pgen_example:
# hom_ref het hom_alt missing n af maf missing_rate carrier_count het_rate (doubles ROUNDed to 6 places)
GENO_STATS_VARIANT = [
    (1, 1, 1, 1, 3, 0.5, 0.5, 0.25, 2, 0.333333),
    (1, 2, 1, 0, 4, 0.5, 0.5, 0.0, 3, 0.5),
    (1, 1, 1, 1, 3, 0.5, 0.5, 0.25, 2, 0.333333),
    (2, 1, 1, 0, 4, 0.375, 0.375, 0.0, 2, 0.25),
]
# test/sql/read_pfile_genotypes_stats.test:114-125 -- sample orient: n, af (4 places), missing_rate, carrier_count
GENO_STATS_SAMPLE = [(4, 0.375, 0.0, 2), (3, 0.3333, 0.25, 2), (4, 0.5, 0.0, 3), (3, 0.6667, 0.25, 2)]
# test/sql/read_pfile_genotypes_stats.test:129-141 -- samples := ['SAMPLE1','SAMPLE3']: hom_ref, het, af, missing_rate
GENO_STATS_SUBSET_1_3 = [(1, 0, 0.5, 0.0), (1, 1, 0.25, 0.0), (0, 1, 0.75, 0.0), (1, 1, 0.25, 0.0)]
# test/sql/read_pfile_genotypes_counts.test:66-78 -- sample orient counts: hom_ref het hom_alt missing per sample
GENO_COUNTS_SAMPLE = [(2, 1, 1, 0), (1, 2, 0, 1), (1, 2, 1, 0), (1, 0, 2, 1)]
# test/sql/read_pfile_genotypes_counts.test:33-47 -- variant orient counts under samples := ['SAMPLE1','SAMPLE3']
GENO_COUNTS_SUBSET_1_3 = [(1, 0, 1, 0), (1, 1, 0, 0), (0, 1, 1, 0), (1, 1, 0, 0)]
# test/sql/read_pfile_sample_counts_sparse.test:37-41 -- rare_small, orient := 'sample', genotypes := 'stats':
# sum(n), sum(carrier_count), round(sum(het) / sum(n), 6)
RARE_SMALL_STATS = (101388, 2171, 0.019203)
