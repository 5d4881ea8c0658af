/* TEST INFRASTRUCTURE ONLY -- see pgen_oracle.h.  Plain C11, scalar, one variant at a time: written to
 * be read, not to be fast.  Reference = teaguesterling/plinking_duck @ b32b0cd; "pgenlib" below is
 * third_party/plink-ng/2.0/include, "spec" is third_party/plink-ng/pgen_spec/pgen_spec.tex. */
#include "pgen_oracle.h"

#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

struct po_file {
  uint8_t* data;      /* whole .pgen file */
  uint64_t size;
  uint32_t V, N, mode;
  int fixed;
  uint64_t data_start;
  uint32_t fixed_width;
  uint32_t const_vrtype;
  uint8_t* vrtypes;   /* [V] (variable-width modes) */
  uint64_t* fpos;     /* [V+1] */
  uint32_t* ldbase;   /* [V] last preceding record that is not LD-compressed (GetLdbaseVidx, pgenlib_read.cc:1846-1894) */
  uint8_t* scratch;   /* [N] LD base codes */
};

static int fail(char* err, size_t errlen, int code, const char* msg) {
  if (err && errlen) snprintf(err, errlen, "%s", msg);
  return code;
}

static uint8_t* slurp(const char* path, uint64_t* size) {
  FILE* fp = fopen(path, "rb");
  if (!fp) return NULL;
  fseek(fp, 0, SEEK_END);
  long n = ftell(fp);
  fseek(fp, 0, SEEK_SET);
  uint8_t* buf = (uint8_t*)malloc((size_t)n + 16);
  if (buf && fread(buf, 1, (size_t)n, fp) != (size_t)n) {
    free(buf);
    buf = NULL;
  }
  fclose(fp);
  if (buf) {
    memset(buf + n, 0, 16);
    *size = (uint64_t)n;
  }
  return buf;
}

static uint32_t le32(const uint8_t* p) { return p[0] | ((uint32_t)p[1] << 8) | ((uint32_t)p[2] << 16) | ((uint32_t)p[3] << 24); }
static uint64_t le64(const uint8_t* p) { return le32(p) | ((uint64_t)le32(p + 4) << 32); }
static uint32_t le_n(const uint8_t* p, uint32_t nbytes) {
  uint32_t v = 0;
  for (uint32_t i = 0; i < nbytes; ++i) v |= (uint32_t)p[i] << (8 * i);
  return v;
}

/* Header: PgfiInitPhase1 (pgenlib_read.cc:689-961) + PgfiInitPhase2Ex (:1099-1680); spec :122-320.
 * Magic 6c 1b <mode>; V, N as little-endian u32; twelfth byte header_ctrl.
 *  - modes 0x02/0x03/0x04: fixed width ceil(N/4) [+ 2N | + 4N], records start at byte 12
 *    (+ ceil(V/8) when nonref flags are stored, header_ctrl >> 6 == 3);
 *  - modes 0x10/0x11 (and 0x20/0x21 with the same block in the .pgi): per variant block of 65536 an
 *    8-byte file offset, then per block the packed vrtypes (4 or 8 bits, header_ctrl bit 2) and the
 *    record lengths (1 + (header_ctrl & 3) bytes), [allele counts], [nonref flags]. */
int po_open(const char* pgen_path, const char* pgi_path, po_file** out, char* err, size_t errlen) {
  *out = NULL;
  po_file* f = (po_file*)calloc(1, sizeof(po_file));
  if (!f) return fail(err, errlen, 6, "out of memory");
  f->data = slurp(pgen_path, &f->size);
  if (!f->data) {
    free(f);
    return fail(err, errlen, 1, "cannot read .pgen");
  }
  uint8_t* hdr = f->data;
  uint64_t hdr_size = f->size;
  uint8_t* pgi = NULL;
  int rc = 0;
  if (f->size < 3 || hdr[0] != 0x6c || hdr[1] != 0x1b) {
    rc = fail(err, errlen, 3, "not a .pgen file");
    goto bad;
  }
  f->mode = hdr[2];
  if ((f->mode & 0xfe) != 0x20 && f->size < 12) {
    rc = fail(err, errlen, 3, "truncated .pgen header");
    goto bad;
  }
  if ((f->mode & 0xfe) == 0x20) {
    if (!pgi_path || !(pgi = slurp(pgi_path, &hdr_size))) {
      rc = fail(err, errlen, 1, "mode 0x20/0x21 needs the .pgen.pgi file");
      goto bad;
    }
    hdr = pgi;
    if (hdr_size < 12 || hdr[0] != 0x6c || hdr[1] != 0x1b || (hdr[2] & 0xfe) != 0x30) {
      rc = fail(err, errlen, 3, "not a .pgen.pgi file");
      goto bad;
    }
  }
  f->V = le32(hdr + 3);
  f->N = le32(hdr + 7);
  const uint32_t ctrl = hdr[11];
  if (!f->V || !f->N) {
    rc = fail(err, errlen, 3, "empty .pgen");
    goto bad;
  }
  const uint64_t geno_bytes = ((uint64_t)f->N + 3) / 4;
  f->scratch = (uint8_t*)malloc(f->N);
  if (f->mode >= 2 && f->mode <= 4) {
    f->fixed = 1;
    f->fixed_width = (uint32_t)(geno_bytes + (f->mode == 3 ? 2ull * f->N : f->mode == 4 ? 4ull * f->N : 0));
    f->const_vrtype = f->mode == 3 ? 0x40 : f->mode == 4 ? 0xc0 : 0;
    f->data_start = 12 + ((ctrl >> 6) == 3 ? ((uint64_t)f->V + 7) / 8 : 0);
    if (f->data_start + (uint64_t)f->V * f->fixed_width != f->size) {
      rc = fail(err, errlen, 3, "unexpected .pgen file size");
      goto bad;
    }
  } else if ((f->mode & 0xfe) == 0x10 || (f->mode & 0xfe) == 0x20) {
    if ((ctrl & 15) >= 8 || ((ctrl >> 4) & 3) > 1) {
      rc = fail(err, errlen, 4, "header_ctrl variant not supported by the oracle");
      goto bad;
    }
    const int wide = (ctrl & 4) != 0;
    const uint32_t len_bytes = 1 + (ctrl & 3);
    const uint32_t ac_bytes = (ctrl >> 4) & 3;
    const uint32_t nblocks = (f->V + 65535) / 65536;
    f->vrtypes = (uint8_t*)malloc(f->V);
    f->fpos = (uint64_t*)malloc(((size_t)f->V + 1) * 8);
    f->ldbase = (uint32_t*)malloc((size_t)f->V * 4);
    uint64_t pos = 12;
    uint64_t fpos = le64(hdr + pos);
    pos += 8ull * nblocks;
    for (uint32_t b = 0; b < nblocks; ++b) {
      const uint32_t v0 = b * 65536u;
      const uint32_t n = f->V - v0 < 65536u ? f->V - v0 : 65536u;
      const uint64_t type_bytes = wide ? n : (n + 1) / 2;
      if (pos + type_bytes + (uint64_t)n * len_bytes > hdr_size) {
        rc = fail(err, errlen, 3, "truncated .pgen header");
        goto bad;
      }
      const uint8_t* tp = hdr + pos;
      const uint8_t* lp = tp + type_bytes;
      for (uint32_t i = 0; i < n; ++i) {
        f->vrtypes[v0 + i] = wide ? tp[i] : (uint8_t)((tp[i >> 1] >> (4 * (i & 1))) & 15);
        f->fpos[v0 + i] = fpos;
        fpos += le_n(lp + (size_t)i * len_bytes, len_bytes);
      }
      pos += type_bytes + (uint64_t)n * len_bytes + (uint64_t)n * ac_bytes + ((ctrl >> 6) == 3 ? (n + 7) / 8 : 0);
    }
    f->fpos[f->V] = fpos;
    if (fpos > f->size) {
      rc = fail(err, errlen, 3, "records run past the end of the file");
      goto bad;
    }
    uint32_t last = 0;
    for (uint32_t v = 0; v < f->V; ++v) {
      const uint32_t t = f->vrtypes[v] & 7;
      if (t == 2 || t == 3) {
        f->ldbase[v] = last;
      } else {
        last = v;
        f->ldbase[v] = v;
      }
    }
  } else {
    rc = fail(err, errlen, 4, "storage mode not supported by the oracle");
    goto bad;
  }
  free(pgi);
  *out = f;
  return 0;
bad:
  free(pgi);
  po_close(f);
  return rc;
}

void po_close(po_file* f) {
  if (!f) return;
  free(f->data);
  free(f->vrtypes);
  free(f->fpos);
  free(f->ldbase);
  free(f->scratch);
  free(f);
}

void po_info(const po_file* f, uint32_t* variant_ct, uint32_t* sample_ct, uint32_t* mode) {
  if (variant_ct) *variant_ct = f->V;
  if (sample_ct) *sample_ct = f->N;
  if (mode) *mode = f->mode;
}
uint32_t po_vrtype(const po_file* f, uint32_t v) { return f->fixed ? f->const_vrtype : f->vrtypes[v]; }
uint64_t po_fpos(const po_file* f, uint32_t v) {
  return f->fixed ? f->data_start + (uint64_t)v * f->fixed_width : f->fpos[v];
}

/* ---- varints and difflists ------------------------------------------------------------------- */

/* GetVint31 (plink2_base.h:4259-4280): base-128 little-endian, < 2^31, at most 5 bytes. */
static int get_vint31(const uint8_t** pp, const uint8_t* end, uint32_t* out) {
  uint32_t val = 0;
  for (int shift = 0; shift < 35; shift += 7) {
    if (*pp >= end) return 0;
    const uint32_t b = *(*pp)++;
    if (shift == 28 && (b & 0xf8)) return 0;
    val |= (b & 127) << shift;
    if (!(b & 128)) {
      *out = val;
      return 1;
    }
  }
  return 0;
}

/* BytesToRepresentNzU32 (plink2_base.h:4098). */
static uint32_t bytes_to_represent(uint32_t n) {
  uint32_t b = 1;
  while (n >>= 8) ++b;
  return b;
}

/* Difflist (spec :354-390; ParseDifflistHeader pgenlib_read.cc:2184-2218, ParseAndSaveDifflist
 * :2220-2274): varint L; ceil(L/64) group-start sample ids (bytes_to_represent(N) bytes each);
 * ceil(L/64) - 1 group byte lengths (stored - 63); [ceil(L/4) bytes of 2-bit values]; then per
 * group the varint deltas between consecutive ids.  L <= N / 8.
 * Calls cb(ctx, sample_id, value, ordinal) per entry; *pp is left after the difflist. */
typedef void (*entry_cb)(void* ctx, uint32_t id, uint32_t geno, uint32_t ordinal);
static int walk_difflist(const uint8_t** pp, const uint8_t* end, uint32_t N, int has_geno, entry_cb cb, void* ctx,
                         uint32_t* len_out) {
  uint32_t len;
  if (!get_vint31(pp, end, &len)) return 0;
  if (len_out) *len_out = len;
  if (!len) return 1;
  if (len > N / 8) return 0;
  const uint32_t idb = bytes_to_represent(N);
  const uint32_t ngroups = (len + 63) / 64;
  const uint8_t* first_ids = *pp;
  const uint64_t index_bytes = (uint64_t)ngroups * (idb + 1) - 1;
  if ((uint64_t)(end - *pp) < index_bytes) return 0;
  *pp += index_bytes;
  const uint8_t* raregeno = NULL;
  if (has_geno) {
    const uint32_t rg = (len + 3) / 4;
    if ((uint64_t)(end - *pp) < rg) return 0;
    raregeno = *pp;
    *pp += rg;
  }
  const uint8_t* p = *pp;  /* the groups' delta streams follow one another */
  for (uint32_t g = 0; g < ngroups; ++g) {
    uint32_t id = le_n(first_ids + (size_t)g * idb, idb);
    const uint32_t cnt = len - g * 64 < 64 ? len - g * 64 : 64;
    for (uint32_t k = 0; k < cnt; ++k) {
      if (id >= N) return 0;
      const uint32_t ord = g * 64 + k;
      const uint32_t geno = raregeno ? (raregeno[ord >> 2] >> (2 * (ord & 3))) & 3 : 0;
      cb(ctx, id, geno, ord);
      if (k + 1 < cnt) {
        uint32_t delta;
        if (!get_vint31(&p, end, &delta)) return 0;
        id += delta;
      }
    }
  }
  *pp = p;
  return 1;
}

static void set_code(void* ctx, uint32_t id, uint32_t geno, uint32_t ord) {
  (void)ord;
  ((uint8_t*)ctx)[id] = (uint8_t)geno;
}

/* ---- record decode ----------------------------------------------------------------------------
 * ReadGenovecSubsetUnsafe (pgenlib_read.cc:2847-2910) without the subsetting; main track only
 * (vrtype & 7), as PgrGet returns it.  Genotype codes (spec :428-436): 0 hom-ref, 1 het, 2 hom-alt,
 * 3 missing, two bits per sample, sample 0 in the low bits.
 *   0  raw 2-bit array                         (ParseNonLdGenovecSubsetUnsafe :2711-2741)
 *   1  1-bit: byte common2 (low = >> 2, high = low + (& 3)), N-bit array, difflist of the rest
 *                                              (ParseOnebitUnsafe :2607-2690)
 *   2  difflist against the LD base record; 3 same, then categories 0 and 2 swapped
 *                                              (LdLoadAndCopyGenovecSubset :2808-2838, GenovecInvertUnsafe)
 *   4, 6, 7  difflist against all-0 / all-2 / all-3;  5  all hom-ref, empty track (:2738-2740) */
static int decode_nonld(const po_file* f, uint32_t v, uint8_t* codes) {
  const uint32_t N = f->N;
  const uint8_t* p = f->data + po_fpos(f, v);
  const uint8_t* end = f->data + po_fpos(f, v + 1);
  const uint32_t t = po_vrtype(f, v) & 7;
  if (t == 0) {
    if ((uint64_t)(end - p) < (N + 3) / 4) return 0;
    for (uint32_t s = 0; s < N; ++s) codes[s] = (p[s >> 2] >> (2 * (s & 3))) & 3;
    return 1;
  }
  if (t == 5) {
    memset(codes, 0, N);
    return 1;
  }
  if (t == 1) {
    const uint32_t bit_bytes = (N + 7) / 8;
    if ((uint64_t)(end - p) < (uint64_t)bit_bytes + 1) return 0;
    const uint32_t low = p[0] >> 2, high = low + (p[0] & 3);
    if (high > 3) return 0;
    const uint8_t* bits = p + 1;
    for (uint32_t s = 0; s < N; ++s) codes[s] = (uint8_t)(((bits[s >> 3] >> (s & 7)) & 1) ? high : low);
    const uint8_t* q = bits + bit_bytes;
    return walk_difflist(&q, end, N, 1, set_code, codes, NULL);
  }
  if (t == 4 || t == 6 || t == 7) {
    memset(codes, (int)(t & 3), N);
    const uint8_t* q = p;
    return walk_difflist(&q, end, N, 1, set_code, codes, NULL);
  }
  return 0;
}

int po_read_codes(po_file* f, uint32_t v, uint8_t* codes) {
  if (v >= f->V) return 2;
  const uint32_t t = po_vrtype(f, v) & 7;
  if (t != 2 && t != 3) return decode_nonld(f, v, codes) ? 0 : 3;
  if (f->fixed) return 3;
  if (!decode_nonld(f, f->ldbase[v], codes)) return 3;
  const uint8_t* q = f->data + f->fpos[v];
  if (!walk_difflist(&q, f->data + f->fpos[v + 1], f->N, 1, set_code, codes, NULL)) return 3;
  if (t == 3) {
    for (uint32_t s = 0; s < f->N; ++s) {
      if (codes[s] == 0) codes[s] = 2;
      else if (codes[s] == 2) codes[s] = 0;
    }
  }
  return 0;
}

/* ---- reductions -------------------------------------------------------------------------------- */

/* PgrGetCounts -> GetBasicGenotypeCounts (pgenlib_read.cc:3624-3739): the reference counts without
 * decoding where it can; the result is by definition the histogram of the decoded codes over the
 * included samples, which is what is computed here. */
int po_counts(po_file* f, uint32_t v0, uint32_t v1, const uint8_t* include, uint32_t* counts) {
  if (v0 > v1 || v1 > f->V) return 2;
  uint8_t* codes = (uint8_t*)malloc(f->N);
  for (uint32_t v = v0; v < v1; ++v) {
    const int rc = po_read_codes(f, v, codes);
    if (rc) {
      free(codes);
      return rc;
    }
    uint32_t* c = counts + (size_t)(v - v0) * 4;
    c[0] = c[1] = c[2] = c[3] = 0;
    for (uint32_t s = 0; s < f->N; ++s) {
      if (!include || ((include[s >> 3] >> (s & 7)) & 1)) c[codes[s]]++;
    }
  }
  free(codes);
  return 0;
}

/* PgrGet + GenoarrToBytesMinus9 (pgenlib_ffi_support.cc:25-43). */
int po_decode_i8(po_file* f, uint32_t v0, uint32_t v1, int8_t* out) {
  if (v0 > v1 || v1 > f->V) return 2;
  uint8_t* codes = (uint8_t*)malloc(f->N);
  for (uint32_t v = v0; v < v1; ++v) {
    const int rc = po_read_codes(f, v, codes);
    if (rc) {
      free(codes);
      return rc;
    }
    int8_t* o = out + (size_t)(v - v0) * f->N;
    for (uint32_t s = 0; s < f->N; ++s) o[s] = codes[s] == 3 ? (int8_t)-9 : (int8_t)codes[s];
  }
  free(codes);
  return 0;
}

static void ignore_entry(void* ctx, uint32_t id, uint32_t geno, uint32_t ord) {
  (void)ctx; (void)id; (void)geno; (void)ord;
}

/* End of the main genotype track of record v = start of aux2 when the record has no multiallelic
 * patch sets (SkipAux1 / main-track layouts, pgenlib_read.cc:6844-6865 and the decoders above). */
static const uint8_t* main_track_end(const po_file* f, uint32_t v) {
  const uint32_t N = f->N;
  const uint8_t* p = f->data + po_fpos(f, v);
  const uint8_t* end = f->data + po_fpos(f, v + 1);
  const uint32_t t = po_vrtype(f, v) & 7;
  const uint8_t* q = p;
  if (t == 0) return p + (N + 3) / 4 <= end ? p + (N + 3) / 4 : NULL;
  if (t == 5) return p;
  if (t == 1) {
    q = p + 1 + (N + 7) / 8;
    if (q > end) return NULL;
  }
  return walk_difflist(&q, end, N, 1, ignore_entry, NULL, NULL) ? q : NULL;
}

/* PgrGetP (ReadGenovecHphaseSubsetUnsafe, pgenlib_read.cc:6874-6962) + UnpackPhasedGenotypes
 * (src/plink_common.cpp:1549-1584): out[(v - v0) * N + s][2] allele pairs.  hom-ref 0,0; hom-alt 1,1;
 * missing -9,-9; het 1,0 iff the aux2 track marks it phased with ALT first, else 0,1.
 * aux2 (vrtype bit 4): het_ct + 1 bits; bit 0 clear -> all hets phased, bits 1.. = phaseinfo;
 * bit 0 set -> bits 1.. = phasepresent flags, and one phaseinfo bit per set flag follows at the next
 * byte boundary (1 + het_ct / 8 bytes in). */
int po_decode_phased(po_file* f, uint32_t v0, uint32_t v1, int8_t* out) {
  if (v0 > v1 || v1 > f->V) return 2;
  uint8_t* codes = (uint8_t*)malloc(f->N);
  for (uint32_t v = v0; v < v1; ++v) {
    int rc = po_read_codes(f, v, codes);
    const uint32_t vt = po_vrtype(f, v);
    const uint8_t* aux2 = NULL;
    if (!rc && (vt & 0x10)) {
      if (vt & 0x08) rc = 4; /* multiallelic + phase: not restated */
      else if (!(aux2 = main_track_end(f, v))) rc = 3;
    }
    if (rc) {
      free(codes);
      return rc;
    }
    uint32_t het_ct = 0;
    for (uint32_t s = 0; s < f->N; ++s) het_ct += codes[s] == 1;
    const uint8_t* end = f->data + po_fpos(f, v + 1);
    const int explicit_pp = aux2 && (aux2[0] & 1);
    const uint8_t* second = aux2 ? aux2 + 1 + het_ct / 8 : NULL;
    if (aux2 && second > end) {
      free(codes);
      return 3;
    }
    int8_t* o = out + (size_t)(v - v0) * f->N * 2;
    uint32_t h = 0, k = 0; /* het ordinal, flagged-het ordinal */
    for (uint32_t s = 0; s < f->N; ++s) {
      int8_t a = -9, b = -9;
      if (codes[s] == 0) { a = 0; b = 0; }
      else if (codes[s] == 2) { a = 1; b = 1; }
      else if (codes[s] == 1) {
        int alt_first = 0;
        if (aux2) {
          const uint32_t bit = 1 + h;
          const int flag = (aux2[bit >> 3] >> (bit & 7)) & 1;
          if (!explicit_pp) alt_first = flag;
          else if (flag) {
            alt_first = (second[k >> 3] >> (k & 7)) & 1;
            ++k;
          }
        }
        ++h;
        a = alt_first ? 1 : 0;
        b = alt_first ? 0 : 1;
      }
      o[2 * s] = a;
      o[2 * s + 1] = b;
    }
  }
  free(codes);
  return 0;
}

/* PlinkMissingScanSample phase 1 (src/plink_missing.cpp:585-611): PgrGetMissingness, then one
 * increment per set bit. */
int po_missing_by_sample(po_file* f, uint32_t v0, uint32_t v1, uint32_t* out) {
  if (v0 > v1 || v1 > f->V) return 2;
  uint8_t* codes = (uint8_t*)malloc(f->N);
  memset(out, 0, (size_t)f->N * 4);
  for (uint32_t v = v0; v < v1; ++v) {
    const int rc = po_read_codes(f, v, codes);
    if (rc) {
      free(codes);
      return rc;
    }
    for (uint32_t s = 0; s < f->N; ++s) out[s] += codes[s] == 3;
  }
  free(codes);
  return 0;
}

/* ---- exact Hardy-Weinberg test -----------------------------------------------------------------
 * What HweExactTestAutosomal computes (src/plink_hardy.cpp:52-79 -> plink2::HweLnP,
 * plink2_stats.cc:1784-2166, then p = clamp(exp(lnp), 0, 1), n == 0 -> 1): the exact test of
 * Wigginton, Cutler & Abecasis (2005) -- given the allele counts, sum the probabilities of all
 * heterozygote counts no more likely than the observed one; mid-p (Graffelman & Moreno 2013) counts
 * the observed probability (and exact ties) half.  HweLnP evaluates the same sums through log
 * factorials to survive huge counts; this restatement walks the two-term recurrences
 *   P(k-2)/P(k) = k (k-1) / (4 (r+1)(c+1)),   P(k+2)/P(k) = 4 r c / ((k+2)(k+1))
 * in units of the observed table's probability, exact to a few ulp per thousand steps.  Ties are
 * recognised with the relative window 2^-38 (HweLnP's kSmallEpsilon-style guard). */
double po_hwe_exact(uint32_t obs_hets, uint32_t obs_hom1, uint32_t obs_hom2, int midp) {
  if (!(obs_hets + obs_hom1 + obs_hom2)) return 1.0;
  const double r0 = obs_hom1 < obs_hom2 ? obs_hom1 : obs_hom2;
  const double c0 = obs_hom1 < obs_hom2 ? obs_hom2 : obs_hom1;
  if (2.0 * r0 + obs_hets < 2.0) return midp ? 0.5 : 1.0;
  const double eps = 3.637978807091713e-12;
  double tail = 1.0, center = 0.0, ties = 1.0;
  double k = obs_hets, r = r0, c = c0, t = 1.0;
  while (k > 1.5) { /* fewer heterozygotes */
    r += 1.0;
    c += 1.0;
    t *= (k * (k - 1.0)) / (4.0 * r * c);
    k -= 2.0;
    if (t > 1.0 + eps) {
      center += t;
      if (!(center < 1.7e308)) return 0.0;
    } else {
      if (t >= 1.0 - eps) ties += 1.0;
      const double before = tail;
      tail += t;
      if (tail <= before) break;
    }
  }
  k = obs_hets;
  r = r0;
  c = c0;
  t = 1.0;
  while (r > 0.5) { /* more heterozygotes */
    k += 2.0;
    t *= (4.0 * r * c) / (k * (k - 1.0));
    r -= 1.0;
    c -= 1.0;
    if (t > 1.0 + eps) {
      center += t;
      if (!(center < 1.7e308)) return 0.0;
    } else {
      if (t >= 1.0 - eps) ties += 1.0;
      const double before = tail;
      tail += t;
      if (tail <= before) break;
    }
  }
  const double total = tail + center;
  if (midp) tail -= 0.5 * ties;
  double p = tail / total;
  if (p < 0.0) p = 0.0;
  if (p > 1.0) p = 1.0;
  return p;
}

/* ---- exact Hardy-Weinberg test, ln p, extended precision -----------------------------------------
 * Same definition as po_hwe_exact (HweLnP's result before LnPToPvalue, plink2_stats.cc:1784-2166;
 * src/plink_hardy.cpp:52-79), but (1) in x87 long double (64-bit mantissa: the ~1e5 recurrence steps of
 * a biobank-sized table cost < 1e-14 relative) and (2) with every quantity carried as mantissa x 2^e with a
 * 64-bit exponent, so the answer is ln p all the way down (HweLnP returns e.g. -40000 where p underflows).
 * This is the yardstick the parity tests use to state what the reference's own log-factorial noise is and
 * what the device kernel's error is.  Tables with probability within 2^-38 of the observed one count as ties. */
typedef struct {
  long double m; /* 0 or in [2^-4000, 2^4000) after xf_norm */
  long long e;
} xf_t;

static xf_t xf_norm(xf_t a) {
  if (a.m == 0.0L) {
    a.e = 0;
    return a;
  }
  while (a.m >= 0x1p4000L) {
    a.m *= 0x1p-4000L;
    a.e += 4000;
  }
  while (a.m < 0x1p-4000L) {
    a.m *= 0x1p4000L;
    a.e -= 4000;
  }
  return a;
}

static xf_t xf_add(xf_t a, xf_t b) {
  if (a.m == 0.0L) return b;
  if (b.m == 0.0L) return a;
  if (a.e < b.e) {
    const xf_t t = a;
    a = b;
    b = t;
  }
  const long long d = a.e - b.e; /* multiple of 4000 */
  if (d >= 12000) return a;
  long double bm = b.m;
  for (long long i = 0; i < d; i += 4000) bm *= 0x1p-4000L;
  a.m += bm;
  return xf_norm(a);
}

/* a <= b * (1 + rel) ? */
static int xf_le1(xf_t a, long double bound) {
  if (a.m == 0.0L) return 1;
  if (a.e != 0) return a.e < 0;
  return a.m <= bound;
}

double po_hwe_lnp_exact(uint32_t obs_hets, uint32_t obs_hom1, uint32_t obs_hom2, int midp) {
  const long double r0 = obs_hom1 < obs_hom2 ? obs_hom1 : obs_hom2;
  const long double c0 = obs_hom1 < obs_hom2 ? obs_hom2 : obs_hom1;
  if (2.0L * r0 + obs_hets < 2.0L) return midp ? -0.6931471805599453094 : 0.0;
  const long double eps = 3.637978807091713e-12L;
  xf_t tail = {1.0L, 0}, center = {0.0L, 0};
  long double ties = 1.0L;
  for (int dir = 0; dir < 2; ++dir) {
    long double k = obs_hets, r = r0, c = c0;
    xf_t t = {1.0L, 0};
    for (;;) {
      if (dir == 0) { /* fewer heterozygotes */
        if (k < 1.5L) break;
        r += 1.0L;
        c += 1.0L;
        t.m *= (k * (k - 1.0L)) / (4.0L * r * c);
        k -= 2.0L;
      } else { /* more heterozygotes */
        if (r < 0.5L) break;
        k += 2.0L;
        t.m *= (4.0L * r * c) / (k * (k - 1.0L));
        r -= 1.0L;
        c -= 1.0L;
      }
      t = xf_norm(t);
      if (!xf_le1(t, 1.0L + eps)) {
        center = xf_add(center, t);
      } else {
        if (t.e == 0 && t.m >= 1.0L - eps) ties += 1.0L;
        const xf_t before = tail;
        tail = xf_add(tail, t);
        if (tail.e == before.e && tail.m == before.m) break; /* unimodal: every later term is smaller */
      }
    }
  }
  const xf_t total = xf_add(tail, center);
  xf_t num = tail;
  if (midp) { /* tail.e == 0: it is a sum of at most 2^31 terms <= 1 + eps */
    num.m -= 0.5L * ties;
    if (num.m <= 0.0L) return -INFINITY;
    num = xf_norm(num);
  }
  const long double lnp = logl(num.m / total.m) + (long double)(num.e - total.e) * 0.693147180559945309417232121458L;
  return lnp > 0.0L ? 0.0 : (double)lnp;
}

/* ---- plink_score ---------------------------------------------------------------------------------
 * The Phase-1 loop of src/plink_score.cpp:582-653 for one weight vector, hardcall records
 * (PgrGetD + Dosage16ToDoublesMinus9 yield 0.0 / 1.0 / 2.0 / -9.0 for them), same operation order,
 * single thread: pass 1 sum_alt / non_missing_ct (:598-606), skip when nothing is observed (:608),
 * then one of the three modes (:612-652). */
int po_score(po_file* f, const uint32_t* vidx, const uint8_t* flip, const double* weights, uint32_t n, int mode,
             double* score, double* dosage_sum, uint32_t* allele_ct) {
  uint8_t* codes = (uint8_t*)malloc(f->N);
  for (uint32_t i = 0; i < n; ++i) {
    if (po_vrtype(f, vidx[i]) & 0x60) {
      free(codes);
      return 4; /* explicit dosages: use the reference oracle */
    }
    const int rc = po_read_codes(f, vidx[i], codes);
    if (rc) {
      free(codes);
      return rc;
    }
    const int fl = flip ? flip[i] : 0;
    const double w = weights[i];
    double sum_alt = 0.0;
    uint32_t nm = 0;
    for (uint32_t s = 0; s < f->N; ++s) {
      if (codes[s] != 3) {
        sum_alt += (double)codes[s];
        ++nm;
      }
    }
    if (!nm) continue;
    const double mean_alt = sum_alt / (double)nm;
    if (mode == 2) {
      const double freq = mean_alt / 2.0;
      const double sd = sqrt(2.0 * freq * (1.0 - freq));
      if (sd == 0.0) continue;
      const double mean_scored = fl ? (2.0 - mean_alt) : mean_alt;
      for (uint32_t s = 0; s < f->N; ++s) {
        if (codes[s] == 3) continue;
        const double d = (double)codes[s];
        const double scored = fl ? (2.0 - d) : d;
        score[s] += w * ((scored - mean_scored) / sd);
        allele_ct[s] += 2;
      }
    } else if (mode == 1) {
      for (uint32_t s = 0; s < f->N; ++s) {
        if (codes[s] == 3) continue;
        const double d = (double)codes[s];
        const double scored = fl ? (2.0 - d) : d;
        score[s] += w * scored;
        dosage_sum[s] += scored;
        allele_ct[s] += 2;
      }
    } else {
      for (uint32_t s = 0; s < f->N; ++s) {
        const double d = codes[s] == 3 ? mean_alt : (double)codes[s];
        const double scored = fl ? (2.0 - d) : d;
        score[s] += w * scored;
        dosage_sum[s] += scored;
        allele_ct[s] += 2;
      }
    }
  }
  free(codes);
  return 0;
}
