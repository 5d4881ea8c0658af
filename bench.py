#!/usr/bin/env python
"""bench.py -- genotype calls decoded+counted per second on B200 vs the reference's CPU scan.

Workload (BASELINE.json configs[1]): plink_hardy over a synthetic fixed-width (mode 0x02) .pgen of
1M variants x 100k samples PER GPU (25 GB of packed 2-bit records each) -- ONE file on tmpfs holding all
N ranges, opened once through the C ABI as ONE context over N GPUs (pd_open_opts.n_devices; the library
partitions the variant range, runs one sub-context + staging ring per GPU and owns the NCCL communicator).
One "step" = one full plink_hardy pass: genotype counts of every variant + exact HWE p-values (K1h) +
results copied to host memory.

  value   : calls/s with the records resident in HBM when the timed region starts (CUDA-event time of the
            call on the library's streams, the slowest GPU's = max over ranks).
  e2e     : the same call from the .pgen FILE on tmpfs (the file the reference arm reads), opened with
            PD_OPEN_PIN_FILE (mapping page-locked once at open): every step copies every record host->device
            (double-buffered cudaMemcpyAsync straight from the page cache) and the results back; wall clock
            around the blocking calls.  e2e.unpinned_file is the default open (pread into the page-locked
            staging ring first: three host memory passes per byte), e2e.pinned_image the same call fed from a
            page-locked in-memory image, e2e.h2d_ceiling_gbs the box's raw concurrent pinned H2D bandwidth.
  roofline: K1h (counts + exact test fused), algorithmic bytes = V * ceil(N/4) per launch and GPU / its
            CUDA-event duration, against the measured HBM copy bandwidth in MEASURED_PEAKS.json.
  score   : (N > 1) one plink_score pass, 16 weight columns over every variant, with the ncclAllReduce of
            the per-sample partial sums inside the timed region.
  cpu_baseline / --impl reference: the reference's own pgenlib (oracle/_ref, PgrGetCounts + HweLnP, one
            reader per thread, fetch_add(128) batches) over the whole 1M x 100k file on tmpfs.

Under torchrun (N > 1) rank 0 drives all N GPUs through the one context, as a DuckDB process would; the other
ranks only join the rendezvous.  Inputs are larger than L2 (25 GB vs 126 MB), so no explicit L2 flush is
needed between steps.
"""
import argparse
import contextlib
import io
import concurrent.futures as cf
import datetime
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

METRIC = "genotype calls decoded+counted/sec"
UNIT = "calls/s"
SEED = 20261104
DTYPE = "u32 popcount (counts) + f64 (HWE)"


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--variants", type=int, default=0, help="override variants per GPU (dev only)")
    ap.add_argument("--samples", type=int, default=100_000)
    ap.add_argument("--skip-cpu", action="store_true")
    ap.add_argument("--skip-score", action="store_true")
    ap.add_argument("--skip-pinned", action="store_true")
    return ap.parse_args()


class ClockSampler:
    """nvidia-smi clocks / throttle reasons DURING the timed region (B200_PROFILING.md recipe)."""

    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.index = index
        self.rows = []
        self.proc = None
        self.thread = None

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "100", "-i",
                 str(self.index)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
        except Exception:
            self.proc = None
            return
        self.thread = threading.Thread(target=self._read, daemon=True)
        self.thread.start()

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append(line.strip())

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, smax, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            f = [x.strip() for x in r.split(",")]
            if len(f) < 7:
                continue
            try:
                sm.append(float(f[0]))
                smax.append(float(f[1]))
            except ValueError:
                continue
            for nm, val in zip(names, f[3:7]):
                if val.lower().startswith("active"):
                    reasons.add(nm)
        return {
            "sm_mhz": float(np.median(sm)) if sm else None,
            "sm_max_mhz": float(max(smax)) if smax else None,
            "samples": len(sm),
            "reasons": sorted(reasons),
        }


def generate_records_gpu(V: int, N: int, seed: int, device, out_u8):
    """V mode-0x02 records (V x ceil(N/4) bytes, no header) into out_u8 (a uint8 tensor on `device`), generated
    on the GPU tile by tile: p_v ~ U(0.05, 0.5), g ~ Binomial(2, p_v), 1 % missing."""
    import torch

    W = (N + 3) // 4
    gen = torch.Generator(device=device)
    gen.manual_seed(seed)
    tile = max(1, min(V, (256 << 20) // max(N, 1)))
    for v0 in range(0, V, tile):
        vt = min(tile, V - v0)
        p = 0.05 + 0.45 * torch.rand((vt, 1), device=device, generator=gen)
        u = torch.rand((vt, W * 4), device=device, generator=gen)
        p2 = p * p
        code = (u < p2).to(torch.uint8) * 2 + ((u >= p2) & (u < p2 + 2 * p * (1 - p))).to(torch.uint8)
        miss = torch.rand((vt, W * 4), device=device, generator=gen) < 0.01
        code = torch.where(miss, torch.full_like(code, 3), code)
        if W * 4 != N:
            code[:, N:] = 0
        c = code.view(vt, W, 4)
        b = c[:, :, 0] | (c[:, :, 1] << 2) | (c[:, :, 2] << 4) | (c[:, :, 3] << 6)
        out_u8[v0 * W: (v0 + vt) * W].copy_(b.reshape(-1), non_blocking=True)
        del p, u, code, miss, c, b
    torch.cuda.synchronize(device)


def generate_image_gpu(V: int, N: int, seed: int, device, out_u8):
    """Whole mode-0x02 file image (12-byte header + records) into out_u8 on `device` (development scripts)."""
    import synth
    import torch

    out_u8[:12].copy_(torch.from_numpy(np.frombuffer(synth.fixed_header(V, N), dtype=np.uint8).copy()))
    generate_records_gpu(V, N, seed, device, out_u8[12:])


def write_fileset(path: str, V_per_gpu: int, N: int, n_gpus: int, seed: int = SEED):
    """One mode-0x02 .pgen of n_gpus x V_per_gpu variants on tmpfs: range r is generated on GPU r (seed + r),
    brought back through a page-locked ring and written with parallel pwrite."""
    import synth
    import torch

    W = (N + 3) // 4
    V_total = V_per_gpu * n_gpus
    fd = os.open(path, os.O_CREAT | os.O_WRONLY | os.O_TRUNC, 0o600)
    try:
        os.pwrite(fd, synth.fixed_header(V_total, N), 0)
        chunk = 1 << 30
        ring = [torch.empty(chunk, dtype=torch.uint8, pin_memory=True) for _ in range(2)]
        piece = 64 << 20
        with cf.ThreadPoolExecutor(max_workers=min(16, os.cpu_count() or 4)) as pool:
            pending = [None, None]
            k = 0
            for r in range(n_gpus):
                dev = torch.device("cuda", r)
                body = torch.empty(V_per_gpu * W, dtype=torch.uint8, device=dev)
                generate_records_gpu(V_per_gpu, N, seed + r, dev, body)
                base = 12 + r * V_per_gpu * W
                for off in range(0, V_per_gpu * W, chunk):
                    n = min(chunk, V_per_gpu * W - off)
                    slot = k & 1
                    if pending[slot]:
                        for f in pending[slot]:
                            f.result()
                    ring[slot][:n].copy_(body[off:off + n])
                    torch.cuda.synchronize(dev)
                    mv = memoryview(ring[slot].numpy())
                    pending[slot] = [pool.submit(os.pwrite, fd, mv[a:min(n, a + piece)], base + off + a)
                                     for a in range(0, n, piece)]
                    k += 1
                del body
                torch.cuda.empty_cache()
            for p in pending:
                if p:
                    for f in p:
                        f.result()
        del ring
    finally:
        os.close(fd)
    return V_total


def read_file_pinned(path: str, nbytes: int):
    """The first nbytes of the file in a page-locked tensor (parallel pread)."""
    import torch

    t = torch.empty(nbytes, dtype=torch.uint8, pin_memory=True)
    mv = memoryview(t.numpy())
    fd = os.open(path, os.O_RDONLY)
    piece = 64 << 20
    try:
        with cf.ThreadPoolExecutor(max_workers=min(16, os.cpu_count() or 4)) as pool:
            list(pool.map(lambda a: os.preadv(fd, [mv[a:min(nbytes, a + piece)]], a), range(0, nbytes, piece)))
    finally:
        os.close(fd)
    return t


def h2d_ceiling(n_gpus: int) -> float:
    """Raw concurrent pinned H2D bandwidth of the box: 1 GiB blocks, all n_gpus devices at once (GB/s, aggregate)."""
    import torch

    gib = 1 << 30
    src = [torch.empty(gib, dtype=torch.uint8, pin_memory=True) for _ in range(n_gpus)]
    dst = [torch.empty(gib, dtype=torch.uint8, device=f"cuda:{i}") for i in range(n_gpus)]
    streams = [torch.cuda.Stream(device=i) for i in range(n_gpus)]
    best = 0.0
    for _ in range(3):
        for i in range(n_gpus):
            torch.cuda.synchronize(i)
        t0 = time.perf_counter()
        for _rep in range(4):
            for i in range(n_gpus):
                with torch.cuda.stream(streams[i]):
                    dst[i].copy_(src[i], non_blocking=True)
        for i in range(n_gpus):
            streams[i].synchronize()
        best = max(best, 4.0 * gib * n_gpus / (time.perf_counter() - t0) / 1e9)
    del src, dst
    torch.cuda.empty_cache()
    return best


def sample_dir():
    return "/dev/shm" if os.path.isdir("/dev/shm") and os.access("/dev/shm", os.W_OK) else "/tmp"


def mem_available_bytes() -> int:
    try:
        for line in open("/proc/meminfo"):
            if line.startswith("MemAvailable:"):
                return int(line.split()[1]) * 1024
    except OSError:
        pass
    return 64 << 30


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            return float(json.load(f)["hbm_gbs"]), "measured"
    return 6650.0, "fallback"


def measured_traffic(alg_bytes, fused):
    """DRAM bytes of one launch of the dominant kernel: the ratio dram bytes / algorithmic bytes of the
    committed `ncu --set full` capture (profiles/traffic.json), scaled to this launch's algorithmic bytes."""
    try:
        with open(os.path.join(ROOT, "profiles", "traffic.json")) as f:
            t = json.load(f)["k_hardy_fixed"]
    except (OSError, KeyError, ValueError):
        return None, None
    if not fused:
        return None, None
    return alg_bytes * t["dram_bytes_per_algorithmic_byte"], f"{t['capture']} ratio x this launch's algorithmic bytes"


def numpy_counts(records: np.ndarray, N: int) -> np.ndarray:
    """(rows, W) uint8 records -> (rows, 4) uint32 genotype counts."""
    lut = np.zeros((256, 4), dtype=np.uint32)
    for b in range(256):
        for s in (0, 2, 4, 6):
            lut[b, (b >> s) & 3] += 1
    c = lut[records].sum(axis=1, dtype=np.uint64)
    pad = records.shape[1] * 4 - N  # trailing pad codes are 0 (hom-ref)
    c[:, 0] -= pad
    return c.astype(np.uint32)


def parity_at_bench_size(path, V_total, N, out_counts, out_p):
    """>= 1 % stratified sample of the variants: counts against a numpy tally of the file's own bytes, P_HWE
    against plink2::HweLnP (oracle/_ref) within the reference's noise and against the extended-precision
    restatement within 1e-12 relative (tests/hwe_check.py).  Plus the size-independent properties over ALL rows."""
    import hwe_check as hc
    from oracle import port, ref

    W = (N + 3) // 4
    assert np.all(out_counts[:V_total].sum(axis=1, dtype=np.uint64) == N), "counts do not add up to N"
    p = out_p[:V_total]
    assert np.all(np.isfinite(p)) and float(p.min()) >= 0.0 and float(p.max()) <= 1.0
    # >= 1 % of the variants up to 2M (N <= 2); beyond that 20 000 variants stratified over the whole range (the
    # numpy tally of the sample runs on rank 0 while every GPU of the box is held)
    n_sample = max(64, min(V_total // 100, 20_000))
    rng = np.random.default_rng(SEED)
    strata = np.linspace(0, V_total, n_sample + 1).astype(np.int64)
    idx = np.unique(np.minimum(strata[:-1] + rng.integers(0, np.maximum(1, np.diff(strata))), V_total - 1))
    idx = np.unique(np.concatenate([idx, [0, V_total - 1]]))
    mm = np.memmap(path, dtype=np.uint8, mode="r", offset=12, shape=(V_total, W))
    exp = np.concatenate([numpy_counts(np.asarray(mm[idx[a:a + 512]]), N) for a in range(0, idx.size, 512)])
    assert np.array_equal(out_counts[idx], exp), "GPU counts disagree with the file's bytes"
    res = {"variants_compared": int(idx.size), "fraction": idx.size / V_total, "counts": "bit-exact vs numpy tally"}
    n_obs = exp[:, :3].sum(axis=1)
    if port.available():
        ex = np.array([port.hwe_lnp_exact(int(c[1]), int(c[0]), int(c[2]), False) for c in exp])
        v = hc.violations_p_vs_exact(p[idx], ex, n_obs)
        assert v.size == 0, "P_HWE vs extended precision: " + hc.describe(v, exp, p[idx], None, ex)
        with np.errstate(divide="ignore"):
            res["p_hwe_max_rel_err_vs_extended_precision"] = float(np.max(np.abs(np.log(p[idx]) - ex)))
    if ref.available():
        rf = ref.hwe_lnp_batch(exp, False)
        v = hc.violations_p_vs_reference(p[idx], rf, n_obs)
        assert v.size == 0, "P_HWE vs HweLnP: " + hc.describe(v, exp, p[idx], rf)
        res["p_hwe"] = f"within {hc.REF_ULPS:g} ulp(2n ln 2n) of plink2::HweLnP (its own noise) on every sampled variant"
    res["p_hwe_min_sampled"] = float(p[idx].min())
    return res


def run_ours(args):
    import torch

    from plinking_duck_b200.pgen import PgenContext

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    n_gpus = world if world > 1 else max(1, args.gpus)
    if world > 1:
        # one process drives every GPU through one context (what the DuckDB extension does); the other ranks
        # of the launcher only keep the rendezvous alive
        import torch.distributed as dist

        dist.init_process_group("gloo", timeout=datetime.timedelta(hours=3))
        if rank != 0:
            dist.barrier()
            dist.destroy_process_group()
            return
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device (the product path has no CPU fallback)")
    if torch.cuda.device_count() < n_gpus:
        raise SystemExit(f"bench.py: --gpus {n_gpus} but only {torch.cuda.device_count()} devices are visible")
    N = args.samples
    V = args.variants or 1_000_000          # per GPU at every N: weak scaling
    W = (N + 3) // 4
    V_total = V * n_gpus
    devices = list(range(n_gpus)) if n_gpus > 1 else None
    path = os.path.join(sample_dir(), f"plinkgpu_bench_{os.getpid()}.pgen")
    line = None
    try:
        t0 = time.perf_counter()
        write_fileset(path, V, N, n_gpus)
        gen_s = time.perf_counter() - t0
        file_bytes = 12 + V_total * W
        out_counts = torch.empty((V_total, 4), dtype=torch.int32, pin_memory=True).numpy().view(np.uint32)
        out_p = torch.empty(V_total, dtype=torch.float64, pin_memory=True).numpy()
        calls_per_step = float(V_total) * float(N)
        ceiling = h2d_ceiling(n_gpus)

        # ---- e2e: the .pgen FILE on tmpfs (the file the reference arm reads) -> H2D -> kernels -> results on host.
        # (1) default open: every step preads the records into the page-locked staging ring first;
        # (2) PD_OPEN_PIN_FILE: the file's mapping is page-locked once at open, every step DMAs straight from
        #     the page cache.  (2) is the headline e2e; (1) and the open / registration time are reported with it.
        e2e_steps = max(1, min(args.steps, 3))

        def time_file_fed(c, steps):
            c.hardy(0, V_total, out_counts=out_counts, out_p=out_p)  # warm-up (allocations, clocks)
            c.reset_stats()
            t0 = time.perf_counter()
            for _ in range(steps):
                c.hardy(0, V_total, out_counts=out_counts, out_p=out_p)
            return (time.perf_counter() - t0) / steps, c.stats()

        t0 = time.perf_counter()
        ctx = PgenContext(path, devices=devices)
        open_s = time.perf_counter() - t0
        un_s, un_st = time_file_fed(ctx, e2e_steps if n_gpus <= 2 else 1)  # the slow leg: one step at N > 2
        file_counts = out_counts[:V_total].copy()
        file_p = out_p[:V_total].copy()
        unpinned = {"value": calls_per_step / un_s, "unit": UNIT, "ms_per_step": un_s * 1e3, "gbs": V_total * W / un_s / 1e9,
                    "host_stage_ms": un_st.host_stage_ms, "host_wait_ms": un_st.host_wait_ms, "open_ms": open_s * 1e3,
                    "how": "default pd_open: pread into the page-locked staging ring, then cudaMemcpyAsync, per step"}
        t0 = time.perf_counter()
        cpin = PgenContext(path, devices=devices, pin_file=True)
        pin_open_s = time.perf_counter() - t0
        pinned_ok = bool(cpin.info().pinned_file)
        e2e_s, st = time_file_fed(cpin, e2e_steps)
        assert np.array_equal(out_counts[:V_total], file_counts) and np.array_equal(out_p[:V_total], file_p)
        cpin.close()
        e2e = {
            "value": calls_per_step / e2e_s, "unit": UNIT, "h2d_bytes_per_step": int(st.h2d_bytes // e2e_steps),
            "d2h_bytes_per_step": V_total * 24, "ms_per_step": e2e_s * 1e3, "steps": e2e_steps,
            "variants": V_total, "gbs": V_total * W / e2e_s / 1e9,
            "source": f"pd_open('{sample_dir()}/...pgen', PD_OPEN_PIN_FILE) on tmpfs (the file the reference arm reads): "
                      f"the file's mapping is page-locked once at open ({pin_open_s:.2f} s, outside the timed region "
                      f"like the reference arm's reader opens); every step DMAs all records from the page cache "
                      f"(cudaMemcpyAsync, double-buffered) and copies the results back, through the C ABI",
            "pinned_file": pinned_ok, "open_ms": pin_open_s * 1e3,
            "host_stage_ms": st.host_stage_ms, "host_wait_ms": st.host_wait_ms,
            "h2d_ceiling_gbs": ceiling,
            "h2d_ceiling_how": f"raw concurrent cudaMemcpyAsync of 1 GiB page-locked blocks to {n_gpus} GPU(s), this run",
            "unpinned_file": unpinned,
        }
        e2e["fraction_of_h2d_ceiling"] = e2e["gbs"] / ceiling
        unpinned["fraction_of_h2d_ceiling"] = unpinned["gbs"] / ceiling

        # ---- e2e from a page-locked in-memory image (no host copy in front of the DMA) ----
        if not args.skip_pinned and n_gpus <= 2:  # at N > 2 the pinned file covers it (and host RAM is the fileset's)
            budget = int(0.45 * mem_available_bytes())
            V_img = int(min(V_total, max(0, budget - 12) // W))
            if V_img >= 1024:
                import synth

                img = read_file_pinned(path, 12 + V_img * W)
                img[:12].copy_(torch.from_numpy(np.frombuffer(synth.fixed_header(V_img, N), dtype=np.uint8).copy()))
                with PgenContext.from_pointer(img.data_ptr(), img.numel(), devices=devices) as cimg:
                    cimg.hardy(0, V_img, out_counts=out_counts, out_p=out_p)
                    t0 = time.perf_counter()
                    for _ in range(e2e_steps):
                        cimg.hardy(0, V_img, out_counts=out_counts, out_p=out_p)
                    s = (time.perf_counter() - t0) / e2e_steps
                assert np.array_equal(out_counts[:V_img], file_counts[:V_img])
                e2e["pinned_image"] = {"value": float(V_img) * N / s, "unit": UNIT, "ms_per_step": s * 1e3,
                                       "variants": V_img, "gbs": V_img * W / s / 1e9,
                                       "fraction_of_h2d_ceiling": V_img * W / s / 1e9 / ceiling}
                del img

        # ---- device-resident: the kernel-side number ----
        ctx.load_resident(0, V_total)
        for _ in range(max(3, args.warmup)):
            ctx.hardy(0, V_total, out_counts=out_counts, out_p=out_p)
        assert np.array_equal(out_counts[:V_total], file_counts), "resident and streamed counts differ"
        assert np.array_equal(out_p[:V_total], file_p), "resident and streamed p-values differ"
        ctx.reset_stats()
        sampler = ClockSampler(0)
        sampler.start()
        call_ms, kern_ms, hwe_ms = [], [], []
        t0 = time.perf_counter()
        for _ in range(args.steps):
            ctx.hardy(0, V_total, out_counts=out_counts, out_p=out_p)
            st = ctx.stats()
            call_ms.append(st.call_ms)
            kern_ms.append(st.kernel_ms)
            hwe_ms.append(st.aux_kernel_ms)
        wall_ms = (time.perf_counter() - t0) * 1e3 / args.steps
        clocks = sampler.stop()
        launches = int(ctx.stats().launches)
        ms_per_step = float(np.mean(call_ms))  # slowest GPU's CUDA-event time of the call (max over ranks)
        value = calls_per_step / (ms_per_step * 1e-3)

        parity = parity_at_bench_size(path, V_total, N, out_counts, out_p)

        # ---- N > 1: plink_score with the all-reduce in the timed region ----
        score = None
        if n_gpus > 1 and not args.skip_score:
            K = 16
            rng = np.random.default_rng(SEED + 1)
            vidx = np.arange(V_total, dtype=np.uint32)
            Wt = rng.standard_normal((V_total, K))
            Wt[:, 0] = 1.0
            Wt[:, 1] = 2.0 * Wt[:, 2]
            het, hom, mis = ctx.sample_counts(0, V_total)
            sc_ar_ms = st_sc = None
            ctx.score(vidx, None, Wt, mode=1)  # warm-up
            t0 = time.perf_counter()
            s, d, a = ctx.score(vidx, None, Wt, mode=1)
            sc_wall = time.perf_counter() - t0
            st_sc = ctx.stats()
            # exact properties (no_mean_imputation, no flips): NAMED_ALLELE_DOSAGE_SUM = het + 2 hom_alt and
            # ALLELE_CT = 2 (V - missing) per sample, bit for bit against the all-reduced integer accumulators;
            # a weight column of ones reproduces the dosage sum; columns are linear in the weights
            assert np.array_equal(d, het.astype(np.float64) + 2.0 * hom.astype(np.float64))
            assert np.array_equal(a, (2 * (V_total - mis.astype(np.int64))).astype(np.uint32))
            assert np.array_equal(s[:, 0], d)
            assert np.allclose(s[:, 1], 2.0 * s[:, 2], rtol=1e-12, atol=1e-9)
            score = {
                "workload": f"plink_score no_mean_imputation, {K} weight columns x {V_total} variants x {N} samples "
                            f"(configs[4] shape: the same calls per GPU as 125k x 500k), records resident",
                "value": calls_per_step / sc_wall, "unit": UNIT, "ms": sc_wall * 1e3,
                "device_ms_slowest_gpu": st_sc.call_ms, "allreduce_ms": st_sc.allreduce_ms,
                "allreduce_bytes_per_gpu": int(st_sc.allreduce_bytes),
                "collective": "ncclAllReduce(sum) of double[N x K] + double[N] + uint32[N], in place in HBM, inside the call",
                "parity": "dosage sums / allele counts bit-exact vs the all-reduced pd_sample_counts; linearity 1e-12",
            }
        ctx.close()

        peak, peak_kind = peaks()
        fused = float(np.mean(hwe_ms)) == 0.0
        k_ms = float(np.mean(kern_ms))
        alg_bytes = float(V) * W  # per GPU and launch
        traffic, traffic_src = measured_traffic(alg_bytes, fused)
        achieved = alg_bytes / (k_ms * 1e-3) / 1e9

        cpu = None
        if n_gpus == 1 and not args.skip_cpu:
            cpu = cpu_baseline(path, V_total, N)
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": n_gpus, "steps": args.steps,
            "warmup": max(3, args.warmup), "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": DTYPE, "data": "synthetic",
            "config": {
                "workload": f"plink_hardy, fixed-width 2-bit .pgen (mode 0x02), {V} variants x {N} samples per GPU "
                            f"({alg_bytes / 1e9:.1f} GB packed), contiguous variant range per GPU",
                "variants_per_gpu": V, "samples": N, "l2": "inputs (>=10 GB) larger than the 126 MB L2, no flush",
                "partition": f"variant-range x{n_gpus}, one pd_ctx over {n_gpus} GPU(s) in one process "
                             f"(pd_open_opts.n_devices)",
                "fileset_bytes": file_bytes, "fileset_generation_s": gen_s,
            },
            "e2e": e2e,
            "gpu_launches": launches,
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                         "traffic": traffic, "traffic_source": traffic_src,
                         "peak_source": f"MEASURED_PEAKS.json ({peak_kind})",
                         "kernel": "k_hardy_fixed (counts + exact test fused)" if fused else "k_count_fixed",
                         "kernel_ms": k_ms, "hwe_kernel_ms": float(np.mean(hwe_ms)),
                         "per": "GPU (slowest of the N)"},
            "wall_ms_per_step": wall_ms,
            "hbm_gbs_all_gpus": achieved * n_gpus,
            "parity": parity,
            "clocks": clocks,
        }
        if score is not None:
            line["score"] = score
        if cpu is not None:
            line["cpu_baseline"] = cpu
    finally:
        if os.path.exists(path):
            os.remove(path)
    if world > 1:
        import torch.distributed as dist

        dist.barrier()
        dist.destroy_process_group()
    if line is not None:
        print(json.dumps(line))


def cpu_scan_lines(path, V, N, reps):
    """The reference's own code path on this box's host cores: T readers, fetch_add(128) batches,
    PgrGetCounts + HweLnP.  Thread counts 16 (the reference's default cap, src/plink_common.cpp:1910) and all
    cores; the reference-flags build (SSE2) and the AVX2 build (the 'best CPU' line)."""
    from oracle import ref

    ncpu = os.cpu_count() or 1
    lines = []
    ref.bench_scan(path, ncpu, 0, min(V, 4096), 1)  # touch pages / warm caches
    for avx2 in (False, True):
        if avx2 and not ref.avx2_available():
            continue
        for threads in sorted({min(16, ncpu), ncpu}):
            times = [ref.bench_scan(path, threads, 0, V, 1, avx2=avx2)[0] for _ in range(reps)]
            lines.append({"build": "x86-64-v3 (AVX2)" if avx2 else "reference flags (-O2, SSE2)", "threads": threads,
                          "value": float(V) * N / min(times), "best_s": min(times), "mean_s": float(np.mean(times))})
    return lines


def cpu_baseline(path, V, N):
    from oracle import ref

    if not ref.available():
        return {"value": None, "unit": UNIT, "cores": 0, "kind": "reference",
                "sample": "oracle/_ref/libpgenref.so not present on this box"}
    lines = cpu_scan_lines(path, V, N, reps=3)
    main = max((ln for ln in lines if not ln["build"].startswith("x86")), key=lambda ln: ln["value"])
    return {
        "value": main["value"], "unit": UNIT, "cores": main["threads"], "kind": "reference",
        "sample": f"all {V} variants x {N} samples of the same tmpfs file, pgenlib PgrGetCounts + HweLnP, "
                  f"{main['threads']} reader threads, best of 3 ({main['best_s']:.3f} s)",
        "lines": lines,
    }


def run_reference(args):
    """The reference's own CPU implementation of the path (oracle/_ref = its pgenlib compiled as is) with
    every host thread, over the whole per-GPU workload (1M x 100k) on tmpfs."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from oracle import ref

    if not ref.available():
        print(json.dumps({"impl": "reference", "unavailable": "oracle/_ref/libpgenref.so was not built"}))
        return
    N = args.samples
    V = args.variants or 1_000_000
    path = os.path.join(sample_dir(), f"plinkgpu_ref_{os.getpid()}.pgen")
    try:
        import torch

        have_gpu = torch.cuda.is_available()
    except Exception:
        have_gpu = False
    try:
        if have_gpu:
            write_fileset(path, V, N, 1)
        else:
            import synth

            V = min(V, 2000)
            synth.write_fixed_pgen(path, synth.gen_hwe_genotypes(V, N, SEED))
        threads = os.cpu_count() or 1
        for _ in range(max(1, min(args.warmup, 2))):
            ref.bench_scan(path, threads, 0, V, 1)
        times = []
        for _ in range(args.steps):
            t, _cs = ref.bench_scan(path, threads, 0, V, 1)
            times.append(t)
    finally:
        if os.path.exists(path):
            os.remove(path)
    ms = float(np.mean(times)) * 1e3
    value = float(V) * N / (ms * 1e-3)
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": DTYPE, "data": "synthetic",
        "config": {"workload": f"plink_hardy, fixed-width 2-bit .pgen (mode 0x02), {V} variants x {N} samples "
                               f"({V * ((N + 3) // 4) / 1e9:.1f} GB packed) per step: one GPU's range of the GPU arm",
                   "variants_per_gpu": V, "samples": N},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": threads, "kind": "reference",
                         "sample": f"{V} variants x {N} samples on tmpfs, pgenlib PgrGetCounts + HweLnP, "
                                   f"{threads} reader threads claiming 128-variant batches"},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }))


def main():
    args = parse_args()
    # stdout carries exactly ONE line, the JSON result: libraries that write to file descriptor 1 (NCCL prints its
    # version there) are sent to stderr for the length of the run
    sys.stdout.flush()
    real_stdout = os.dup(1)
    os.dup2(2, 1)
    buf = io.StringIO()
    try:
        with contextlib.redirect_stdout(buf):
            if args.impl == "reference":
                run_reference(args)
            else:
                run_ours(args)
    finally:
        sys.stdout.flush()
        os.dup2(real_stdout, 1)
        os.close(real_stdout)
        out = buf.getvalue()
        if out:
            sys.stdout.write(out)
            sys.stdout.flush()


if __name__ == "__main__":
    main()
