#!/usr/bin/env python
"""bench.py -- genotype calls decoded+counted per second on B200 vs the reference's CPU scan.

Workload (BASELINE.json configs[1]): plink_hardy over a synthetic fixed-width (mode 0x02) .pgen of
1M variants x 100k samples (25 GB of packed 2-bit records) on one B200.  One "step" = one full
plink_hardy pass: genotype counts of every variant (K1) + exact HWE p-values (K8) + results copied
to host memory.

  value  : calls/s with the records resident in HBM when the timed region starts (CUDA-event time
           of the whole call on the library's stream, max over ranks).
  e2e    : the same pass through the C ABI from a page-locked HOST image of the file: every step
           re-stages all record bytes host->device (double-buffered cudaMemcpyAsync) and copies the
           results back; wall clock around the blocking calls.
  roofline: K1 count kernel, algorithmic bytes = V * ceil(N/4) per launch / CUDA-event duration,
           against the measured HBM copy bandwidth in MEASURED_PEAKS.json.
  cpu_baseline / --impl reference: the reference's own pgenlib (oracle/_ref, PgrGetCounts + HweLnP,
           one reader per thread, fetch_add(128) batches) on a bounded sample of the same file.

Inputs are larger than L2 (25 GB vs 126 MB), so no explicit L2 flush is needed between steps.
"""
import argparse
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


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--variants", type=int, default=0, help="override variants per rank (dev only)")
    ap.add_argument("--samples", type=int, default=100_000)
    ap.add_argument("--cpu-sample-variants", type=int, default=0)
    ap.add_argument("--e2e-variants", type=int, default=250_000, help="host-image variants per rank when N > 1")
    ap.add_argument("--e2e-cap-variants", type=int, default=0,
                    help="dev only: cap the host image at N = 1 (e.g. for the 125 GB configs[2] file)")
    ap.add_argument("--skip-cpu", action="store_true")
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


def generate_image_gpu(V: int, N: int, seed: int, device, out_u8):
    """Mode-0x02 file image (header + V records) into out_u8 (a uint8 tensor on `device` or on the host),
    generated on the GPU tile by tile: p_v ~ U(0.05, 0.5), g ~ Binomial(2, p_v), 1 % missing."""
    import torch
    import synth

    W = (N + 3) // 4
    hdr = np.frombuffer(synth.fixed_header(V, N), dtype=np.uint8)
    out_u8[:12].copy_(torch.from_numpy(hdr.copy()))
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
        out_u8[12 + v0 * W: 12 + (v0 + vt) * W].copy_(b.reshape(-1), non_blocking=True)
        del p, u, code, miss, c, b
    torch.cuda.synchronize(device)


def host_image_of(dev_image, V_host: int, N: int):
    """Page-locked host copy of the first V_host records of a device image, as its own mode-0x02 file."""
    import torch
    import synth

    W = (N + 3) // 4
    nbytes = 12 + V_host * W
    host = torch.empty(nbytes, dtype=torch.uint8, pin_memory=True)
    host.copy_(dev_image[:nbytes])
    host[:12].copy_(torch.from_numpy(np.frombuffer(synth.fixed_header(V_host, N), dtype=np.uint8).copy()))
    return host


def write_sample_file(image_u8, V_sample: int, N: int, path: str):
    """First V_sample records of the image as their own mode-0x02 file (page-cache / tmpfs resident)."""
    import synth

    W = (N + 3) // 4
    with open(path, "wb") as f:
        f.write(synth.fixed_header(V_sample, N))
        body = image_u8[12: 12 + V_sample * W].numpy()
        f.write(memoryview(body))


def sample_dir():
    return "/dev/shm" if os.path.isdir("/dev/shm") and os.access("/dev/shm", os.W_OK) else "/tmp"


def cpu_reference_run(path: str, V_sample: int, N: int, threads: int, reps: int):
    """Times the reference's own code path: T readers, fetch_add(128) batches, PgrGetCounts + HweLnP."""
    from oracle import ref

    ref.bench_scan(path, threads, 0, min(V_sample, 2048), 1)  # touch pages / warm caches
    best = None
    times = []
    for _ in range(reps):
        t, _cs = ref.bench_scan(path, threads, 0, V_sample, 1)
        times.append(t)
        best = t if best is None else min(best, t)
    return times


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


def bind_to_gpu_numa_node(local_rank: int):
    """Run this rank on the host cores next to its GPU, so that the page-locked host image is first-touched
    on that NUMA node and every rank's H2D traffic stays on its own socket's memory controllers and PCIe
    root.  Best effort: returns the node (or None when sysfs does not say)."""
    try:
        import torch

        bus = torch.cuda.get_device_properties(local_rank).pci_bus_id
        dom = torch.cuda.get_device_properties(local_rank).pci_domain_id
        dev = torch.cuda.get_device_properties(local_rank).pci_device_id
        path = f"/sys/bus/pci/devices/{dom:04x}:{bus:02x}:{dev:02x}.0/numa_node"
        node = int(open(path).read().strip())
        if node < 0:
            return None
        cpus = []
        for part in open(f"/sys/devices/system/node/node{node}/cpulist").read().strip().split(","):
            a, _, b = part.partition("-")
            cpus.extend(range(int(a), int(b or a) + 1))
        allowed = os.sched_getaffinity(0)
        cpus = [c for c in cpus if c in allowed]
        if not cpus:
            return None
        os.sched_setaffinity(0, cpus)
        return node
    except Exception:  # noqa: BLE001 -- sysfs layout differs between boxes; the bench runs either way
        return None


def run_ours(args):
    import torch
    import torch.distributed as dist

    from plinking_duck_b200.pgen import PgenContext

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device (the product path has no CPU fallback)")
    numa_node = bind_to_gpu_numa_node(local_rank) if world > 1 else None
    torch.cuda.set_device(local_rank)
    device = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=device)

    def barrier():
        torch.cuda.synchronize(device)
        if world > 1:
            dist.barrier()

    def max_over_ranks(x: float) -> float:
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device=device)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def sum_over_ranks(x: float) -> float:
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device=device)
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
        return float(t.item())

    N = args.samples
    # One rank = one contiguous variant range of the fileset (no data-path collective for hardy); every
    # rank holds the same amount of work at every N (weak scaling): V variants generated straight into HBM.
    V = args.variants or 1_000_000
    W = (N + 3) // 4
    nbytes = 12 + V * W
    dev_image = torch.empty(nbytes + 256, dtype=torch.uint8, device=device)
    generate_image_gpu(V, N, SEED + rank, device, dev_image)
    torch.cuda.empty_cache()
    # e2e: the same file as a page-locked HOST image.  At N = 1 it is the whole file (25 GB); with several
    # ranks sharing one host's RAM each rank stages the first 250k variants (6.25 GB) of its range.
    V_host = V if world == 1 else min(V, args.e2e_variants)
    if args.e2e_cap_variants:
        V_host = min(V_host, args.e2e_cap_variants)
    image = host_image_of(dev_image, V_host, N)

    out_counts = torch.empty((V, 4), dtype=torch.int32, pin_memory=True).numpy().view(np.uint32)
    out_p = torch.empty(V, dtype=torch.float64, pin_memory=True).numpy()

    calls_per_step = float(V) * float(N)

    # ---- e2e: host image -> staged H2D -> kernels -> results on host, every step ----
    ctx_host = PgenContext.from_pointer(image.data_ptr(), image.numel(), device=local_rank)
    e2e_steps = max(1, min(args.steps, 3))
    ctx_host.hardy(0, V_host, out_counts=out_counts, out_p=out_p)  # warm-up (allocations, clocks)
    first_counts = out_counts[:1024].copy()
    barrier()
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        ctx_host.hardy(0, V_host, out_counts=out_counts, out_p=out_p)
    barrier()
    e2e_s = max_over_ranks((time.perf_counter() - t0) / e2e_steps)
    e2e_stats = ctx_host.stats()
    h2d_per_step = int(e2e_stats.record_bytes)
    d2h_per_step = V_host * 24
    e2e_launches = int(e2e_stats.launches)
    ctx_host.close()

    # ---- device-resident: the kernel-side number ----
    ctx = PgenContext.from_device_pointer(dev_image.data_ptr(), nbytes, device=local_rank)
    for _ in range(max(3, args.warmup)):
        ctx.hardy(0, V, out_counts=out_counts, out_p=out_p)
    assert np.array_equal(out_counts[:1024], first_counts)
    ctx.reset_stats()
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    barrier()
    call_ms, kern_ms, hwe_ms = [], [], []
    t0 = time.perf_counter()
    for _ in range(args.steps):
        ctx.hardy(0, V, out_counts=out_counts, out_p=out_p)
        st = ctx.stats()
        call_ms.append(st.call_ms)
        kern_ms.append(st.kernel_ms)
        hwe_ms.append(st.aux_kernel_ms)
    barrier()
    wall_ms = (time.perf_counter() - t0) * 1e3 / args.steps
    clocks = sampler.stop() if rank == 0 else None
    launches = int(ctx.stats().launches)
    ms_per_step = max_over_ranks(float(np.mean(call_ms)))
    wall_ms = max_over_ranks(wall_ms)
    total_calls = sum_over_ranks(calls_per_step)
    value = total_calls / (ms_per_step * 1e-3)
    e2e_value = sum_over_ranks(float(V_host) * float(N)) / e2e_s

    # ---- sanity: spot-check counts against numpy on the host image ----
    body = image[12: 12 + 64 * W].numpy().reshape(64, W)
    two = np.stack([(body >> s) & 3 for s in (0, 2, 4, 6)], axis=2).reshape(64, W * 4)[:, :N]
    exp = np.stack([(two == k).sum(1) for k in range(4)], axis=1).astype(np.uint32)
    assert np.array_equal(out_counts[:64], exp), "GPU counts disagree with the host image"
    # size-independent properties over the whole range: the four counts of every variant add up to N,
    # every p-value is a probability
    assert np.all(out_counts[:V].reshape(V, 4).sum(axis=1, dtype=np.uint64) == N), "counts do not add up to N"
    assert np.all(np.isfinite(out_p[:V])) and float(out_p[:V].min()) >= 0.0 and float(out_p[:V].max()) <= 1.0

    peak, peak_kind = peaks()
    traffic, traffic_src = measured_traffic(float(V) * W, fused=float(np.mean(hwe_ms)) == 0.0)
    k_ms = float(np.mean(kern_ms))
    alg_bytes = float(V) * W
    achieved = alg_bytes / (k_ms * 1e-3) / 1e9

    line = None
    if rank == 0:
        cpu = None
        if world == 1 and not args.skip_cpu:
            from oracle import ref

            if ref.available():
                threads = os.cpu_count() or 1
                Vs = args.cpu_sample_variants or min(V, 200_000)
                path = os.path.join(sample_dir(), f"plinkgpu_bench_sample_{os.getpid()}.pgen")
                try:
                    write_sample_file(image, Vs, N, path)
                    times = cpu_reference_run(path, Vs, N, threads, reps=3)
                finally:
                    if os.path.exists(path):
                        os.remove(path)
                cpu = {
                    "value": float(Vs) * N / min(times), "unit": UNIT, "cores": threads, "kind": "reference",
                    "sample": f"first {Vs} variants x {N} samples of the same file on tmpfs (page-cache hot), "
                              f"pgenlib PgrGetCounts + HweLnP, {threads} reader threads, best of 3 "
                              f"({min(times):.3f} s)",
                }
            else:
                cpu = {"value": None, "unit": UNIT, "cores": 0, "kind": "reference",
                       "sample": "oracle/_ref/libpgenref.so not present on this box"}
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": max(3, args.warmup), "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "u32 popcount (counts) + f64 (HWE)", "data": "synthetic",
            "config": {
                "workload": f"plink_hardy, fixed-width 2-bit .pgen (mode 0x02), {V} variants x {N} samples per GPU "
                            f"({alg_bytes / 1e9:.1f} GB packed), contiguous variant range per rank",
                "variants_per_gpu": V, "samples": N, "l2": "inputs (>=10 GB) larger than the 126 MB L2, no flush",
                "partition": f"variant-range x{world}",
                "host_numa": "each rank bound to its GPU's NUMA node" if numa_node is not None else "unbound",
            },
            "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": h2d_per_step,
                    "d2h_bytes_per_step": d2h_per_step, "ms_per_step": e2e_s * 1e3, "steps": e2e_steps,
                    "variants_per_gpu": V_host,
                    "source": "page-locked host image of the .pgen, staged per step through the C ABI"},
            "gpu_launches": launches,
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                         "traffic": traffic, "traffic_source": traffic_src,
                         "peak_source": f"MEASURED_PEAKS.json ({peak_kind})",
                         "kernel": "k_hardy_fixed (counts + exact test fused)" if float(np.mean(hwe_ms)) == 0.0 else "k_count_fixed",
                         "kernel_ms": k_ms, "hwe_kernel_ms": float(np.mean(hwe_ms))},
            "wall_ms_per_step": wall_ms,
            "hbm_gbs_all_gpus": achieved * world,
            "clocks": clocks,
        }
        if cpu is not None:
            line["cpu_baseline"] = cpu
    ctx.close()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    if line is not None:
        print(json.dumps(line))


def run_reference(args):
    """The reference's own CPU implementation of the path (oracle/_ref = its pgenlib compiled as is) with
    every host thread, on a bounded sample of the same workload."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from oracle import ref

    if not ref.available():
        print(json.dumps({"impl": "reference", "unavailable": "oracle/_ref/libpgenref.so was not built"}))
        return
    N = args.samples
    Vs = args.cpu_sample_variants or 200_000
    if args.variants:
        Vs = min(Vs, args.variants)
    W = (N + 3) // 4
    path = os.path.join(sample_dir(), f"plinkgpu_ref_sample_{os.getpid()}.pgen")
    try:
        import torch

        have_gpu = torch.cuda.is_available()
    except Exception:
        have_gpu = False
    try:
        if have_gpu:
            import torch

            dev = torch.device("cuda", int(os.environ.get("LOCAL_RANK", "0")))
            image = torch.empty(12 + Vs * W, dtype=torch.uint8)
            generate_image_gpu(Vs, N, SEED, dev, image)
            image.numpy().tofile(path)
        else:
            import synth

            Vs = min(Vs, 2000)
            synth.write_fixed_pgen(path, synth.gen_hwe_genotypes(Vs, N, SEED))
        threads = os.cpu_count() or 1
        for _ in range(max(1, min(args.warmup, 2))):
            ref.bench_scan(path, threads, 0, Vs, 1)
        times = []
        for _ in range(args.steps):
            t, _cs = ref.bench_scan(path, threads, 0, Vs, 1)
            times.append(t)
    finally:
        if os.path.exists(path):
            os.remove(path)
    ms = float(np.mean(times)) * 1e3
    value = float(Vs) * N / (ms * 1e-3)
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "u32 popcount (counts) + f64 (HWE)", "data": "synthetic",
        "config": {"workload": f"plink_hardy, fixed-width 2-bit .pgen (mode 0x02), bounded sample of {Vs} variants x "
                               f"{N} samples per step (the GPU arm scans 1M variants)", "samples": N},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": threads, "kind": "reference",
                         "sample": f"{Vs} variants x {N} samples on tmpfs, pgenlib PgrGetCounts + HweLnP, "
                                   f"{threads} reader threads claiming 128-variant batches"},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }))


def main():
    args = parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
