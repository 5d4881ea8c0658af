"""Variant-range partition of one .pgen over the GPUs of a box (one process per GPU).

What the reference does with up to 16 scan threads and `fetch_add` batch claims
(src/plink_freq.cpp:434-441) plus mutex merges of per-thread accumulators
(src/plink_missing.cpp:614-619, src/plink_score.cpp:657-664) becomes: contiguous variant ranges of
equal record BYTES, one per rank; per-variant outputs need no communication (disjoint row slices);
per-sample accumulators are summed with ONE all-reduce over NVLink (NCCL) on the device buffers the
library leaves in HBM (pd_device_result), or over gloo in the CPU tests.
"""
from __future__ import annotations

import ctypes as C
from typing import List, Optional, Sequence, Tuple

import numpy as np


def partition_variants(fpos: np.ndarray, vrtypes: Optional[np.ndarray], world: int,
                       v_begin: int = 0, v_end: Optional[int] = None) -> List[Tuple[int, int]]:
    """Split [v_begin, v_end) into `world` contiguous ranges of (nearly) equal record bytes.
    Range starts are moved back to the closest variant that is not LD-compressed (vrtype & 6 != 2), so no
    rank's first record depends on a record owned by the previous rank more than necessary; the library
    re-reads a missing LD base itself (ExtendToBase), so any cut is correct -- this only avoids the re-read."""
    fpos = np.asarray(fpos, dtype=np.uint64)
    V = fpos.size - 1
    v_end = V if v_end is None else v_end
    if not (0 <= v_begin <= v_end <= V):
        raise ValueError("variant range out of bounds")
    if world <= 0:
        raise ValueError("world must be positive")
    cuts = [v_begin]
    total = int(fpos[v_end]) - int(fpos[v_begin])
    for r in range(1, world):
        target = int(fpos[v_begin]) + (total * r) // world
        v = int(np.searchsorted(fpos[v_begin:v_end + 1], np.uint64(target), side="left")) + v_begin
        v = min(max(v, cuts[-1]), v_end)
        if vrtypes is not None:
            while v > cuts[-1] and v < v_end and (int(vrtypes[v]) & 6) == 2:
                v -= 1
        cuts.append(v)
    cuts.append(v_end)
    return [(cuts[i], cuts[i + 1]) for i in range(world)]


def partition_list(vidx: np.ndarray, fpos: np.ndarray, world: int) -> List[Tuple[int, int]]:
    """Split an ascending variant LIST into `world` contiguous chunks of (nearly) equal record bytes
    (the scored-variant list of plink_score, src/plink_score.cpp:407).  Returns index ranges into vidx."""
    vidx = np.asarray(vidx, dtype=np.int64)
    fpos = np.asarray(fpos, dtype=np.uint64)
    if vidx.size == 0:
        return [(0, 0)] * world
    w = (fpos[vidx + 1] - fpos[vidx]).astype(np.float64)
    cum = np.concatenate([[0.0], np.cumsum(w)])
    cuts = [0]
    for r in range(1, world):
        i = int(np.searchsorted(cum, cum[-1] * r / world, side="left"))
        cuts.append(min(max(i, cuts[-1]), vidx.size))
    cuts.append(vidx.size)
    return [(cuts[i], cuts[i + 1]) for i in range(world)]


class _DevArray:
    """A raw device pointer dressed as __cuda_array_interface__ so torch can all-reduce it in place."""

    def __init__(self, ptr: int, n: int, typestr: str):
        self.__cuda_array_interface__ = {"data": (int(ptr), False), "shape": (int(n),), "typestr": typestr,
                                         "version": 2, "strides": None}


class ShardedPgen:
    """One rank's view: `local` does the work on this rank's GPU (a PgenContext), `dist` sums across ranks.

    local: object with variant_ct, sample_ct, index(), counts(v0, v1), missing_by_sample(v0, v1),
           sample_counts(v0, v1), score(vidx, flip, W, mode) -- PgenContext on a GPU; the CPU tests pass an
           oracle-backed stand-in to exercise the partition / reduction logic under gloo.
    """

    def __init__(self, local, rank: int = 0, world: int = 1, group=None, device=None, prepartitioned: bool = False):
        """prepartitioned: `local` already holds only this rank's contiguous variant range (one shard file per
        rank): every call works on the whole local range and only the reductions span the group."""
        self.local = local
        self.rank, self.world, self.group = rank, world, group
        self.device = device  # torch.device of this rank's GPU, or None for CPU (gloo) reductions
        self.prepartitioned = prepartitioned
        self.vrtypes, self.fpos = local.index()
        self.V = int(self.fpos.size - 1)
        self.ranges = [(0, self.V)] * world if prepartitioned else partition_variants(self.fpos, self.vrtypes, world)
        self.v0, self.v1 = self.ranges[rank]

    # ---- reductions ----
    def _allreduce_host(self, arr: np.ndarray) -> np.ndarray:
        if self.world == 1:
            return arr
        import torch
        import torch.distributed as dist

        t = torch.from_numpy(np.ascontiguousarray(arr))
        if arr.dtype == np.uint32:
            t = torch.from_numpy(np.ascontiguousarray(arr).view(np.int32))
        if self.device is not None:
            t = t.to(self.device)
        dist.all_reduce(t, op=dist.ReduceOp.SUM, group=self.group)
        out = t.cpu().numpy()
        return out.view(np.uint32) if arr.dtype == np.uint32 else out

    def _allreduce_device(self, which: int, typestr: str, itemsize: int) -> np.ndarray:
        """All-reduce a result the library left in HBM, in place, then read it back once."""
        import torch
        import torch.distributed as dist

        from . import _native as nat

        ptr, nbytes = C.c_void_p(), C.c_uint64()
        self.local._check(self.local._lib.pd_device_result(self.local._h, which, C.byref(ptr), C.byref(nbytes)))
        t = torch.as_tensor(_DevArray(ptr.value, nbytes.value // itemsize, typestr), device=self.device)
        if self.world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.SUM, group=self.group)
        return t.cpu().numpy()

    def _on_gpu(self) -> bool:
        return self.device is not None and hasattr(self.local, "_h")

    # ---- per-variant outputs: disjoint row ranges, no collective on the data path ----
    def counts_local(self) -> np.ndarray:
        return self.local.counts(self.v0, self.v1)

    def counts(self) -> np.ndarray:
        """Counts of ALL variants on every rank (convenience for tests / small outputs): each rank fills its
        own rows of a zero matrix, one sum all-reduce stitches them."""
        full = np.zeros((self.V, 4), dtype=np.uint32)
        full[self.v0:self.v1] = self.counts_local()
        return self._allreduce_host(full.reshape(-1)).reshape(self.V, 4)

    # ---- per-sample accumulators: one all-reduce ----
    def missing_by_sample(self) -> np.ndarray:
        if self._on_gpu():
            from . import _native as nat

            self.local._check(self.local._lib.pd_missing_by_sample(self.local._h, self.v0, self.v1, None))
            n = self.local.sample_ct
            return self._allreduce_device(nat.PD_RESULT_SAMPLE_COUNTS, "<i4", 4).view(np.uint32)[2 * n:3 * n].copy()
        return self._allreduce_host(self.local.missing_by_sample(self.v0, self.v1))

    def sample_counts(self):
        if self._on_gpu():
            from . import _native as nat

            self.local._check(self.local._lib.pd_sample_counts(self.local._h, self.v0, self.v1, None, None, None))
            n = self.local.sample_ct
            a = self._allreduce_device(nat.PD_RESULT_SAMPLE_COUNTS, "<i4", 4).view(np.uint32)
            return a[:n].copy(), a[n:2 * n].copy(), a[2 * n:3 * n].copy()
        het, hom, mis = self.local.sample_counts(self.v0, self.v1)
        s = self._allreduce_host(np.concatenate([het, hom, mis]))
        n = het.size
        return s[:n], s[n:2 * n], s[2 * n:]

    def score(self, vidx: Sequence[int], flip, weights: np.ndarray, mode: int = 0):
        """plink_score over an ascending scored-variant list, K weight columns: the list is cut into
        contiguous chunks of equal record bytes, each rank scores its chunk, partial sums are all-reduced."""
        vidx = np.ascontiguousarray(vidx, dtype=np.uint32)
        W = np.ascontiguousarray(weights, dtype=np.float64)
        if W.ndim == 1:
            W = W.reshape(-1, 1)
        K = W.shape[1]
        i0, i1 = (0, vidx.size) if self.prepartitioned else partition_list(vidx, self.fpos, self.world)[self.rank]
        fl = None if flip is None else np.ascontiguousarray(flip, dtype=np.uint8)[i0:i1]
        n = self.local.sample_ct
        if self._on_gpu():
            from . import _native as nat

            def p(a):
                return a.ctypes.data_as(C.c_void_p) if a is not None else None

            v_loc, w_loc = vidx[i0:i1].copy(), W[i0:i1].copy()
            self.local._check(self.local._lib.pd_score(self.local._h, p(v_loc), p(fl), p(w_loc), int(i1 - i0), K,
                                                       int(mode), None, None, None))
            s = self._allreduce_device(nat.PD_RESULT_SCORE, "<f8", 8).reshape(n, K)
            d = self._allreduce_device(nat.PD_RESULT_DOSAGE_SUM, "<f8", 8)
            a = self._allreduce_device(nat.PD_RESULT_ALLELE_CT, "<i4", 4).view(np.uint32)
            return s, d, a
        s, d, a = self.local.score(vidx[i0:i1], fl, W[i0:i1], mode)
        flat = self._allreduce_host(np.concatenate([s.reshape(-1), d]))
        a = self._allreduce_host(a)
        return flat[:n * K].reshape(n, K), flat[n * K:], a
