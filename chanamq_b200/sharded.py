"""Binding-slice sharding of ONE exchange over the GPUs of a box (BASELINE.json configs[3]).

Binding i lives on rank i mod G; every rank routes ALL keys of a batch against its slice; the
per-slice CSR results are all-gathered (NCCL over NVLink) and merged per message into the sorted,
de-duplicated queue list on every rank (chq_merge_slices_device).  Every rank declares the queue
names of the WHOLE exchange, so queue ids agree across slices.

The reference has no counterpart (one exchange is one actor on one thread,
chana-mq-server/.../entity/ExchangeEntity.scala:107); this is the layout the north star prescribes
for tables that should be spread over several GPUs.  When the table fits one GPU — 1 M bindings
compile to ~110 MB — sharding by KEYS (chanamq_b200.GpuMatcher per rank, no collective at all) moves
far fewer bytes and is what bench.py times.

The collective is injected (`Comm`): torch.distributed for real runs (nccl on GPUs, gloo in the CPU
tests), so the host logic — slicing, declaring, gather layout, merge call — is testable without a GPU.
"""
from typing import NamedTuple

import numpy as np


def slice_of(i, world):
    """rank that owns binding i"""
    return i % world


def slice_bindings(key_blob, key_off, queue_blob, queue_off, durable, rank, world):
    """-> the packed bindings (same layout) that rank `rank` of `world` owns: indices rank, rank+G, ..."""
    n = len(key_off) - 1
    idx = np.arange(rank, n, world)
    kb, ko = take_fast(np.asarray(key_blob), np.asarray(key_off), idx)
    qb, qo = take_fast(np.asarray(queue_blob), np.asarray(queue_off), idx)
    return kb, ko, qb, qo, np.asarray(durable)[idx]


def take_fast(blob, off, idx):
    """gather of variable-length strings without a Python loop"""
    lens = (off[1:] - off[:-1])[idx].astype(np.int64)
    new_off = np.zeros(len(idx) + 1, dtype=np.uint32)
    new_off[1:] = np.cumsum(lens).astype(np.uint32)
    total = int(new_off[-1]) if len(idx) else 0
    if total == 0:
        return np.zeros(1, dtype=np.uint8), new_off
    starts = off[:-1][idx].astype(np.int64)
    # position p of the output belongs to string j = searchsorted(new_off, p, 'right') - 1
    owner = np.repeat(np.arange(len(idx)), lens)
    src = starts[owner] + (np.arange(total) - new_off[:-1].astype(np.int64)[owner])
    return blob[src], new_off


class SliceResult(NamedTuple):
    offsets: np.ndarray
    queue_ids: np.ndarray
    any_durable: np.ndarray


def merge_slices_host(offsets, ids, durable):
    """numpy statement of what chq_merge_slices_device computes (used by the CPU tests of the plumbing):
    offsets [G][n+1], ids list of G arrays, durable [G][n] -> union per message, sorted, unique"""
    g, n1 = offsets.shape
    n = n1 - 1
    out_off = np.zeros(n + 1, dtype=np.uint32)
    chunks = []
    for m in range(n):
        parts = [ids[r][offsets[r, m]:offsets[r, m + 1]] for r in range(g)]
        u = np.unique(np.concatenate(parts)) if parts else np.zeros(0, np.uint32)
        chunks.append(u.astype(np.uint32))
        out_off[m + 1] = out_off[m] + len(u)
    return SliceResult(out_off, np.concatenate(chunks) if chunks else np.zeros(0, np.uint32), durable.max(axis=0).astype(np.uint8))


class ShardedGpuMatcher:
    """One rank of a binding-sliced exchange.  Needs torch.distributed initialised (nccl) and a CUDA device."""

    def __init__(self, exchange_type, mode, rank, world, device):
        import torch

        from . import GpuMatcher

        self.torch = torch
        self.rank, self.world, self.device = rank, world, device
        self.local = GpuMatcher(exchange_type, mode, device=device)

    def load(self, key_blob, key_off, queue_blob, queue_off, durable):
        """every rank is given the WHOLE binding list; it keeps its slice and declares every queue name"""
        import ctypes as C

        from ._lib import check

        kb, ko, qb, qo, du = slice_bindings(key_blob, key_off, queue_blob, queue_off, durable, self.rank, self.world)
        self.local.bind_many(kb, ko, qb, qo, du)
        qb_all = np.ascontiguousarray(queue_blob, dtype=np.uint8)
        qo_all = np.ascontiguousarray(queue_off, dtype=np.uint32)
        check(self.local.L.chq_declare_queues(self.local.h, qb_all.ctypes.data, qo_all.ctypes.data, len(qo_all) - 1))
        self.local.commit()
        # does any queue have bindings on two slices?  If not, the merge need not de-duplicate (CHQ_MERGE_DISJOINT)
        raw, seen, self.disjoint = qb_all.tobytes(), {}, True
        for i in range(len(qo_all) - 1):
            name = raw[qo_all[i]:qo_all[i + 1]]
            if seen.setdefault(name, i % self.world) != i % self.world:
                self.disjoint = False
                break

    def route_device(self, d_blob, d_off, n, slice_cap, out_cap):
        """d_blob uint8 / d_off int32 torch tensors on this rank's device holding the SAME n keys on every rank
        -> (offsets int32[n+1], ids int32[total], any_durable uint8[n]) torch tensors, identical on every rank"""
        torch, dist = self.torch, self.torch.distributed
        dev = d_blob.device
        st = torch.cuda.current_stream().cuda_stream
        loc_off = torch.empty(n + 1, dtype=torch.int32, device=dev)
        loc_ids = torch.empty(slice_cap, dtype=torch.int32, device=dev)
        loc_dur = torch.empty(n, dtype=torch.uint8, device=dev)
        loc_err = torch.empty(n, dtype=torch.uint8, device=dev)
        tot = torch.zeros(4, dtype=torch.int64, device=dev)
        self.local.route_device(d_blob.data_ptr(), d_off.data_ptr(), n, loc_off.data_ptr(), loc_ids.data_ptr(), slice_cap,
                                loc_dur.data_ptr(), loc_err.data_ptr(), tot.data_ptr(), tot.data_ptr() + 8, st)
        g = self.world
        all_off = torch.empty(g * (n + 1), dtype=torch.int32, device=dev)
        all_ids = torch.empty(g * slice_cap, dtype=torch.int32, device=dev)
        all_dur = torch.empty(g * n, dtype=torch.uint8, device=dev)
        dist.all_gather_into_tensor(all_off, loc_off)
        dist.all_gather_into_tensor(all_ids, loc_ids)  # padded to slice_cap: one collective, no size exchange
        dist.all_gather_into_tensor(all_dur, loc_dur)
        out_off = torch.empty(n + 1, dtype=torch.int32, device=dev)
        out_ids = torch.empty(out_cap, dtype=torch.int32, device=dev)
        out_dur = torch.empty(n, dtype=torch.uint8, device=dev)
        from ._lib import check

        check(self.local.L.chq_merge_slices_device2(self.device, all_off.data_ptr(), all_ids.data_ptr(), slice_cap,
                                                    all_dur.data_ptr(), g, n, out_off.data_ptr(), out_ids.data_ptr(), out_cap,
                                                    out_dur.data_ptr(), tot.data_ptr() + 16, tot.data_ptr() + 24,
                                                    1 if self.disjoint else 0, st))
        torch.cuda.synchronize()
        s_local, s_merge = int(tot[1]) & 0xFFFFFFFF, int(tot[3]) & 0xFFFFFFFF
        if s_local or s_merge:
            raise RuntimeError("sharded route: status local %d merge %d (capacity too small?)" % (s_local, s_merge))
        return out_off, out_ids[: int(tot[2])], out_dur, loc_err

    def route_device_owned(self, d_blob, d_off, n, slice_cap, out_cap):
        """The same routing with the MERGE sharded too: rank g owns the messages [g*n/G, (g+1)*n/G).  Every rank still
        routes all n keys against its binding slice, but then only the small arrays (offsets, durable flags) are
        all-gathered; the queue ids travel by all-to-all — each rank receives, from every slice, just the ids of the
        messages it owns — and each rank merges its own range.  Every id crosses NVLink at most once and is merged
        once (the all-gather form moves and merges everything on every rank).
        -> (m0, m1, offsets int32[m1-m0+1], ids int32[total], any_durable uint8[m1-m0]) for this rank's range"""
        torch, dist = self.torch, self.torch.distributed
        dev = d_blob.device
        st = torch.cuda.current_stream().cuda_stream
        g = self.world
        marks = []

        def mk(name):   # stage marks for bench_sharded.py (self.trace = True): CUDA events on the current stream
            if getattr(self, "trace", False):
                e = torch.cuda.Event(enable_timing=True)
                e.record()
                marks.append((name, e))

        mk("start")
        loc_off = torch.empty(n + 1, dtype=torch.int32, device=dev)
        loc_ids = torch.empty(slice_cap, dtype=torch.int32, device=dev)
        loc_dur = torch.empty(n, dtype=torch.uint8, device=dev)
        loc_err = torch.empty(n, dtype=torch.uint8, device=dev)
        tot = torch.zeros(4, dtype=torch.int64, device=dev)
        self.local.route_device(d_blob.data_ptr(), d_off.data_ptr(), n, loc_off.data_ptr(), loc_ids.data_ptr(), slice_cap,
                                loc_dur.data_ptr(), loc_err.data_ptr(), tot.data_ptr(), tot.data_ptr() + 8, st)
        mk('local route')
        all_off = torch.empty(g * (n + 1), dtype=torch.int32, device=dev)
        all_dur = torch.empty(g * n, dtype=torch.uint8, device=dev)
        dist.all_gather_into_tensor(all_off, loc_off)
        dist.all_gather_into_tensor(all_dur, loc_dur)
        mk('all-gather offsets+flags')
        all_off = all_off.view(g, n + 1)
        bounds = [(r * n) // g for r in range(g + 1)]                      # message ranges of the owners
        cuts = all_off[:, bounds].to(torch.int64).cpu()                    # [slice][owner boundary]: one small read-back
        if int(tot[1].item()) & 0xFFFFFFFF:
            raise RuntimeError("sharded route: local status %d (slice_cap too small?)" % (int(tot[1].item()) & 0xFFFFFFFF))
        m0, m1 = bounds[self.rank], bounds[self.rank + 1]
        send = [int(cuts[self.rank, r + 1] - cuts[self.rank, r]) for r in range(g)]      # my slice's ids for owner r
        recv = [int(cuts[r, self.rank + 1] - cuts[r, self.rank]) for r in range(g)]      # slice r's ids for my range
        cap = max(max(recv), 1)
        got = torch.empty(g * cap, dtype=torch.int32, device=dev)
        base = int(cuts[self.rank, 0])
        ins, outs, at = [], [], base
        for r in range(g):
            ins.append(loc_ids[at:at + send[r]])
            at += send[r]
            outs.append(got[r * cap:r * cap + recv[r]])
        dist.all_to_all(outs, ins)
        mk('all-to-all ids')
        nn = m1 - m0
        sub_off = (all_off[:, m0:m1 + 1] - all_off[:, m0:m0 + 1]).contiguous()
        sub_dur = all_dur.view(g, n)[:, m0:m1].contiguous()
        mk('rebase')
        out_off = torch.empty(nn + 1, dtype=torch.int32, device=dev)
        out_ids = torch.empty(out_cap, dtype=torch.int32, device=dev)
        out_dur = torch.empty(max(nn, 1), dtype=torch.uint8, device=dev)
        from ._lib import check

        check(self.local.L.chq_merge_slices_device2(self.device, sub_off.data_ptr(), got.data_ptr(), cap, sub_dur.data_ptr(), g, nn,
                                                    out_off.data_ptr(), out_ids.data_ptr(), out_cap, out_dur.data_ptr(),
                                                    tot.data_ptr() + 16, tot.data_ptr() + 24, 1 if self.disjoint else 0, st))
        mk("merge")
        torch.cuda.synchronize()
        if marks:
            self.last_stages = {b[0]: a[1].elapsed_time(b[1]) for a, b in zip(marks, marks[1:])}
        s_merge = int(tot[3]) & 0xFFFFFFFF
        if s_merge:
            raise RuntimeError("sharded route: merge status %d (out_cap too small?)" % s_merge)
        return m0, m1, out_off, out_ids[: int(tot[2])], out_dur[:nn]
