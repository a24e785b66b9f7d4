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

Two drivers of the same layout live here:
  ShardedExchange     the product path: libchanagpu.so's chq_shard_* (NCCL and the peer reads inside the library,
                      one call per batch, each rank returns the finished result of its own range of messages);
  ShardedGpuMatcher   the all-gather form spelled out with torch.distributed between two C-ABI calls (every rank ends
                      with the whole batch's result) — kept as the readable statement of the data flow; the CPU tests
                      run its host logic over gloo with the oracle in place of the kernels.
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


class ShardedExchange:
    """One rank of a binding-sliced exchange through the C ABI (chq_shard_*, include/chanamq_gpu.h): one process per
    GPU.  `unique_id` = bytes from ShardedExchange.unique_id() on one rank, handed to the others by the caller."""

    @staticmethod
    def unique_id():
        import ctypes as C

        from ._lib import CHQ_OK, ChqError, load_library

        L = load_library()
        buf = (C.c_uint8 * 128)()
        rc = L.chq_shard_unique_id(buf)
        if rc != CHQ_OK:
            raise ChqError(rc, (L.chq_shard_last_error() or b"").decode())
        return bytes(buf)

    def __init__(self, unique_id, rank, world, device, exchange_type, mode):
        import ctypes as C

        from ._lib import load_library

        self.L = load_library()
        self.rank, self.world, self.device = rank, world, device
        self.h = C.c_void_p()
        idb = (C.c_uint8 * 128).from_buffer_copy(unique_id)
        self._check(self.L.chq_shard_create(C.byref(self.h), idb, rank, world, device, exchange_type, mode))

    def _check(self, rc):
        from ._lib import CHQ_OK, ChqError

        if rc != CHQ_OK:
            text = (self.L.chq_shard_last_error() or b"").decode("utf-8", "replace") or (self.L.chq_last_error() or b"").decode()
            raise ChqError(rc, text)

    def close(self):
        if self.h:
            self.L.chq_shard_destroy(self.h)
            self.h = None

    def bind_many(self, key_blob, key_off, queue_blob, queue_off, durable, first_index=0):
        """the WHOLE binding list on every rank: binding i is kept by rank (first_index + i) mod world"""
        kb = np.ascontiguousarray(key_blob, dtype=np.uint8)
        ko = np.ascontiguousarray(key_off, dtype=np.uint32)
        qb = np.ascontiguousarray(queue_blob, dtype=np.uint8)
        qo = np.ascontiguousarray(queue_off, dtype=np.uint32)
        du = np.ascontiguousarray(durable, dtype=np.uint8)
        self._check(self.L.chq_shard_bind_many(self.h, kb.ctypes.data, ko.ctypes.data, qb.ctypes.data, qo.ctypes.data,
                                               du.ctypes.data, len(ko) - 1, first_index))

    def commit(self):
        self._check(self.L.chq_shard_commit(self.h))

    @property
    def disjoint(self):
        return bool(self.L.chq_shard_is_disjoint(self.h))

    def reserve(self, max_keys, slice_cap, pull_cap):
        self._check(self.L.chq_shard_reserve(self.h, max_keys, slice_cap, pull_cap))

    def route_device(self, d_blob, d_off, n, d_offsets, d_ids, ids_cap, d_durable, d_err, d_total, d_status, stream=0):
        """raw device pointers; asynchronous on `stream` -> (first_message, end_message) of this rank's range"""
        import ctypes as C

        m0, m1 = C.c_uint32(), C.c_uint32()
        self._check(self.L.chq_shard_route_device(self.h, d_blob, d_off, n, d_offsets, d_ids, ids_cap, d_durable, d_err, d_total,
                                                  d_status, C.byref(m0), C.byref(m1), stream))
        return m0.value, m1.value

    def stage_ms(self, enable):
        """-> ({stage: ms summed since the last call}, batches); then switch recording on/off"""
        import ctypes as C

        ms, n = (C.c_double * 4)(), C.c_uint32()
        self._check(self.L.chq_shard_stage_ms(self.h, int(enable), ms, C.byref(n)))
        return dict(zip(("local_route", "all_gather", "pull", "merge"), list(ms))), n.value
