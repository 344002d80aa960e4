"""Host side of the sharded multi-GPU outer sync (one process per GPU).

`SyncContext` owns an `nd_sync_ctx` (include/nanodiloco_b200.h): an NCCL
communicator of its own over the ranks of a torch process group, side streams
and events.  torch.distributed is used only as plumbing -- to hand the NCCL
unique id from rank 0 to the other ranks.  It replaces, for this path, the
process group the reference creates at nanodiloco/training_utils/utils.py:41-43
and uses per tensor at nanodiloco/diloco/diloco.py:49.
"""
from __future__ import annotations

import ctypes
from ctypes import c_int64 as _c_i64

import torch
import torch.distributed as dist

from . import _lib


def shard_index_map(n_padded: int, world: int, rank: int, bucket: int) -> torch.Tensor:
    """Global element index of every element of `rank`'s shard, in shard order.

    Bucket b covers [b*bucket, min((b+1)*bucket, n_padded)); it is cut into `world` equal
    slices and rank r owns slice r.  Pure host logic (CPU tensor), shared with the tests."""
    parts = []
    off = 0
    while off < n_padded:
        length = min(bucket, n_padded - off)
        assert length % world == 0
        sl = length // world
        parts.append(torch.arange(off + rank * sl, off + (rank + 1) * sl, dtype=torch.int64))
        off += length
    return torch.cat(parts) if parts else torch.zeros(0, dtype=torch.int64)


def hybrid_index_map(n_padded: int, n_nvls: int, world: int, rank: int, bucket: int) -> torch.Tensor:
    """Global element indices of `rank`'s shard under the hybrid layout (nd_outer_sync_hybrid): its slices
    of the in-switch region [0, n_nvls) in bucket order, then its contiguous piece of the peer-kernel
    region [n_nvls, n_padded).  Pure host logic (CPU tensor)."""
    a = shard_index_map(n_nvls, world, rank, bucket)
    sB = (n_padded - n_nvls) // world
    b = torch.arange(n_nvls + rank * sB, n_nvls + (rank + 1) * sB, dtype=torch.int64)
    return torch.cat([a, b])


class SyncContext:
    def __init__(self, world: int, rank: int, *, bucket_elems: int, n_streams: int = 2, group=None,
                 device=None, ag_comm=None):
        """ag_comm: give the all-gathers a communicator of their own (default: on, ND_NCCL_AG_COMM=0 turns
        it off) so that they do not queue behind the next bucket's reduce-scatter."""
        import os
        L = _lib.load()
        self._nccl_bufs = []
        self.world, self.rank, self.group = world, rank, group
        self.device = device if device is not None else torch.device("cuda", torch.cuda.current_device())
        uid = ctypes.create_string_buffer(_lib.ND_UNIQUE_ID_BYTES)
        if world > 1:
            if rank == 0:
                _lib.check(L.nd_nccl_unique_id(uid), "nd_nccl_unique_id")
            box = [uid.raw]
            src = dist.get_global_rank(group, 0) if group is not None else 0
            dist.broadcast_object_list(box, src=src, group=group)
            uid = ctypes.create_string_buffer(box[0], _lib.ND_UNIQUE_ID_BYTES)
        handle = ctypes.c_void_p()
        with torch.cuda.device(self.device):
            _lib.check(L.nd_sync_ctx_create(uid, world, rank, int(bucket_elems), int(n_streams),
                                            ctypes.byref(handle)), "nd_sync_ctx_create")
        self._h = handle
        self.bucket = int(L.nd_sync_bucket_elems(self._h))
        if ag_comm is None:
            ag_comm = os.environ.get("ND_NCCL_AG_COMM", "1") == "1"
        self.ag_comm = bool(ag_comm) and world > 1
        if self.ag_comm:
            uid2 = ctypes.create_string_buffer(_lib.ND_UNIQUE_ID_BYTES)
            if rank == 0:
                _lib.check(L.nd_nccl_unique_id(uid2), "nd_nccl_unique_id")
            box = [uid2.raw]
            dist.broadcast_object_list(box, src=dist.get_global_rank(group, 0) if group is not None else 0, group=group)
            uid2 = ctypes.create_string_buffer(box[0], _lib.ND_UNIQUE_ID_BYTES)
            with torch.cuda.device(self.device):
                _lib.check(L.nd_sync_ctx_add_ag_comm(self._h, uid2), "nd_sync_ctx_add_ag_comm")

    def alloc(self, numel: int, dtype=torch.float32) -> torch.Tensor:
        """A zeroed device tensor in memory from NCCL's allocator, registered with this context's
        communicators (nd_nccl_mem_alloc + nd_sync_register): what `local` and `pg` should live in for the
        reduce-scatter / all-gather to run on them without staging copies.  Freed by close()."""
        L = _lib.load()
        nbytes = int(numel) * torch.empty(0, dtype=dtype).element_size()
        ptr = ctypes.c_void_p()
        with torch.cuda.device(self.device):
            _lib.check(L.nd_nccl_mem_alloc(max(nbytes, 16), ctypes.byref(ptr)), "nd_nccl_mem_alloc")
            self._nccl_bufs.append(ptr.value)
            _lib.check(L.nd_sync_register(self._h, ptr.value, max(nbytes, 16)), "nd_sync_register")
        raw = _RawDeviceBuffer(ptr.value, nbytes)
        t = torch.as_tensor(raw, device=self.device).view(dtype)
        t._nd_keepalive = (raw, self)
        t.zero_()
        return t

    def set_profiling(self, on: bool) -> None:
        _lib.check(_lib.load().nd_sync_set_profiling(self._h, int(bool(on))), "nd_sync_set_profiling")

    def timeline(self):
        """Per-bucket stage times of the last outer_sync (needs set_profiling(True) before it): a list of
        dicts with milliseconds since the call started -- k1 = pseudo-gradient kernel, rs = reduce-scatter,
        k2 = shard update, ag = all-gather, refresh = master replica copy -- and each collective's bus GB/s
        (bytes every rank sends = (W-1)/W of the bucket, over the time that collective took)."""
        L = _lib.load()
        n = _c_i64(0)
        _lib.check(L.nd_sync_timeline(self._h, None, 0, ctypes.byref(n)), "nd_sync_timeline")
        if n.value == 0:
            return []
        buf = (ctypes.c_float * (7 * n.value))()
        _lib.check(L.nd_sync_timeline(self._h, buf, n.value, ctypes.byref(n)), "nd_sync_timeline")
        rows = []
        for b in range(n.value):
            k1s, k1e, rse, k2s, k2e, age, rfe = buf[7 * b: 7 * b + 7]
            rows.append({"bucket": b, "k1": (k1s, k1e), "rs": (k1e, rse), "k2": (k2s, k2e), "ag": (k2e, age),
                         "refresh": (age, rfe)})
        return rows

    def padded_elems(self, n: int) -> int:
        return int(_lib.load().nd_sync_padded_elems(self._h, n))

    def shard_elems(self, n: int) -> int:
        return int(_lib.load().nd_sync_shard_elems(self._h, n))

    def outer_sync(self, local, master, mom_shard, pg, *, lr, momentum, dampening=0.0, weight_decay=0.0,
                   nesterov=True, first_step=False, stream=None) -> None:
        _lib._require_cuda(local, master, mom_shard, pg)
        _lib.check(_lib.load().nd_outer_sync(self._h, _lib._ptr(local), _lib._ptr(master), _lib._ptr(mom_shard),
                                             _lib._ptr(pg), local.numel(), lr, momentum, dampening, weight_decay,
                                             int(nesterov), int(first_step), _lib._stream(stream)),
                   "nd_outer_sync")

    def outer_step_offloaded(self, master_host, staging, chunk_elems, mom, local, *, lr, momentum,
                             dampening=0.0, weight_decay=0.0, nesterov=True, first_step=False, stream=None) -> None:
        """nd_outer_step_offloaded: last-sync copy in pinned host memory (reference semantics)."""
        if master_host.is_cuda or not master_host.is_pinned():
            raise _lib.NDError("master_host must be page-locked host memory")
        _lib._require_cuda(staging, mom, local)
        _lib.check(_lib.load().nd_outer_step_offloaded(
            self._h, master_host.data_ptr(), _lib._ptr(staging), int(chunk_elems), _lib._ptr(mom), _lib._ptr(local),
            _lib._dtype_code(local), local.numel(), lr, momentum, dampening, weight_decay, int(nesterov),
            int(first_step), _lib._stream(stream)), "nd_outer_step_offloaded")

    def last_ms(self) -> float:
        ms = ctypes.c_float()
        _lib.check(_lib.load().nd_sync_last_ms(self._h, ctypes.byref(ms)), "nd_sync_last_ms")
        return float(ms.value)

    def gather_shards(self, shard: torch.Tensor, n_padded: int) -> torch.Tensor:
        """Collective over the torch group: the full flat buffer assembled from every rank's shard."""
        parts = [torch.empty_like(shard) for _ in range(self.world)]
        dist.all_gather(parts, shard, group=self.group)
        full = torch.empty(n_padded, dtype=shard.dtype, device=shard.device)
        for r, part in enumerate(parts):
            full[shard_index_map(n_padded, self.world, r, self.bucket).to(shard.device)] = part
        return full

    def close(self) -> None:
        if getattr(self, "_h", None):
            torch.cuda.synchronize(self.device)
            _lib.load().nd_sync_ctx_destroy(self._h)     # deregisters the buffers before they are freed
            self._h = None
            for p in self._nccl_bufs:
                _lib.load().nd_nccl_mem_free(p)
            self._nccl_bufs = []

    def __del__(self):  # pragma: no cover
        try:
            self.close()
        except Exception:
            pass


class _RawDeviceBuffer:
    """A device allocation owned by the C library, exposed to torch through the CUDA array interface."""

    def __init__(self, ptr: int, nbytes: int):
        self.ptr, self.nbytes = ptr, nbytes
        self.__cuda_array_interface__ = {"shape": (nbytes,), "typestr": "|u1", "data": (ptr, False), "version": 2}


class PeerContext:
    """Host side of the peer-memory outer sync (include/nanodiloco_b200.h, `nd_outer_sync_p2p`).

    Allocates this rank's shared working-parameter buffer, exchanges CUDA IPC handles over the torch
    process group (plumbing only) and maps the peers' buffers.  `local(dtype)` is the torch view of
    the shared buffer: parameters re-homed into it are readable and writable by every rank's kernel.
    """

    def __init__(self, world: int, rank: int, n_padded: int, *, elem_bytes: int = 4, group=None, device=None,
                 replicas: int = 1, exchange: str = "auto", fold=None):
        """exchange: "pull" (the owner's kernel reads the remote copies), "push" (every worker writes its
        slices into the owners' staging slots first; needs (world/replicas)/world of the working copy's
        size in extra peer memory per rank) or "auto" (ND_P2P_MODE if set, else pull).
        fold: False brackets the pull kernel with two barrier launches instead of folding the rendezvous into it."""
        import os
        L = _lib.load()
        self.world, self.rank, self.group = world, rank, group
        self.device = device if device is not None else torch.device("cuda", torch.cuda.current_device())
        self.n_padded = int(n_padded)
        if exchange == "auto":
            exchange = os.environ.get("ND_P2P_MODE", "pull")
        if exchange not in ("pull", "push"):
            raise ValueError("exchange must be 'pull', 'push' or 'auto'")
        self.exchange = exchange if world > 1 else "pull"
        local_bytes = (self.n_padded * elem_bytes + 255) // 256 * 256
        staging_bytes = 0
        if self.exchange == "push":
            staging_bytes = (world // int(replicas)) * (self.n_padded // world) * elem_bytes
        self.nbytes = local_bytes + staging_bytes
        ptr = ctypes.c_void_p()
        handle = ctypes.create_string_buffer(_lib.ND_IPC_HANDLE_BYTES)
        with torch.cuda.device(self.device):
            _lib.check(L.nd_peer_alloc(self.nbytes, ctypes.byref(ptr), handle), "nd_peer_alloc")
            self._ptr = ptr.value
            handles = [None] * world
            if world > 1:
                dist.all_gather_object(handles, handle.raw, group=group)
            else:
                handles[0] = handle.raw
            blob = ctypes.create_string_buffer(b"".join(handles), world * _lib.ND_IPC_HANDLE_BYTES)
            h = ctypes.c_void_p()
            _lib.check(L.nd_peer_ctx_create(world, rank, self._ptr, self.nbytes, blob, ctypes.byref(h)),
                       "nd_peer_ctx_create")
        self._h = h
        self.replicas = int(replicas)
        if self.replicas != 1:
            _lib.check(L.nd_peer_set_replicas(self._h, self.replicas), "nd_peer_set_replicas")
        if fold is not None:      # None: the library's default (one launch; ND_P2P_FOLD=0 for barrier launches)
            _lib.check(L.nd_peer_set_fold(self._h, int(bool(fold))), "nd_peer_set_fold")
        if staging_bytes:
            _lib.check(L.nd_peer_set_staging(self._h, local_bytes, staging_bytes), "nd_peer_set_staging")
            _lib.check(L.nd_peer_set_mode(self._h, _lib.ND_PEER_PUSH), "nd_peer_set_mode")
        self._local_bytes = self.n_padded * elem_bytes
        self._raw = _RawDeviceBuffer(self._ptr, self.nbytes)
        self._bytes = torch.as_tensor(self._raw, device=self.device)

    def local(self, dtype=torch.float32) -> torch.Tensor:
        return self._bytes[: self._local_bytes].view(dtype)

    def peer_view(self, w: int, dtype=torch.float32) -> torch.Tensor:
        """Rank w's buffer as mapped into this process (tests / debugging)."""
        p = ctypes.c_void_p()
        _lib.check(_lib.load().nd_peer_buffer(self._h, w, ctypes.byref(p)), "nd_peer_buffer")
        raw = _RawDeviceBuffer(p.value, self._local_bytes)
        t = torch.as_tensor(raw, device=self.device).view(dtype)
        t._nd_keepalive = (raw, self)
        return t

    def shard_range(self, rank=None):
        r = self.rank if rank is None else rank
        s = self.n_padded // self.world
        return r * s, (r + 1) * s

    def barrier(self, stream=None) -> None:
        _lib.check(_lib.load().nd_peer_barrier(self._h, _lib._stream(stream)), "nd_peer_barrier")

    def outer_sync(self, master_shard, mom_shard, *, dtype=torch.float32, lr, momentum, dampening=0.0,
                   weight_decay=0.0, nesterov=True, first_step=False, sumsq_out=None, scratch_buf=None,
                   stream=None) -> None:
        _lib._require_cuda(master_shard, mom_shard, sumsq_out, scratch_buf)
        code = _lib.ND_BF16 if dtype == torch.bfloat16 else _lib.ND_F32
        _lib.check(_lib.load().nd_outer_sync_p2p(self._h, _lib._ptr(master_shard), _lib._ptr(mom_shard), code,
                                                 self.n_padded, lr, momentum, dampening, weight_decay,
                                                 int(nesterov), int(first_step), _lib._ptr(sumsq_out),
                                                 _lib._ptr(scratch_buf), _lib._stream(stream)),
                   "nd_outer_sync_p2p")

    def outer_sync_offloaded(self, master_host_shard, staging, chunk_elems, mom_shard, *, dtype=torch.float32,
                             lr, momentum, dampening=0.0, weight_decay=0.0, nesterov=True, first_step=False,
                             stream=None) -> None:
        """nd_outer_sync_p2p_offloaded: this rank's master shard in pinned host memory."""
        if master_host_shard.is_cuda or not master_host_shard.is_pinned():
            raise _lib.NDError("master_host_shard must be page-locked host memory")
        _lib._require_cuda(staging, mom_shard)
        code = _lib.ND_BF16 if dtype == torch.bfloat16 else _lib.ND_F32
        _lib.check(_lib.load().nd_outer_sync_p2p_offloaded(
            self._h, master_host_shard.data_ptr(), _lib._ptr(staging), int(chunk_elems), _lib._ptr(mom_shard), code,
            self.n_padded, lr, momentum, dampening, weight_decay, int(nesterov), int(first_step),
            _lib._stream(stream)), "nd_outer_sync_p2p_offloaded")

    def last_ms(self) -> float:
        ms = ctypes.c_float()
        _lib.check(_lib.load().nd_peer_last_ms(self._h, ctypes.byref(ms)), "nd_peer_last_ms")
        return float(ms.value)

    def inner_step_replicas(self, params_ctx: "PeerContext", exp_avg, exp_avg_sq, *, dtype=torch.float32, lr, beta1=0.9,
                            beta2=0.999, eps=1e-8, weight_decay=1e-2, step, max_norm=1.0, sumsq_total_out=None,
                            stream=None) -> None:
        """nd_inner_step_replicas with THIS context holding the gradient buffers and `params_ctx` the working
        copies: gradient average over the replicas of this rank's worker + clip + AdamW on the owned part,
        new parameters written to every replica."""
        _lib._require_cuda(exp_avg, exp_avg_sq, sumsq_total_out)
        code = _lib.ND_BF16 if dtype == torch.bfloat16 else _lib.ND_F32
        _lib.check(_lib.load().nd_inner_step_replicas(
            params_ctx._h, self._h, _lib._ptr(exp_avg), _lib._ptr(exp_avg_sq), code, self.n_padded, lr, beta1, beta2, eps,
            weight_decay, int(step), max_norm, _lib._ptr(sumsq_total_out), _lib._stream(stream)),
            "nd_inner_step_replicas")

    def last_ms2(self):
        """(total, wait) device milliseconds of the last outer_sync: `wait` is the first flag barrier, i.e.
        the time spent waiting for the slowest worker to arrive."""
        tot, wait = ctypes.c_float(), ctypes.c_float()
        _lib.check(_lib.load().nd_peer_last_ms2(self._h, ctypes.byref(tot), ctypes.byref(wait)), "nd_peer_last_ms2")
        return float(tot.value), float(wait.value)

    def gather_shards(self, shard: torch.Tensor) -> torch.Tensor:
        """Collective over the torch group: contiguous shards -> the full flat buffer."""
        parts = [torch.empty_like(shard) for _ in range(self.world)]
        if self.world > 1:
            dist.all_gather(parts, shard, group=self.group)
        else:
            parts[0] = shard
        return torch.cat(parts)

    def close(self) -> None:
        if getattr(self, "_h", None):
            torch.cuda.synchronize(self.device)
            if self.world > 1 and dist.is_initialized():
                dist.barrier(group=self.group)      # nobody unmaps while a peer may still touch the buffer
            _lib.load().nd_peer_ctx_destroy(self._h)
            self._h = None
            # the buffer itself: every torch view handed out by local()/peer_view() dangles from here on
            self._bytes = None
            _lib.load().nd_peer_free(self._ptr)
            self._ptr = None


class NvlsContext:
    """Host side of `nd_outer_sync_nvls`: the cross-worker sum is taken inside the NVSwitch.

    Needs two buffers that every rank maps AND that have a multicast address: the working
    parameters and the pseudo-gradient staging.  They come from torch's symmetric memory
    (`torch.distributed._symmetric_memory`: CUDA VMM allocations bound to a multicast object) --
    plumbing, like the process group itself; the kernels that use the multicast addresses are ours.
    """

    def __init__(self, world: int, rank: int, n_padded: int, *, bucket_elems: int = 1 << 25, group=None,
                 device=None, nvls_fraction: float = 1.0):
        import torch.distributed._symmetric_memory as symm
        q = 4 * world
        self.bucket = (int(bucket_elems) + q - 1) // q * q
        # hybrid: the first n_nvls elements go through the switch, the rest through the peer kernel
        if not 0.0 <= nvls_fraction <= 1.0:
            raise ValueError("nvls_fraction must be in [0, 1]")
        self.n_nvls = min(int(n_padded), int(int(n_padded) * nvls_fraction) // q * q)
        if nvls_fraction >= 1.0:
            self.n_nvls = int(n_padded)
        L = _lib.load()
        self.world, self.rank, self.group = world, rank, group
        self.device = device if device is not None else torch.device("cuda", torch.cuda.current_device())
        self.n_padded = int(n_padded)
        payload = (self.n_padded * 4 + 255) // 256 * 256
        gname = (group if group is not None else dist.group.WORLD).group_name
        with torch.cuda.device(self.device):
            self._buf_local = symm.empty(payload + 4096, dtype=torch.uint8, device=self.device)
            self._buf_pg = symm.empty(payload, dtype=torch.uint8, device=self.device)
            self._buf_local.zero_()
            self._buf_pg.zero_()
            h_local = symm.rendezvous(self._buf_local, gname)
            h_pg = symm.rendezvous(self._buf_pg, gname)
            if not h_local.multicast_ptr or not h_pg.multicast_ptr:
                raise _lib.NDError("this system offers no NVLS multicast address for symmetric memory; use sync='p2p'")
            self._handles = (h_local, h_pg)
            self.mc_local, self.mc_pg = int(h_local.multicast_ptr), int(h_pg.multicast_ptr)
            bases = (ctypes.c_void_p * world)(*[int(p) for p in h_local.buffer_ptrs])
            h = ctypes.c_void_p()
            _lib.check(L.nd_peer_ctx_create_from_ptrs(world, rank, bases, payload, ctypes.byref(h)),
                       "nd_peer_ctx_create_from_ptrs")
            torch.cuda.synchronize(self.device)
        dist.barrier(group=group)           # every rank's flag page is zero before anyone signals
        self._h = h

    def local(self) -> torch.Tensor:
        return self._buf_local[: self.n_padded * 4].view(torch.float32)

    def pg(self) -> torch.Tensor:
        return self._buf_pg[: self.n_padded * 4].view(torch.float32)

    def shard_range(self, rank=None):
        r = self.rank if rank is None else rank
        s = self.n_padded // self.world
        return r * s, (r + 1) * s

    def outer_sync(self, master, mom_shard, *, lr, momentum, dampening=0.0, weight_decay=0.0, nesterov=True,
                   first_step=False, stream=None) -> None:
        _lib._require_cuda(master, mom_shard)
        _lib.check(_lib.load().nd_outer_sync_hybrid(
            self._h, self.local().data_ptr(), self.pg().data_ptr(), self.mc_local, self.mc_pg, _lib._ptr(master),
            _lib._ptr(mom_shard), self.n_padded, self.n_nvls, self.bucket, lr, momentum, dampening, weight_decay,
            int(nesterov), int(first_step), _lib._stream(stream)), "nd_outer_sync_hybrid")

    def last_ms(self) -> float:
        ms = ctypes.c_float()
        _lib.check(_lib.load().nd_peer_last_ms(self._h, ctypes.byref(ms)), "nd_peer_last_ms")
        return float(ms.value)

    def gather_shards(self, shard: torch.Tensor) -> torch.Tensor:
        """Collective: this layout's shards -> flat buffer."""
        parts = [torch.empty_like(shard) for _ in range(self.world)]
        dist.all_gather(parts, shard, group=self.group)
        full = torch.empty(self.n_padded, dtype=shard.dtype, device=shard.device)
        for r, part in enumerate(parts):
            full[self.shard_indices(r).to(shard.device)] = part
        return full

    def shard_indices(self, rank=None) -> torch.Tensor:
        """Global element indices of rank's shard, in shard order: its slices of the in-switch region
        (bucket-interleaved) followed by its contiguous piece of the peer-kernel region."""
        r = self.rank if rank is None else rank
        return hybrid_index_map(self.n_padded, self.n_nvls, self.world, r, self.bucket)

    def close(self) -> None:
        if getattr(self, "_h", None):
            torch.cuda.synchronize(self.device)
            if dist.is_initialized():
                dist.barrier(group=self.group)
            _lib.load().nd_peer_ctx_destroy(self._h)
            self._h = None
