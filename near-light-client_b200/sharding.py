"""Host logic of the coset-sharded commit.

A commit of N = n * 2^rate_bits leaves splits by LDE coset: after the index bit reversal coset k occupies the
contiguous leaf range [bitrev_r(k) * n, +n), so rank g of G = 2^s ranks owns the leaves [g, g+1) * N / G, i.e.
2^(rate_bits - s) whole cosets and 2^(cap_height - s) whole cap subtrees.  Each rank computes its shard with
nlc_commit_from_values_shard (no communication), and the only exchange on the path is the all-gather of the
cap entries, 32 bytes each: NCCL over NVLink on GPUs, gloo in the CPU tests of this logic.

When there are more ranks than cosets (starky commits at rate_bits 1 on 4 or 8 GPUs) the commit
shards twice: by COLUMN for the transforms (rank p owns the columns column_split(W, G)[p]: nlc_lde_columns_dev), then
one all-to-all hands rank g the leaf rows [g, g+1) * N / G of every column, and the tree is built by leaf ROW exactly
as for a coset shard (nlc_commit_from_lde_rows_dev).  pack_rows / exchange_rows below are that exchange:
8 * W * N * (G-1) / G bytes in total, uniform all-to-all through NVSwitch.  Two fused forms skip work:
nlc_lde_columns_scatter_dev stores the last NTT pass straight into the send layout (no pack_rows pass), or -- with
PeerRows -- straight into the other GPUs' receive buffers over NVLink (no send buffer, no NCCL all-to-all; one
stream_barrier separates the transforms from the tree).
"""
from __future__ import annotations

import numpy as np


class ShardPlanError(ValueError):
    pass


def check_plan(rate_bits: int, cap_height: int, n_shards: int) -> int:
    """Validate a shard count; returns log2(n_shards)."""
    if n_shards < 1 or n_shards & (n_shards - 1):
        raise ShardPlanError(f"n_shards = {n_shards} must be a power of two")
    s = n_shards.bit_length() - 1
    if s > rate_bits:
        raise ShardPlanError(f"coset sharding needs n_shards <= 2^rate_bits = {1 << rate_bits} (got {n_shards}); "
                             "shard independent commits across the remaining GPUs instead")
    if s > cap_height:
        raise ShardPlanError(f"coset sharding needs n_shards <= 2^cap_height = {1 << cap_height} (got {n_shards})")
    return s


def check_row_plan(cap_height: int, n_shards: int) -> int:
    """Leaf-row sharding (8e.3) only needs whole cap subtrees per shard; returns log2(n_shards)."""
    if n_shards < 1 or n_shards & (n_shards - 1):
        raise ShardPlanError(f"n_shards = {n_shards} must be a power of two")
    s = n_shards.bit_length() - 1
    if s > cap_height:
        raise ShardPlanError(f"leaf-row sharding needs n_shards <= 2^cap_height = {1 << cap_height} (got {n_shards})")
    return s


def column_split(W: int, n_shards: int):
    """[(first column, count)] per rank: contiguous, in rank order, sizes differing by at most one."""
    base, extra = divmod(W, n_shards)
    out, c0 = [], 0
    for p in range(n_shards):
        nc = base + (1 if p < extra else 0)
        out.append((c0, nc))
        c0 += nc
    return out


def uniform_column_split(W: int, n_shards: int):
    """[(first column, count)] per rank with EQUAL strides: rank p owns [p * per, min(W, (p + 1) * per)), per =
    ceil(W / n_shards) -- the layout of the gathered coefficient matrix of a coset-sharded commit (n_shards * per
    columns, the tail padding unused).  Returns (per, split)."""
    per = -(-W // n_shards)
    return per, [(min(W, p * per), max(0, min(W, (p + 1) * per) - min(W, p * per))) for p in range(n_shards)]


def pack_rows(lde_cols, n_shards: int):
    """lde_cols: this rank's columns, (nc, N) in leaf order (torch tensor or ndarray).  Returns the send buffer of the
    all-to-all: (n_shards, nc, N / n_shards) contiguous, block g = the leaves of shard g."""
    nc, N = lde_cols.shape
    v = lde_cols.reshape(nc, n_shards, N // n_shards)
    if isinstance(v, np.ndarray):
        return np.ascontiguousarray(v.transpose(1, 0, 2))
    return v.permute(1, 0, 2).contiguous()


def exchange_rows(send, W: int, group=None):
    """all_to_all of pack_rows' buffers (torch tensors on the backend's device).  Returns (W, N / G): every column's
    leaves of this rank's shard, column-major -- the layout nlc_commit_from_lde_rows_dev hashes."""
    import torch
    import torch.distributed as dist
    G = dist.get_world_size(group)
    per = send.shape[2]
    split = column_split(W, G)
    recv = torch.empty((W, per), dtype=send.dtype, device=send.device)
    dist.all_to_all_single(recv.view(-1), send.view(-1), output_split_sizes=[nc * per for _, nc in split],
                           input_split_sizes=[send.shape[1] * per] * G, group=group)
    return recv


def leaf_range(log_n: int, rate_bits: int, shard_idx: int, n_shards: int):
    """[first, last) leaf (index-bit-reversed LDE row) owned by a shard."""
    per = (1 << (log_n + rate_bits)) // n_shards
    return shard_idx * per, (shard_idx + 1) * per


def cosets_of(rate_bits: int, shard_idx: int, n_shards: int):
    """The cosets k (LDE point i = m * 2^r + k) a shard evaluates: those with bitrev_r(k) in its range."""
    per = (1 << rate_bits) // n_shards
    out = []
    for k in range(1 << rate_bits):
        kr = int(format(k, f"0{rate_bits}b")[::-1], 2) if rate_bits else 0
        if shard_idx * per <= kr < (shard_idx + 1) * per:
            out.append(k)
    return out


def cap_slice(cap_height: int, shard_idx: int, n_shards: int) -> slice:
    per = (1 << cap_height) // n_shards
    return slice(shard_idx * per, (shard_idx + 1) * per)


def rank_and_world(group=None):
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()):
        raise ShardPlanError("from_values_sharded without shard_idx needs an initialised torch.distributed group")
    return dist.get_rank(group), dist.get_world_size(group)


def gather_cap(local_cap: np.ndarray, cap_height: int, group=None, device=None) -> np.ndarray:
    """all_gather of the shards' cap entries -> the full MerkleCap (2^cap_height x 4), identical on every rank."""
    import torch
    import torch.distributed as dist
    world = dist.get_world_size(group)
    per = (1 << cap_height) // world
    local = np.ascontiguousarray(local_cap, dtype=np.uint64).reshape(per, 4)
    backend = dist.get_backend(group)
    t = torch.from_numpy(local.view(np.int64).copy())
    if backend == "nccl":
        t = t.cuda(device if device is not None and device >= 0 else torch.cuda.current_device())
    out = torch.empty((world * per, 4), dtype=torch.int64, device=t.device)
    dist.all_gather_into_tensor(out, t, group=group)
    return out.cpu().numpy().view(np.uint64).reshape(1 << cap_height, 4)


class PeerRows:
    """The receive buffers of a row-sharded commit, writable by every rank of the node over NVLink (CUDA IPC).
    Collective over `group`: every rank allocates W * per words, the 64-byte handles travel by all_gather_object,
    every rank maps the others' buffers.  ptrs[g] = rank g's buffer as seen from this process (own buffer for own
    rank).  close_peers() unmaps the others' buffers (after the transforms); free() releases the own one."""

    def __init__(self, ctx, words: int, group=None):
        import ctypes as C
        import torch.distributed as dist
        from . import _ffi
        self._ctx, self._ffi = ctx, _ffi
        self.rank, self.world = dist.get_rank(group), dist.get_world_size(group)
        own, handle = _ffi.vp(), C.create_string_buffer(64)
        ctx._check(_ffi.lib().nlc_peer_alloc(ctx._h, words, C.byref(own), handle))
        self.own = own.value
        handles = [None] * self.world
        dist.all_gather_object(handles, handle.raw, group=group)
        self.ptrs, self._opened = [], []
        for g, h in enumerate(handles):
            if g == self.rank:
                self.ptrs.append(self.own)
                continue
            p = _ffi.vp()
            ctx._check(_ffi.lib().nlc_peer_open(ctx._h, h, C.byref(p)))
            self.ptrs.append(p.value)
            self._opened.append(p.value)

    def close_peers(self):
        for p in self._opened:
            self._ctx._check(self._ffi.lib().nlc_peer_close(self._ctx._h, self._ffi.vp(p)))
        self._opened = []

    def free(self):
        self.close_peers()
        if self.own and self._ctx._h:
            self._ffi.lib().nlc_peer_free(self._ctx._h, self._ffi.vp(self.own))
        self.own = None

    def __del__(self):
        try:
            self.free()
        except Exception:
            pass


def stream_barrier(device, group=None):
    """A barrier in stream order (a 1-word all-reduce on the current stream): what every rank enqueued before it
    has completed on every rank when it completes."""
    import torch
    import torch.distributed as dist
    t = torch.zeros(1, dtype=torch.int32, device=device)
    dist.all_reduce(t, group=group)
    return t


def assemble(parts):
    """Concatenate per-shard arrays (cap entries, digests or leaf rows) in shard order: the single-GPU array."""
    return np.concatenate(list(parts), axis=0)
