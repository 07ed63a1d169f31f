"""Context-sharded decode: one sequence's KV pages spread over several GPUs.

Sequence sharding (``sharding.py``) covers batches of sequences; it cannot help a single 128 Ki+
token sequence, whose every decode step must stream the whole context through ONE GPU's HBM.  Here
the logical blocks of each sequence are dealt round-robin (in chunks of ``chunk_blocks`` pages) to
the ranks; each rank keeps the pages it owns in its own pool under rank-local block ids, runs the
paged-attention partition pass over them and exchanges one (max, sum, out) triple per (seq, head)
with its peers through NVLink peer stores inside the merge kernel
(``csrc/context_shard.cu``; C ABI ``b200_paged_attention_ctx_shard``).  No NCCL on the data path:
``torch.distributed`` only carries the 64-byte CUDA-IPC handles at start-up.

The merge is the reference's own partition-merge formula (attention_kernels.cu:565-669) applied
across ranks; there is no reference twin of the cross-GPU step (the reference keeps a sequence on
one device, decoding/worker/cache_engine.py:15-103).
"""
import ctypes
from dataclasses import dataclass
from typing import List, Optional, Sequence

import torch
import torch.distributed as dist

from . import _lib

PARTITION_SIZE = 512  # paged_attn.py:14


@dataclass(frozen=True)
class ContextSharding:
    """Host-side policy: which rank owns which logical block of a sequence.

    Logical block ``b`` belongs to rank ``(b // chunk_blocks) % world``; inside the owner it is
    local block ``(b // (chunk_blocks * world)) * chunk_blocks + b % chunk_blocks``.  Local order
    preserves global order, so the only partially filled page of a sequence is the last local page
    of its owner, which is what the kernels' tail masking expects."""
    world: int
    block_size: int
    chunk_blocks: int = 1

    def owner_of_block(self, b: int) -> int:
        return (b // self.chunk_blocks) % self.world

    def owner_of_token(self, pos: int) -> int:
        return self.owner_of_block(pos // self.block_size)

    def local_block(self, b: int) -> int:
        c = self.chunk_blocks
        return (b // (c * self.world)) * c + b % c

    def local_num_blocks(self, num_blocks: int, rank: int) -> int:
        c, w = self.chunk_blocks, self.world
        full_rounds, rest = divmod(num_blocks, c * w)
        return full_rounds * c + min(max(rest - rank * c, 0), c)

    def local_len(self, seq_len: int, rank: int) -> int:
        """Tokens of a ``seq_len``-token sequence stored on ``rank``."""
        bs = self.block_size
        nb = (seq_len + bs - 1) // bs
        n = self.local_num_blocks(nb, rank) * bs
        if nb and seq_len % bs and self.owner_of_block(nb - 1) == rank:
            n -= bs - seq_len % bs
        return n

    def local_slot(self, pos: int, local_block_table: Sequence[int]) -> int:
        """Slot (for ``slot_mapping``) of token ``pos`` inside its owner's pool."""
        lb = self.local_block(pos // self.block_size)
        return int(local_block_table[lb]) * self.block_size + pos % self.block_size

    def split_block_table(self, block_table: Sequence[int], seq_len: int) -> List[List[int]]:
        """Deal the logical blocks of one sequence to the ranks (per-rank lists in local order)."""
        nb = (seq_len + self.block_size - 1) // self.block_size
        out: List[List[int]] = [[] for _ in range(self.world)]
        for b in range(nb):
            out[self.owner_of_block(b)].append(int(block_table[b]))
        return out

    def shard_batch(self, block_tables: torch.Tensor, seq_lens: Sequence[int], rank: int):
        """(local block_tables int32 [B, maxb], local seq_lens int32 [B], max local len) of ``rank``
        for a batch described by global tables/lengths (CPU tensors)."""
        rows = [self.split_block_table(block_tables[i].tolist(), int(n))[rank]
                for i, n in enumerate(seq_lens)]
        maxb = max(1, max((len(r) for r in rows), default=1))
        bt = torch.zeros(len(rows), maxb, dtype=torch.int32)
        for i, r in enumerate(rows):
            if r:
                bt[i, :len(r)] = torch.tensor(r, dtype=torch.int32)
        lens = torch.tensor([self.local_len(int(n), rank) for n in seq_lens], dtype=torch.int32)
        return bt, lens, int(lens.max()) if len(rows) else 0


def _check(rc: int) -> None:
    if rc != 0:
        raise RuntimeError(_lib.last_error())


class ContextShardGroup:
    """One rank's end of the exchange: its own peer-visible buffer plus its view of every peer's.

    ``create`` (one process per GPU; handles travel over ``torch.distributed``) or ``create_local``
    (several virtual ranks inside one process on one device, each driven from its own CUDA stream
    -- how the single-GPU tests and ``tools/ctx_shard_bench.py --local`` exercise the protocol)."""

    def __init__(self, rank: int, world: int, max_pairs: int, head_size: int, device: int,
                 own_buf: int, bufs: List[int], opened: List[int], owns: bool):
        self.rank, self.world, self.max_pairs, self.head_size = rank, world, max_pairs, head_size
        self.device = device
        self._own = own_buf
        self._bufs = (ctypes.c_void_p * world)(*bufs)
        self._opened = opened
        self._owns = owns
        self.epoch = 0
        # True: the epoch lives in the exchange buffer (CUDA-graph capturable); all ranks must agree
        self.device_epoch = True
        self._scratch = {}   # grow-only per (B, H, D, dtype); must pre-exist any CUDA-graph capture
        # failure mirror: 4 pinned host words the exchange kernel itself writes when a wait times out
        # (zero-copy; no device-to-host copy on the per-step path)
        self._mirror = torch.zeros(4, dtype=torch.int32).pin_memory() if torch.cuda.is_available() \
            else torch.zeros(4, dtype=torch.int32)
        if torch.cuda.is_available() and own_buf:
            _check(_lib.lib.b200_ctx_exchange_set_host_mirror(own_buf, world, self._mirror.data_ptr(), device))

    # ---- construction ----
    @staticmethod
    def _alloc(world: int, max_pairs: int, head_size: int, device: int) -> int:
        nbytes = _lib.lib.b200_ctx_exchange_bytes(world, max_pairs, head_size)
        if nbytes <= 0:
            raise ValueError(f"bad exchange geometry world={world} max_pairs={max_pairs} head_size={head_size}")
        p = ctypes.c_void_p()
        _check(_lib.lib.b200_peer_alloc(ctypes.byref(p), nbytes, device))
        return int(p.value)

    @classmethod
    def create(cls, max_pairs: int, head_size: int, device: int,
               group: Optional[dist.ProcessGroup] = None) -> "ContextShardGroup":
        rank, world = dist.get_rank(group), dist.get_world_size(group)
        own = cls._alloc(world, max_pairs, head_size, device)
        handle = ctypes.create_string_buffer(64)
        _check(_lib.lib.b200_peer_export(own, handle, device))
        handles: List[Optional[bytes]] = [None] * world
        dist.all_gather_object(handles, handle.raw, group=group)
        bufs, opened = [], []
        for r in range(world):
            if r == rank:
                bufs.append(own)
                continue
            p = ctypes.c_void_p()
            _check(_lib.lib.b200_peer_open(handles[r], ctypes.byref(p), device))
            bufs.append(int(p.value))
            opened.append(int(p.value))
        dist.barrier(group=group)
        return cls(rank, world, max_pairs, head_size, device, own, bufs, opened, True)

    @classmethod
    def create_local(cls, world: int, max_pairs: int, head_size: int,
                     device: int) -> List["ContextShardGroup"]:
        bufs = [cls._alloc(world, max_pairs, head_size, device) for _ in range(world)]
        return [cls(r, world, max_pairs, head_size, device, bufs[r], bufs, [], True)
                for r in range(world)]

    def close(self) -> None:
        for p in self._opened:
            _lib.lib.b200_peer_close(p, self.device)
        self._opened = []
        if self._owns and self._own:
            _lib.lib.b200_peer_free(self._own, self.device)
            self._own = 0

    def status(self) -> int:
        """0, or 1 + the rank whose publication this rank timed out waiting for (synchronises)."""
        s = ctypes.c_int32(0)
        _check(_lib.lib.b200_ctx_exchange_status(self._own, self.world, ctypes.byref(s), self.device))
        return int(s.value)

    def set_timeout(self, milliseconds: int) -> None:
        """How long this rank waits for a peer's publication (default 2000 ms; first calls pay lazy
        module loads and IPC mapping).  Stored in this group's exchange buffer."""
        _check(_lib.lib.b200_ctx_exchange_set_timeout(self._own, self.world, int(milliseconds), self.device))

    def clear_status(self) -> None:
        _check(_lib.lib.b200_ctx_exchange_clear_status(self._own, self.world, self.device))
        self._mirror.zero_()

    def check(self, wait: bool = False) -> None:
        """Raise if a previous call timed out waiting for a peer (its output rows were NaN-poisoned
        by the kernel).  The kernel writes the failure into pinned host words itself, so this is a plain
        host-memory read: ``paged_attention`` calls it on entry and a failure of call n surfaces at the
        first call after that kernel has run.  ``wait=True`` synchronises the device first."""
        if wait:
            torch.cuda.synchronize(self.device)
        if int(self._mirror[0]) != 0:
            raise RuntimeError(
                f"context-sharded attention: rank {self.rank} timed out waiting for rank "
                f"{int(self._mirror[0]) - 1} to publish epoch {int(self._mirror[3])}; the outputs of "
                "that call are NaN on this rank (ContextShardGroup.clear_status() to re-arm)")

    # ---- the op ----
    def scratch(self, query: torch.Tensor, max_local_seq_len: int):
        B, H, D = query.shape
        P = max(1, (max_local_seq_len + PARTITION_SIZE - 1) // PARTITION_SIZE)
        key = (B, H, D, query.dtype)
        cur = self._scratch.get(key)
        if cur is None or cur[0].shape[2] < P:
            es = torch.empty(B, H, P, dtype=torch.float32, device=query.device)
            cur = (es, torch.empty_like(es), torch.empty(B, H, P, D, dtype=query.dtype, device=query.device))
            self._scratch[key] = cur
        return cur

    def paged_attention(self, out: torch.Tensor, query: torch.Tensor, key_cache: torch.Tensor,
                        value_cache: torch.Tensor, num_kv_heads: int, scale: float,
                        block_tables: torch.Tensor, local_seq_lens: torch.Tensor, block_size: int,
                        max_local_seq_len: int, kv_cache_dtype: str = "auto", k_scale: float = 1.0,
                        v_scale: float = 1.0, phases: int = 3) -> None:
        """Same argument meaning as ``ops.paged_attention_v2`` with the scratch managed here;
        block tables / lengths describe THIS rank's pages.  Collective: every rank of the group must
        call it, in the same order.  ``phases=1`` (partition pass + publish) followed later by
        ``phases=2`` (wait + merge, same arguments) is the split-phase form of the default 3."""
        assert query.is_cuda and out.is_contiguous() and query.stride(-1) == 1
        self.check()
        B, H, D = query.shape
        assert B * H <= self.max_pairs and D == self.head_size
        es, ml, tmp = self.scratch(query, max_local_seq_len)
        kv = _lib.lib.b200_parse_kv_cache_dtype(kv_cache_dtype.encode())
        if kv < 0:
            raise RuntimeError(f"Unsupported data type of kv cache: {kv_cache_dtype}")
        if phases != 2:
            self.epoch += 1  # calls issued (the kernels keep the protocol epoch on the device)
        f = ctypes.c_float
        rc = _lib.lib.b200_paged_attention_ctx_shard(
            out.data_ptr(), es.data_ptr(), ml.data_ptr(), tmp.data_ptr(), query.data_ptr(),
            key_cache.data_ptr(), value_cache.data_ptr(), block_tables.data_ptr(),
            local_seq_lens.data_ptr(), B, H, num_kv_heads, D, block_size, int(max_local_seq_len),
            block_tables.shape[1], es.shape[2], query.stride(0), key_cache.stride(0),
            key_cache.stride(1), f(scale), f(k_scale), f(v_scale), _lib._DT[query.dtype], kv,
            self._bufs, self.rank, self.world, self.max_pairs, 0 if self.device_epoch else (self.epoch & 0xFFFFFFFF), phases,
            query.device.index if query.device.index is not None else torch.cuda.current_device(),
            torch.cuda.current_stream(query.device).cuda_stream)
        _check(rc)
