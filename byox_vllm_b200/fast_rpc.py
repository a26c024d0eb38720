"""SURVEY §8f-2, RPC half: the per-step record rank 0 ships to the tensor-parallel workers.

The reference pickles `[method_name, list[Sequence], is_prefill]` into a 1 MiB shared-memory block
every step (engine/model_runner.py:44, :85-100). A `Sequence` has no compact state
(engine/sequence.py:14-77: every token id of the prompt and of the completion travels on every
decode step), so the pickle grows with batch x context, overflows the block at the BASELINE batch
sizes (1.5-6 MiB at configs 3/4) and costs milliseconds per step on both sides.

What the workers actually read is tiny: a decode step needs `last_token, num_tokens, block_table`
per sequence (model_runner.py:206-226), a prefill step `num_tokens, num_cached_tokens, block_table`
and the token ids that are not cached (:160-204). This module replaces the transport, not the
protocol (same shared-memory block name, same events, same `call()` dispatch):

  wire format   int32 header {magic, payload bytes, kind, batch, aux, 0, 0, 0} + 32 flag bytes
                (flag r = 1 once rank r has attached the block through THIS module), then
                kind 1 (decode)   last_token[B] | num_tokens[B] | table_len[B] | blocks[sum]  int32
                                  - exactly fast_meta.pack_decode_record, so a worker hands it to
                                  its DecodeMetaStager without building a single Python object
                kind 2 (prefill)  num_tokens[S] | num_cached[S] | table_len[S] | blocks[sum] |
                                  token ids of the uncached parts                       int32
                kind 0 (anything else: "exit", unusual arguments)  the reference's pickle
  block size    from the engine's own limits (max_num_seqs, max_model_len, block size,
                max_num_batched_tokens), not a constant
  ModelRunner.run   same body as model_runner.py:57-65 without its two debug f-strings: the second
                one formats `logits.shape`, and logits is None on every rank > 0, so the unmodified
                reference dies on its first step at any tensor-parallel size > 1.

  mixed installs    the block is NOT readable by the reference's _read_shm (and the reference's
                pickles are not readable here). A spawned rank that did not re-run install() - the
                launcher called it under `if __name__ == "__main__":` - would unpickle garbage and
                die while rank 0 hangs in NCCL. So every call starts with MAGIC (a reader that finds
                anything else says so), every bound worker raises its flag byte when it attaches the
                block, and rank 0's first _write_shm waits for all flags and otherwise raises with
                the instruction to call install() at module top level. (Falling back to the pickle
                transport would not help such a rank: the reference's own `run` cannot take a step
                on ranks > 0, see above.)

bind() / unbind() rebinding is done by install.install(fast_rpc=True) (default on). The transport
(_write_shm / _read_shm, the block size) only runs when world_size > 1; `ModelRunner.run` and the
`__init__` wrapper that records the engine limits are replaced at every tensor-parallel size, 1
included. Pure host code: no CUDA, no torch tensors on the wire.
"""
from __future__ import annotations

import pickle
import time
from itertools import chain

import numpy as np

KIND_PICKLE, KIND_DECODE, KIND_PREFILL = 0, 1, 2
MAGIC = 0x42323A02  # "B2:" + wire-format version 2; first word of every call
HEADER_INTS = 16  # magic, payload bytes, kind, batch, aux (prefill: uncached token ids), 3 spare,
HEADER_BYTES = HEADER_INTS * 4  # then bytes 32..63: one "attached through fast_rpc" flag per rank
FLAGS_AT, MAX_RANKS = 32, 32
ATTACH_TIMEOUT_S = 60.0


# ---- records ------------------------------------------------------------------------------------
def pack_decode(seqs, out: np.ndarray) -> int:
    """Write the decode record of `seqs` into the int32 array `out` (after the header); returns
    the payload length in int32 words. Layout = fast_meta.pack_decode_record."""
    B = len(seqs)
    lens = [len(s.block_table) for s in seqs]
    total = 3 * B + sum(lens)
    if total > out.size:
        raise OverflowError(f"decode record of {total} words exceeds the {out.size}-word block")
    out[:B] = [s.last_token for s in seqs]
    out[B:2 * B] = [s.num_tokens for s in seqs]
    out[2 * B:3 * B] = lens
    out[3 * B:total] = np.fromiter(chain.from_iterable(s.block_table for s in seqs),
                                   dtype=np.int32, count=total - 3 * B)
    return total


def pack_prefill(seqs, out: np.ndarray) -> tuple[int, int]:
    """Prefill record; returns (payload words, number of uncached token ids)."""
    S = len(seqs)
    lens = [len(s.block_table) for s in seqs]
    n_ids = sum(s.num_tokens - s.num_cached_tokens for s in seqs)
    total = 3 * S + sum(lens) + n_ids
    if total > out.size:
        raise OverflowError(f"prefill record of {total} words exceeds the {out.size}-word block")
    out[:S] = [s.num_tokens for s in seqs]
    out[S:2 * S] = [s.num_cached_tokens for s in seqs]
    out[2 * S:3 * S] = lens
    at = 3 * S + sum(lens)
    out[3 * S:at] = np.fromiter(chain.from_iterable(s.block_table for s in seqs), dtype=np.int32,
                                count=at - 3 * S)
    out[at:total] = np.fromiter(
        chain.from_iterable(s.token_ids[s.num_cached_tokens:s.num_tokens] for s in seqs),
        dtype=np.int32, count=n_ids)
    return total, n_ids


class SeqView:
    """What ModelRunner's metadata builders read of a Sequence (engine/sequence.py), rebuilt on a
    worker from a record: lengths, block table and - for prefill - the uncached token ids.
    `seq[seq.num_cached_tokens:]` is the only indexing the builders do (model_runner.py:172)."""

    __slots__ = ("last_token", "num_tokens", "num_cached_tokens", "block_table", "block_size",
                 "_tail", "temperature")

    def __init__(self, num_tokens, block_table, block_size, num_cached_tokens=0, last_token=0,
                 tail=None):
        self.num_tokens = int(num_tokens)
        self.block_table = block_table
        self.block_size = block_size
        self.num_cached_tokens = int(num_cached_tokens)
        self.last_token = int(last_token)
        self._tail = tail
        self.temperature = 1.0  # workers never sample (model_runner.py:60)

    def __len__(self):
        return self.num_tokens

    def __getitem__(self, key):
        if (isinstance(key, slice) and key.start == self.num_cached_tokens and key.stop is None
                and key.step is None and self._tail is not None):
            return self._tail
        raise IndexError("a worker-side sequence view only holds its uncached token ids")

    @property
    def num_cached_blocks(self):  # sequence.py:58-60
        return self.num_cached_tokens // self.block_size

    @property
    def num_blocks(self):  # sequence.py:62-64
        return (self.num_tokens + self.block_size - 1) // self.block_size

    @property
    def last_block_num_tokens(self):  # sequence.py:66-68
        return self.num_tokens - (self.num_blocks - 1) * self.block_size


def _tables(lens, blocks):
    out, at = [], 0
    for n in lens:
        out.append(blocks[at:at + n])
        at += n
    return out


def unpack_decode(rec: np.ndarray, batch: int, block_size: int) -> list[SeqView]:
    lens = rec[2 * batch:3 * batch].tolist()
    tables = _tables(lens, rec[3 * batch:3 * batch + sum(lens)].tolist())
    return [SeqView(n, t, block_size, last_token=tok)
            for tok, n, t in zip(rec[:batch].tolist(), rec[batch:2 * batch].tolist(), tables)]


def unpack_prefill(rec: np.ndarray, n_seqs: int, n_ids: int, block_size: int) -> list[SeqView]:
    S = n_seqs
    num_tokens, cached = rec[:S].tolist(), rec[S:2 * S].tolist()
    lens = rec[2 * S:3 * S].tolist()
    at = 3 * S + sum(lens)
    tables = _tables(lens, rec[3 * S:at].tolist())
    ids = rec[at:at + n_ids].tolist()
    out, p = [], 0
    for n, c, t in zip(num_tokens, cached, tables):
        tail = ids[p:p + n - c]
        p += n - c
        out.append(SeqView(n, t, block_size, num_cached_tokens=c,
                           last_token=tail[-1] if tail else 0, tail=tail))
    return out


# ---- the shared-memory block --------------------------------------------------------------------
def block_bytes(max_num_seqs: int, max_model_len: int, block_size: int,
                max_num_batched_tokens: int) -> int:
    """Size of the block from the engine's own limits: the largest decode record, the largest
    prefill record, and room for a pickled fallback call."""
    pages = (max_model_len + block_size - 1) // block_size
    decode = max_num_seqs * (3 + pages)
    prefill = max_num_seqs * (3 + pages) + max_num_batched_tokens
    return HEADER_BYTES + 4 * max(decode, prefill) + (1 << 16)


def write_call(buf, method_name: str, args: tuple) -> int:
    """Serialise one `call(method_name, *args)` into the block; returns the bytes used."""
    words = np.frombuffer(buf, dtype=np.int32)
    head, body = words[:HEADER_INTS], words[HEADER_INTS:]
    kind, batch, aux, n = KIND_PICKLE, 0, 0, 0
    if method_name == "run" and len(args) == 2 and isinstance(args[1], bool) and len(args[0]):
        seqs, is_prefill = args
        try:
            if is_prefill:
                n, aux = pack_prefill(seqs, body)
                kind = KIND_PREFILL
            else:
                n = pack_decode(seqs, body)
                kind = KIND_DECODE
            batch = len(seqs)
            n *= 4
        except (AttributeError, OverflowError):  # not Sequence-like / beyond the sized block
            kind = KIND_PICKLE
    if kind == KIND_PICKLE:
        data = pickle.dumps([method_name, *args])
        n = len(data)
        if HEADER_BYTES + n > len(buf):
            raise OverflowError(f"pickled call of {n} bytes exceeds the {len(buf)}-byte block")
        buf[HEADER_BYTES:HEADER_BYTES + n] = data
    head[:5] = (MAGIC, n, kind, batch, aux)  # bytes 32..63 are the workers' flags: not ours
    return HEADER_BYTES + n


def read_call(buf, block_size: int):
    """-> (method_name, args). Compact records come back as the worker-side method names
    `_b200_run_decode` / `_b200_run_prefill` with a private copy of the record."""
    words = np.frombuffer(buf, dtype=np.int32)
    magic, n, kind, batch, aux = words[:5].tolist()
    if magic != MAGIC:
        raise RuntimeError(
            f"nanovllm shared-memory block starts with {magic & 0xFFFFFFFF:#010x}, not the fast_rpc "
            f"magic {MAGIC:#010x}: rank 0 writes the reference's pickle format (or another "
            "fast_rpc version). Call byox_vllm_b200.install() at module top level of the "
            "launcher so that every spawned rank and rank 0 bind the same transport.")
    if kind == KIND_PICKLE:
        method_name, *args = pickle.loads(bytes(buf[HEADER_BYTES:HEADER_BYTES + n]))
        return method_name, args
    rec = words[HEADER_INTS:HEADER_INTS + n // 4].copy()  # rank 0 reuses the block next step
    if kind == KIND_DECODE:
        return "_b200_run_decode", [rec, batch]
    if kind == KIND_PREFILL:
        return "_b200_run_prefill", [rec, batch, aux]
    raise RuntimeError(f"unknown record kind {kind} in the nanovllm shared-memory block")


# ---- ModelRunner drop-ins (bound by install.install(fast_rpc=True)) -------------------------------
_ORIG: dict = {}
_STATS = {"steps": 0, "bytes": 0, "host_s": 0.0}


def mark_attached(buf, rank: int) -> None:
    """A worker that attached the block through this module says so (flag byte of its rank)."""
    if 0 < rank < MAX_RANKS:
        buf[FLAGS_AT + rank] = 1


def missing_ranks(buf, world_size: int) -> list[int]:
    return [r for r in range(1, min(world_size, MAX_RANKS)) if buf[FLAGS_AT + r] != 1]


def wait_for_workers(buf, world_size: int, timeout_s: float | None = None) -> None:
    """Rank 0, before its first call: every worker must have attached through fast_rpc. Workers
    attach right after the barrier rank 0 has just left (model_runner.py:43-49), so this returns
    at once in a consistent install; otherwise it raises instead of letting rank 0 hang in NCCL."""
    deadline = time.monotonic() + (ATTACH_TIMEOUT_S if timeout_s is None else timeout_s)
    while True:
        missing = missing_ranks(buf, world_size)
        if not missing:
            return
        if time.monotonic() >= deadline:
            raise RuntimeError(
                f"fast_rpc: tensor-parallel rank(s) {missing} did not attach the shared-memory "
                "block through byox_vllm_b200.fast_rpc: they run the reference's transport and "
                "cannot read this format. Call byox_vllm_b200.install() at module top level of the "
                "launcher script (not under `if __name__ == '__main__':`) so that spawned ranks "
                "re-apply it, or pass install(fast_rpc=False) on every rank.")
        time.sleep(0.002)


def _write_shm(self, method_name, *args):
    """ModelRunner._write_shm (model_runner.py:93-100) with the compact wire format."""
    assert self.world_size > 1 and self.rank == 0
    if not getattr(self, "_b200_rpc_workers_ok", False):
        wait_for_workers(self.shm.buf, self.world_size)
        self._b200_rpc_workers_ok = True
    t0 = time.perf_counter()
    used = write_call(self.shm.buf, method_name, args)
    _STATS["host_s"] += time.perf_counter() - t0
    _STATS["steps"] += 1
    _STATS["bytes"] += used
    for event in self.event:
        event.set()


def _read_shm(self):
    """ModelRunner._read_shm (model_runner.py:85-91)."""
    assert self.world_size > 1 and self.rank > 0
    self.event.wait()
    method_name, args = read_call(self.shm.buf, self.block_size)
    self.event.clear()
    return method_name, args


def _run(self, seqs, is_prefill):
    """ModelRunner.run (model_runner.py:57-65) without its two debug f-strings (:58, :63): the
    second formats `logits.shape` and logits is None on ranks > 0."""
    import sys

    mr = sys.modules[type(self).__module__]
    input_ids, positions = (self._prepare_prefill(seqs) if is_prefill
                            else self._prepare_decode(seqs))
    temperatures = self._prepare_sample(seqs) if self.rank == 0 else None
    logits = self._run_model(input_ids, positions, is_prefill)
    token_ids = self.sampler(logits, temperatures).tolist() if self.rank == 0 else None
    mr.reset_context()
    return token_ids


def _b200_run_decode(self, rec, batch):
    """Worker side of a decode step: the record goes straight to the device-side metadata builder
    when install(fast_meta=True) bound it (no Python object per sequence), else through light
    sequence views into the reference's own _prepare_decode."""
    import sys

    from . import fast_meta as fm

    mr = sys.modules[type(self).__module__]
    staged = None
    if getattr(type(self), "_prepare_decode", None) is fm.prepare_decode:
        staged = fm.prepare_decode_from_record(self, rec, batch)
    if staged is None:
        staged = self._prepare_decode(unpack_decode(rec, batch, self.block_size))
    input_ids, positions = staged
    self._run_model(input_ids, positions, False)
    mr.reset_context()
    return None


def _b200_run_prefill(self, rec, n_seqs, n_ids):
    import sys

    mr = sys.modules[type(self).__module__]
    seqs = unpack_prefill(rec, n_seqs, n_ids, self.block_size)
    input_ids, positions = self._prepare_prefill(seqs)
    self._run_model(input_ids, positions, True)
    mr.reset_context()
    return None


def _sized_shared_memory(mr_module):
    """SharedMemory subclass for model_runner.py:44: the creator sizes the block from the engine
    limits stashed by the wrapped ModelRunner.__init__ instead of the constant 2**20."""
    base = _ORIG["SharedMemory"]

    class SizedSharedMemory(base):
        def __init__(self, name=None, create=False, size=0):
            if create:
                size = max(size, mr_module._b200_rpc_block_bytes)
            super().__init__(name=name, create=create, size=size)
            if not create:  # a worker (model_runner.py:48): tell rank 0 this rank speaks fast_rpc
                mark_attached(self.buf, getattr(mr_module, "_b200_rpc_rank", 0))

    return SizedSharedMemory


def bind(mr_module) -> None:
    """Rebind the transport of nanovllm.engine.model_runner (module object of the reference)."""
    cls = mr_module.ModelRunner
    if "_write_shm" in _ORIG:
        return
    _ORIG.update(_write_shm=cls._write_shm, _read_shm=cls._read_shm, run=cls.run,
                 __init__=cls.__init__, SharedMemory=mr_module.SharedMemory)
    orig_init = cls.__init__

    def __init__(self, config, rank, event):
        mr_module._b200_rpc_block_bytes = block_bytes(
            config.max_num_seqs, config.max_model_len, config.kvcache_block_size,
            config.max_num_batched_tokens)
        mr_module._b200_rpc_rank = rank  # one ModelRunner per process (llm_engine.py:24-31)
        orig_init(self, config, rank, event)

    mr_module._b200_rpc_block_bytes = 1 << 20
    mr_module.SharedMemory = _sized_shared_memory(mr_module)
    cls.__init__ = __init__
    cls._write_shm = _write_shm
    cls._read_shm = _read_shm
    cls.run = _run
    cls._b200_run_decode = _b200_run_decode
    cls._b200_run_prefill = _b200_run_prefill


def unbind(mr_module) -> None:
    if "_write_shm" not in _ORIG:
        return
    cls = mr_module.ModelRunner
    cls._write_shm = _ORIG.pop("_write_shm")
    cls._read_shm = _ORIG.pop("_read_shm")
    cls.run = _ORIG.pop("run")
    cls.__init__ = _ORIG.pop("__init__")
    mr_module.SharedMemory = _ORIG.pop("SharedMemory")
    for name in ("_b200_run_decode", "_b200_run_prefill"):
        if hasattr(cls, name):
            delattr(cls, name)


def stats() -> dict:
    """Calls written by this process (rank 0), their total bytes and the host seconds spent
    serialising them: per-step RPC size and cost."""
    return dict(_STATS)
