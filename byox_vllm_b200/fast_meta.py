"""decode-step metadata built on the device.

The reference forms the decode metadata on the host every step — five Python lists, five pinned
tensors, five H2D copies (engine/model_runner.py:206-226, :152-158) — and then refreshes the
CUDA-graph input buffers with seven more small device ops (:242-254). Here the host packs ONE
compact int32 record

    last_token[B] | num_tokens[B] | table_len[B] | blocks[sum(table_len)]

into a pinned staging buffer, copies it with one cudaMemcpyAsync, and one launch of
b200attn_prepare_decode (csrc/prepare_meta.cu) writes input_ids / positions / slot_mapping /
context_lens / block_tables straight into the graph input buffers, bit-identically.

  pack_decode_record(...)      host packing (numpy), shared by the tests
  DecodeMetaStager             pinned double buffer + device record + the launch
  pack_prefill_record / PrefillMetaStager   the same for the per-token lists of a prefill step
                               (b200attn_prepare_prefill: positions, slot_mapping, cu_seqlens)
  prepare_decode / run_model / prepare_prefill   drop-ins for ModelRunner._prepare_decode /
                               _run_model / _prepare_prefill (install.install(fast_meta=True))
"""
from __future__ import annotations

from itertools import chain

import numpy as np
import torch

from . import _lib


def pack_decode_record(last_tokens, num_tokens, block_tables, out: np.ndarray | None = None):
    """-> (record int32[3B + sum(len)], table_width). `block_tables`: one list per sequence."""
    B = len(last_tokens)
    lens = [len(t) for t in block_tables]
    total = 3 * B + sum(lens)
    if out is None:
        out = np.empty(total, dtype=np.int32)
    assert out.dtype == np.int32 and out.size >= total
    out[:B] = last_tokens
    out[B:2 * B] = num_tokens
    out[2 * B:3 * B] = lens
    out[3 * B:total] = np.fromiter(chain.from_iterable(block_tables), dtype=np.int32,
                                   count=total - 3 * B)
    return out[:total], (max(lens) if lens else 0)


class DecodeMetaStager:
    """Owns the staging buffers for one set of graph input buffers (ModelRunner.graph_vars)."""

    def __init__(self, input_ids: torch.Tensor, positions: torch.Tensor,
                 slot_mapping: torch.Tensor, context_lens: torch.Tensor,
                 block_tables: torch.Tensor, block_size: int):
        for t in (input_ids, positions, slot_mapping, context_lens, block_tables):
            if not t.is_cuda:
                raise RuntimeError("DecodeMetaStager needs CUDA graph-input buffers; no CPU path")
        assert input_ids.dtype == torch.int64 and positions.dtype == torch.int64
        assert slot_mapping.dtype == torch.int32 and context_lens.dtype == torch.int32
        assert block_tables.dtype == torch.int32 and block_tables.dim() == 2
        assert block_tables.stride(1) == 1
        self.capacity = min(input_ids.numel(), positions.numel(), slot_mapping.numel(),
                            context_lens.numel(), block_tables.size(0))
        self.max_width = block_tables.size(1)
        self.block_size = block_size
        self.bufs = (input_ids, positions, slot_mapping, context_lens, block_tables)
        n = self.capacity * (3 + self.max_width)
        self.device = input_ids.device
        self._dev = torch.empty(n, dtype=torch.int32, device=self.device)
        # two pinned buffers: the copy of step t may still be in flight while step t+1 is packed
        self._pinned = [torch.empty(n, dtype=torch.int32, device="cpu", pin_memory=True)
                        for _ in range(2)]
        self._np = [p.numpy() for p in self._pinned]
        self._events = [torch.cuda.Event(), torch.cuda.Event()]
        self._used = [False, False]
        self._turn = 0

    def fits(self, batch: int, table_width: int) -> bool:
        return 0 < batch <= min(self.capacity, 1024) and table_width <= self.max_width

    def _next_pinned(self) -> int:
        i = self._turn
        self._turn ^= 1
        if self._used[i]:
            self._events[i].synchronize()
        return i

    def stage(self, last_tokens, num_tokens, block_tables) -> tuple[int, int]:
        """Pack, copy and expand. Returns (batch, table_width). Asynchronous on the current
        stream."""
        i = self._next_pinned()
        rec, width = pack_decode_record(last_tokens, num_tokens, block_tables, self._np[i])
        return self._launch(i, rec.size, len(last_tokens), width)

    def stage_record(self, rec: np.ndarray, batch: int) -> tuple[int, int]:
        """The same from an already packed record (fast_rpc: what a tensor-parallel worker gets
        from rank 0 IS this record, so no per-sequence Python object is ever built)."""
        width = int(rec[2 * batch:3 * batch].max()) if batch else 0
        i = self._next_pinned()
        self._np[i][:rec.size] = rec
        return self._launch(i, rec.size, batch, width)

    def _launch(self, i: int, n: int, B: int, width: int) -> tuple[int, int]:
        if not self.fits(B, width):
            raise ValueError(f"batch {B} x table width {width} exceeds the staged buffers")
        self._dev[:n].copy_(self._pinned[i][:n], non_blocking=True)
        self._events[i].record()
        self._used[i] = True
        input_ids, positions, slot_mapping, context_lens, block_tables_d = self.bufs
        lib = _lib.load()
        rc = lib.b200attn_prepare_decode(
            self._dev.data_ptr(), B, width, self.block_size, input_ids.data_ptr(),
            positions.data_ptr(), slot_mapping.data_ptr(), context_lens.data_ptr(),
            block_tables_d.data_ptr(), block_tables_d.stride(0), self.capacity,
            torch.cuda.current_stream(self.device).cuda_stream)
        _lib.check("b200attn_prepare_decode", rc)
        return B, width


def pack_prefill_record(num_tokens, num_cached, block_tables, out: np.ndarray | None = None):
    """-> record int32[3S + sum(len)] = num_tokens | num_cached | table_len | blocks."""
    S = len(num_tokens)
    lens = [len(t) for t in block_tables]
    total = 3 * S + sum(lens)
    if out is None:
        out = np.empty(total, dtype=np.int32)
    assert out.dtype == np.int32 and out.size >= total
    out[:S] = num_tokens
    out[S:2 * S] = num_cached
    out[2 * S:3 * S] = lens
    out[3 * S:total] = np.fromiter(chain.from_iterable(block_tables), dtype=np.int32,
                                   count=total - 3 * S)
    return out[:total]


class _PinnedRing:
    """Two growable pinned staging buffers with a reuse event each (the H2D copy of step t may
    still be in flight while step t+1 is being packed)."""

    def __init__(self, dtype):
        self.dtype = dtype
        self.bufs = [None, None]
        self.events = [None, None]
        self.turn = 0

    def next(self, n: int) -> torch.Tensor:
        i = self.turn
        self.turn ^= 1
        if self.events[i] is not None:
            self.events[i].synchronize()
        if self.bufs[i] is None or self.bufs[i].numel() < n:
            # explicit device: ModelRunner.__init__ warms up under torch.set_default_device("cuda")
            self.bufs[i] = torch.empty(max(n, 4096), dtype=self.dtype, device="cpu",
                                       pin_memory=True)
        self._last = i
        return self.bufs[i]

    def mark(self) -> None:
        if self.events[self._last] is None:
            self.events[self._last] = torch.cuda.Event()
        self.events[self._last].record()


class PrefillMetaStager:
    """ModelRunner._prepare_prefill's per-token lists (positions, slot_mapping, cu_seqlens) built
    by b200attn_prepare_prefill from one compact record. Output tensors are fresh per step, as in
    the reference (prefill is not graph-captured)."""

    def __init__(self, device: torch.device, block_size: int):
        if torch.device(device).type != "cuda":
            raise RuntimeError("PrefillMetaStager needs a CUDA device; there is no CPU path")
        self.device = torch.device(device)
        self.block_size = block_size
        self._rec = _PinnedRing(torch.int32)

    def stage(self, num_tokens, num_cached, block_tables):
        """-> (positions int64[Nq], slot_mapping int32[Ns], cu_seqlens_q, cu_seqlens_k) on the
        device, asynchronous on the current stream."""
        S, bs = len(num_tokens), self.block_size
        if S > 1024:
            raise ValueError(f"{S} sequences exceed the 1024 the device builder handles")
        total_q = sum(num_tokens) - sum(num_cached)
        total_slots = sum(n - (c // bs) * bs
                          for n, c, t in zip(num_tokens, num_cached, block_tables) if t)
        n_rec = 3 * S + sum(map(len, block_tables))
        pinned = self._rec.next(n_rec)
        pack_prefill_record(num_tokens, num_cached, block_tables, pinned.numpy())
        dev = self.device
        rec = torch.empty(n_rec, dtype=torch.int32, device=dev)
        rec.copy_(pinned[:n_rec], non_blocking=True)
        self._rec.mark()
        positions = torch.empty(total_q, dtype=torch.int64, device=dev)
        slot_mapping = torch.empty(total_slots, dtype=torch.int32, device=dev)
        cu_q = torch.empty(S + 1, dtype=torch.int32, device=dev)
        cu_k = torch.empty(S + 1, dtype=torch.int32, device=dev)
        rc = _lib.load().b200attn_prepare_prefill(
            rec.data_ptr(), S, bs, positions.data_ptr(), total_q, slot_mapping.data_ptr(),
            total_slots, cu_q.data_ptr(), cu_k.data_ptr(),
            torch.cuda.current_stream(dev).cuda_stream)
        _lib.check("b200attn_prepare_prefill", rc)
        return positions, slot_mapping, cu_q, cu_k


# ---- drop-ins for the reference's ModelRunner (bound by install.py) ----------------------------
_ORIG = {}


def _stager_of(runner):
    st = getattr(runner, "_b200_meta_stager", None)
    if st is None and getattr(runner, "graph_vars", None) is not None:
        gv = runner.graph_vars
        st = DecodeMetaStager(gv["input_ids"], gv["positions"], gv["slot_mapping"],
                              gv["context_lens"], gv["block_tables"], runner.block_size)
        runner._b200_meta_stager = st
    return st


def prepare_decode(self, seqs):
    """ModelRunner._prepare_decode (model_runner.py:206-226) writing the graph inputs directly."""
    import sys

    from . import attention as _attn

    # the Context module install() bound (nanovllm.utils.context), else our standalone mirror
    set_context = sys.modules[_attn.get_context.__module__].set_context
    st = None if self.enforce_eager else _stager_of(self)
    bs = len(seqs)
    tables = [seq.block_table for seq in seqs]
    if st is None or not st.fits(bs, max(map(len, tables))):
        return _ORIG["_prepare_decode"](self, seqs)
    _, width = st.stage([seq.last_token for seq in seqs], [seq.num_tokens for seq in seqs], tables)
    gv = self.graph_vars
    set_context(False, slot_mapping=gv["slot_mapping"][:bs], context_lens=gv["context_lens"][:bs],
                block_tables=gv["block_tables"][:bs, :width])
    self._b200_meta_staged = bs
    return gv["input_ids"][:bs], gv["positions"][:bs]


def prepare_decode_from_record(self, rec: np.ndarray, batch: int):
    """prepare_decode for a tensor-parallel worker that received the packed decode record of
    fast_rpc instead of Sequence objects. Returns (input_ids, positions) or None when the staged
    buffers cannot take the batch (the caller then rebuilds sequence views)."""
    import sys

    from . import attention as _attn

    set_context = sys.modules[_attn.get_context.__module__].set_context
    st = None if self.enforce_eager else _stager_of(self)
    if st is None or batch <= 0:
        return None
    width = int(rec[2 * batch:3 * batch].max())
    if not st.fits(batch, width):
        return None
    st.stage_record(rec, batch)
    gv = self.graph_vars
    set_context(False, slot_mapping=gv["slot_mapping"][:batch],
                context_lens=gv["context_lens"][:batch],
                block_tables=gv["block_tables"][:batch, :width])
    self._b200_meta_staged = batch
    return gv["input_ids"][:batch], gv["positions"][:batch]


def prepare_prefill(self, seqs):
    """ModelRunner._prepare_prefill (model_runner.py:160-204): the per-token lists (positions,
    slot_mapping) and the cu_seqlens come from one device launch; input_ids (the token ids
    themselves), max_seqlen_q/k and the prefix-cache block tables are formed as the reference
    forms them."""
    import sys

    from . import attention as _attn

    set_context = sys.modules[_attn.get_context.__module__].set_context
    self._b200_meta_staged = 0
    if len(seqs) > 1024:
        return _ORIG["_prepare_prefill"](self, seqs)
    st = getattr(self, "_b200_prefill_stager", None)
    if st is None:
        st = PrefillMetaStager(torch.device("cuda", torch.cuda.current_device()), self.block_size)
        self._b200_prefill_stager = st
        self._b200_ids_ring = _PinnedRing(torch.int64)
    num_tokens = [len(seq) for seq in seqs]
    num_cached = [seq.num_cached_tokens for seq in seqs]
    positions, slot_mapping, cu_q, cu_k = st.stage(num_tokens, num_cached,
                                                   [seq.block_table for seq in seqs])
    total_q = positions.numel()
    ring = self._b200_ids_ring
    pinned = ring.next(total_q)
    pinned.numpy()[:total_q] = np.fromiter(
        chain.from_iterable(seq[seq.num_cached_tokens:] for seq in seqs), dtype=np.int64,
        count=total_q)
    input_ids = torch.empty(total_q, dtype=torch.int64, device=positions.device)
    input_ids.copy_(pinned[:total_q], non_blocking=True)
    ring.mark()
    max_q = max(n - c for n, c in zip(num_tokens, num_cached))
    max_k = max(num_tokens)
    block_tables = None
    if sum(num_tokens) > total_q:  # prefix cache (model_runner.py:188-189)
        block_tables = self._prepare_block_tables(seqs)
    set_context(True, cu_q, cu_k, max_q, max_k, slot_mapping, None, block_tables)
    return input_ids, positions


@torch.inference_mode()
def run_model(self, input_ids, positions, is_prefill):
    """ModelRunner._run_model (model_runner.py:228-254): when prepare_decode above already filled
    the graph inputs, replay the graph without the seven refresh copies."""
    staged = getattr(self, "_b200_meta_staged", 0)
    self._b200_meta_staged = 0
    if is_prefill or not staged:
        return _ORIG["_run_model"](self, input_ids, positions, is_prefill)
    if staged != input_ids.size(0):
        # prepare_decode handed out views of the graph inputs: the reference's _run_model would
        # fill_(-1) / zero_() those very buffers before copying them onto themselves
        raise RuntimeError(f"decode metadata was staged for {staged} rows but _run_model got "
                           f"{input_ids.size(0)}: call _prepare_decode again for this batch")
    bs = staged
    graph = self.graphs[next(x for x in self.graph_bs if x >= bs)]
    graph.replay()
    return self.model.compute_logits(self.graph_vars["outputs"][:bs])


def bind(model_runner_cls) -> None:
    if "_prepare_decode" not in _ORIG:
        _ORIG["_prepare_decode"] = model_runner_cls._prepare_decode
        _ORIG["_run_model"] = model_runner_cls._run_model
        _ORIG["_prepare_prefill"] = model_runner_cls._prepare_prefill
    model_runner_cls._prepare_decode = prepare_decode
    model_runner_cls._run_model = run_model
    model_runner_cls._prepare_prefill = prepare_prefill


def unbind(model_runner_cls) -> None:
    if "_prepare_decode" in _ORIG:
        model_runner_cls._prepare_decode = _ORIG.pop("_prepare_decode")
        model_runner_cls._run_model = _ORIG.pop("_run_model")
        model_runner_cls._prepare_prefill = _ORIG.pop("_prepare_prefill")
