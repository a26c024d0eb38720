"""device-side decode metadata. CPU part: the host packing; GPU part: the kernel
against the restated builders (tools/metadata_ref.py, pinned by the reference's own expectations
in tests/test_prepare_funcs.py:83-143,169-206 — see tests/test_metadata.py) and against the
reference's OWN ModelRunner._prepare_decode / _prepare_prefill (baseline/_ref, through
ModelRunner.__new__) — bit-exact."""
import random

import numpy as np
import pytest
import torch

from byox_vllm_b200 import fast_meta as fm
from tools import metadata_ref as md


def test_pack_decode_record_layout():
    rec, width = fm.pack_decode_record([7, 9, 11], [5, 300, 17], [[3], [8, 2], [4]])
    assert rec.dtype == np.int32 and width == 2
    assert rec.tolist() == [7, 9, 11, 5, 300, 17, 1, 2, 1, 3, 8, 2, 4]
    buf = np.full(64, -7, dtype=np.int32)
    rec2, _ = fm.pack_decode_record([7, 9, 11], [5, 300, 17], [[3], [8, 2], [4]], buf)
    assert rec2.tolist() == rec.tolist() and rec2.base is buf and buf[13] == -7
    rec, width = fm.pack_decode_record([], [], [])
    assert rec.size == 0 and width == 0


def _run(dev, last, ntok, tables, block_size, capacity, max_width, stale=12345):
    ids = torch.full((capacity,), stale, dtype=torch.int64, device=dev)
    pos = torch.full((capacity,), stale, dtype=torch.int64, device=dev)
    slot = torch.full((capacity,), stale, dtype=torch.int32, device=dev)
    lens = torch.full((capacity,), stale, dtype=torch.int32, device=dev)
    bt = torch.full((capacity, max_width), stale, dtype=torch.int32, device=dev)
    st = fm.DecodeMetaStager(ids, pos, slot, lens, bt, block_size)
    B, W = st.stage(last, ntok, tables)
    torch.cuda.synchronize()
    return ids.cpu(), pos.cpu(), slot.cpu(), lens.cpu(), bt.cpu(), B, W, st


def _check(dev, last, ntok, tables, block_size, capacity, max_width):
    stale = 12345
    ids, pos, slot, lens, bt, B, W, _ = _run(dev, last, ntok, tables, block_size, capacity,
                                             max_width, stale)
    want = md.prepare_decode([md.SeqState(n, list(t)) for n, t in zip(ntok, tables)], block_size)
    assert W == want["block_tables"].size(1)
    assert torch.equal(ids[:B], torch.tensor(last, dtype=torch.int64))
    assert torch.equal(pos[:B], want["positions"])
    assert torch.equal(slot[:B], want["slot_mapping"])
    assert torch.equal(lens[:B], want["context_lens"])
    assert torch.equal(bt[:B, :W], want["block_tables"])
    # what _run_model's fill_(-1) / zero_() leave in the padding rows (model_runner.py:248-251)
    assert bool((slot[B:] == -1).all()) and bool((lens[B:] == 0).all())
    # untouched: ids / positions of padding rows, table columns >= W and padding table rows
    assert bool((ids[B:] == stale).all()) and bool((pos[B:] == stale).all())
    assert bool((bt[:, W:] == stale).all()) and bool((bt[B:] == stale).all())


@pytest.mark.gpu
def test_reference_expectations_block16(cuda_dev):
    """The cases of the reference's tests/test_prepare_funcs.py (block_size 16; its mocks leave -1
    as the LAST table entry, which makes the slot negative: block_table[-1] is what counts)."""
    _check(cuda_dev, [5], [5], [[0, -1, -1, -1]], 16, 8, 4)            # :83-101  slot -12
    _check(cuda_dev, [3, 7], [3, 4], [[0, -1, -1, -1], [1, -1, -1, -1]], 16, 8, 6)  # :104-143
    _check(cuda_dev, [4, 9], [4, 5], [[0, -1, -1, -1], [1, -1, -1, -1]], 16, 2, 4)  # :169-206


@pytest.mark.gpu
@pytest.mark.parametrize("batch,block_size,max_len,capacity", [
    (1, 256, 4096, 512), (3, 256, 1000, 16), (256, 256, 4096, 512), (512, 256, 4096, 512),
    (511, 64, 40960, 512), (1000, 256, 2048, 1024)])
def test_prepare_decode_matches_reference_builders(cuda_dev, batch, block_size, max_len, capacity):
    rng = random.Random(batch)
    ntok = [rng.randint(1, max_len) for _ in range(batch)]
    # page-boundary rows: exactly full page, first token of a new page
    for i, n in enumerate((block_size, block_size + 1, 1, max_len)):
        if i < batch:
            ntok[i] = n
    tables = [[rng.randrange(1 << 22) for _ in range((n + block_size - 1) // block_size)]
              for n in ntok]
    last = [rng.randint(0, 151935) for _ in range(batch)]
    _check(cuda_dev, last, ntok, tables, block_size, capacity,
           (max_len + block_size - 1) // block_size + 2)


@pytest.mark.gpu
def test_stager_reuse_and_model_runner_dropins(cuda_dev):
    """Two consecutive steps through the same stager (pinned double buffer), then the ModelRunner
    drop-ins on a stand-in runner: the Context carries slices of the graph inputs, the graph is
    replayed without the refresh copies, and an oversized batch falls back to the original."""
    from byox_vllm_b200 import attention as attn_mod
    from byox_vllm_b200.context import reset_context

    dev = cuda_dev
    cap, width = 16, 8
    gv = dict(input_ids=torch.zeros(cap, dtype=torch.int64, device=dev),
              positions=torch.zeros(cap, dtype=torch.int64, device=dev),
              slot_mapping=torch.zeros(cap, dtype=torch.int32, device=dev),
              context_lens=torch.zeros(cap, dtype=torch.int32, device=dev),
              block_tables=torch.zeros(cap, width, dtype=torch.int32, device=dev),
              outputs=torch.arange(cap, dtype=torch.float32, device=dev).view(cap, 1))

    class Seq:
        def __init__(self, n, table, tok):
            self.num_tokens, self.block_table, self.last_token = n, table, tok

    class Graph:
        replays = 0

        def replay(self):
            Graph.replays += 1

    class Runner:
        enforce_eager = False
        block_size = 256
        graph_vars = gv
        graph_bs = [1, 2, 4, 8, 16]
        graphs = {b: Graph() for b in graph_bs}
        model = type("M", (), {"compute_logits": staticmethod(lambda x: x)})()

    calls = []
    fm._ORIG["_prepare_decode"] = lambda self, seqs: calls.append("prepare") or ("ids", "pos")
    fm._ORIG["_run_model"] = lambda self, i, p, pre: calls.append("run") or "orig"
    try:
        r = Runner()
        for step in range(3):
            seqs = [Seq(300 + step, [5, 9], 11 + step), Seq(256 + step, [7] + [3] * (step > 0), 4),
                    Seq(17, [2], 6)]
            ids, pos = fm.prepare_decode(r, seqs)
            ctx = attn_mod.get_context()
            torch.cuda.synchronize()
            want = md.prepare_decode([md.SeqState(s.num_tokens, s.block_table) for s in seqs], 256)
            assert ids.data_ptr() == gv["input_ids"].data_ptr() and ids.size(0) == 3
            assert ids.tolist() == [s.last_token for s in seqs]
            assert torch.equal(pos.cpu(), want["positions"])
            assert not ctx.is_prefill
            assert ctx.slot_mapping.data_ptr() == gv["slot_mapping"].data_ptr()
            assert torch.equal(ctx.slot_mapping.cpu(), want["slot_mapping"])
            assert torch.equal(ctx.context_lens.cpu(), want["context_lens"])
            assert torch.equal(ctx.block_tables.cpu(), want["block_tables"])
            assert gv["slot_mapping"][3:].tolist() == [-1] * (cap - 3)
            out = fm.run_model(r, ids, pos, False)
            assert Graph.replays == step + 1 and out.shape == (3, 1) and calls == []
        # prefill and unstaged calls go to the original
        assert fm.run_model(r, ids, pos, True) == "orig"
        assert fm.run_model(r, ids, pos, False) == "orig"  # nothing staged any more
        big = [Seq(5, [1], 2) for _ in range(cap + 1)]
        assert fm.prepare_decode(r, big) == ("ids", "pos")
        assert calls == ["run", "run", "prepare"]
    finally:
        fm._ORIG.clear()
        reset_context()


def test_pack_prefill_record_layout():
    rec = fm.pack_prefill_record([5, 300], [0, 256], [[3], [8, 2]])
    assert rec.dtype == np.int32 and rec.tolist() == [5, 300, 0, 256, 1, 2, 3, 8, 2]
    assert fm.pack_prefill_record([], [], []).size == 0


def _check_prefill(dev, seqs, block_size):
    st = fm.PrefillMetaStager(dev, block_size)
    want = md.prepare_prefill(seqs, block_size)
    for _ in range(3):  # the pinned ring is reused across steps
        pos, slot, cu_q, cu_k = st.stage([s.num_tokens for s in seqs],
                                         [s.num_cached_tokens for s in seqs],
                                         [s.block_table for s in seqs])
        torch.cuda.synchronize()
        assert pos.dtype == torch.int64 and slot.dtype == torch.int32
        assert torch.equal(pos.cpu(), want["positions"])
        assert torch.equal(slot.cpu(), want["slot_mapping"])
        assert torch.equal(cu_q.cpu(), want["cu_seqlens_q"])
        assert torch.equal(cu_k.cpu(), want["cu_seqlens_k"])


@pytest.mark.gpu
def test_prepare_prefill_reference_expectations_block16(cuda_dev):
    """The prefill cases of the reference's tests/test_prepare_funcs.py:17-80,146-166 (block 16),
    including a cached prefix that is not a whole block (slots cover the block, positions do
    not)."""
    S = lambda n, t, c=0: md.SeqState(n, list(t), c)
    _check_prefill(cuda_dev, [S(5, [0])], 16)
    _check_prefill(cuda_dev, [S(3, [0]), S(5, [1])], 16)
    _check_prefill(cuda_dev, [S(6, [0, 1], 3)], 16)
    _check_prefill(cuda_dev, [S(5, [0], 2), S(6, [1], 1)], 16)


@pytest.mark.gpu
def test_prepare_prefill_block256_batches(cuda_dev):
    S = lambda n, t, c=0: md.SeqState(n, list(t), c)
    _check_prefill(cuda_dev, [S(600, [4, 6, 0], 512), S(300, [1, 2])], 256)
    _check_prefill(cuda_dev, [S(4096, []), S(4096, [])], 256)  # warm-up: no tables, no slots
    _check_prefill(cuda_dev, [S(256, [9]), S(257, [3, 5]), S(1, [7]), S(513, [2, 4, 6], 512)], 256)
    rng = random.Random(3)
    seqs = []
    for _ in range(48):  # ~16k uncached tokens, some sequences with cached prefixes
        n = rng.randint(1, 700)
        cached = rng.choice([0, 0, 256, 512]) if n > 512 else (256 if n > 300 and rng.random() < .5
                                                                else 0)
        nb = (n + 255) // 256
        seqs.append(S(n, [rng.randrange(1 << 20) for _ in range(nb)], cached))
    _check_prefill(cuda_dev, seqs, 256)
    _check_prefill(cuda_dev, [S(16384, [rng.randrange(1 << 20) for _ in range(64)])], 256)
    many = [S(rng.randint(1, 40), [rng.randrange(1 << 20)]) for _ in range(1024)]
    _check_prefill(cuda_dev, many, 256)


# ---- against the reference's OWN builders (baseline/_ref, unmodified) --------------------------------
def _reference_runner():
    """ModelRunner.__new__ (no __init__: no process group, no model): the reference's
    _prepare_decode / _prepare_prefill need only `block_size` (model_runner.py:152-226)."""
    import sys
    from pathlib import Path

    ref = Path(__file__).resolve().parents[1] / "baseline" / "_ref"
    if not (ref / "nanovllm").is_dir():
        pytest.skip("baseline/_ref (the installed reference) is not present")
    if str(ref) not in sys.path:
        sys.path.insert(0, str(ref))
    from byox_vllm_b200 import install

    # only the attention module is swapped (so that flash_attn / triton are not needed to import
    # the engine); the reference's builders stay as they are
    install.install(fuse_qkv=False, fast_meta=False, fast_rpc=False)
    from nanovllm.engine.model_runner import ModelRunner
    from nanovllm.engine.sequence import Sequence
    from nanovllm.utils.context import get_context, reset_context

    return ModelRunner, Sequence, get_context, reset_context, install


def _ref_seq(Sequence, rng, n, block_size, cached_blocks=0):
    s = Sequence([rng.randrange(30000) for _ in range(n)], block_size)
    s.block_table = [rng.randrange(1 << 20) for _ in range(s.num_blocks)]
    s.num_cached_tokens = cached_blocks * block_size
    return s


@pytest.mark.gpu
@pytest.mark.parametrize("block_size", [16, 256])
def test_device_builders_match_reference_model_runner(cuda_dev, block_size):
    """The decode and prefill metadata built on the device (fast_meta stagers, one copy + one
    launch) against what the reference's own ModelRunner._prepare_decode / _prepare_prefill put
    into the Context for the same Sequence objects: bit-exact, field by field."""
    ModelRunner, Sequence, get_context, reset_context, install = _reference_runner()
    try:
        runner = ModelRunner.__new__(ModelRunner)
        runner.block_size = block_size
        rng = random.Random(block_size)
        for B in (1, 3, 17, 64):
            seqs = [_ref_seq(Sequence, rng, rng.randint(1, 40 * block_size), block_size)
                    for _ in range(B)]
            ids_r, pos_r = ModelRunner._prepare_decode(runner, seqs)
            ctx = get_context()
            W = ctx.block_tables.size(1)
            cap = B + 5
            gv = [torch.full((cap,), 99, dtype=torch.int64, device=cuda_dev),
                  torch.full((cap,), 99, dtype=torch.int64, device=cuda_dev),
                  torch.full((cap,), 99, dtype=torch.int32, device=cuda_dev),
                  torch.full((cap,), 99, dtype=torch.int32, device=cuda_dev),
                  torch.full((cap, W + 3), 99, dtype=torch.int32, device=cuda_dev)]
            st = fm.DecodeMetaStager(*gv, block_size)
            b, w = st.stage([s.last_token for s in seqs], [s.num_tokens for s in seqs],
                            [s.block_table for s in seqs])
            torch.cuda.synchronize()
            assert (b, w) == (B, W)
            assert torch.equal(gv[0][:B], ids_r) and torch.equal(gv[1][:B], pos_r)
            assert torch.equal(gv[2][:B], ctx.slot_mapping)
            assert torch.equal(gv[3][:B], ctx.context_lens)
            assert torch.equal(gv[4][:B, :W], ctx.block_tables)
            reset_context()
        for B in (1, 4, 23):
            seqs = []
            for _ in range(B):
                n = rng.randint(1, 12 * block_size)
                seqs.append(_ref_seq(Sequence, rng, n, block_size,
                                     cached_blocks=rng.randint(0, (n - 1) // block_size)))
            ids_r, pos_r = ModelRunner._prepare_prefill(runner, seqs)
            ctx = get_context()
            pst = fm.PrefillMetaStager(cuda_dev, block_size)
            pos, slot, cu_q, cu_k = pst.stage([s.num_tokens for s in seqs],
                                              [s.num_cached_tokens for s in seqs],
                                              [s.block_table for s in seqs])
            torch.cuda.synchronize()
            assert torch.equal(pos, pos_r)
            assert torch.equal(slot, ctx.slot_mapping)
            assert torch.equal(cu_q, ctx.cu_seqlens_q) and torch.equal(cu_k, ctx.cu_seqlens_k)
            assert ids_r.tolist() == [t for s in seqs for t in s[s.num_cached_tokens:]]
            reset_context()
    finally:
        install.uninstall()


@pytest.mark.gpu
def test_bound_drop_ins_on_reference_model_runner(cuda_dev):
    """install(fast_meta=True) rebinds ModelRunner._prepare_decode / _run_model; on a runner with
    graph buffers (ModelRunner.__new__ + the attributes _capture_cudagraph would set) a decode step
    fills the graph inputs with exactly what the reference's own builder + refresh copies
    (model_runner.py:206-254) would have put there, and replays the graph."""
    ModelRunner, Sequence, get_context, reset_context, install = _reference_runner()
    install.uninstall()
    install.install(fuse_qkv=False, fast_meta=True, fast_rpc=False)
    try:
        from nanovllm.engine.model_runner import ModelRunner as MR

        assert MR._prepare_decode is fm.prepare_decode and MR._run_model is fm.run_model
        orig_prepare = fm._ORIG["_prepare_decode"]
        cap, width, bs = 16, 12, 256
        dev = cuda_dev
        gv = dict(input_ids=torch.zeros(cap, dtype=torch.int64, device=dev),
                  positions=torch.zeros(cap, dtype=torch.int64, device=dev),
                  slot_mapping=torch.zeros(cap, dtype=torch.int32, device=dev),
                  context_lens=torch.zeros(cap, dtype=torch.int32, device=dev),
                  block_tables=torch.zeros(cap, width, dtype=torch.int32, device=dev),
                  outputs=torch.arange(cap, dtype=torch.float32, device=dev).view(cap, 1))
        replays = []

        class Graph:
            def replay(self):
                replays.append({k: v.clone() for k, v in gv.items()})

        runner = MR.__new__(MR)
        runner.block_size, runner.enforce_eager, runner.rank, runner.world_size = bs, False, 0, 1
        runner.graph_vars, runner.graph_bs = gv, [1, 2, 4, 8, 16]
        runner.graphs = {b: Graph() for b in runner.graph_bs}
        runner.model = type("M", (), {"compute_logits": staticmethod(lambda x: x)})()
        rng = random.Random(9)
        for B in (3, 8, 1, 11):
            seqs = [_ref_seq(Sequence, rng, rng.randint(1, 10 * bs), bs) for _ in range(B)]
            ids, pos = MR._prepare_decode(runner, seqs)
            out = MR._run_model(runner, ids, pos, False)
            torch.cuda.synchronize()
            assert out.shape == (B, 1)
            seen = replays[-1]
            # the reference's own path for the same sequences
            ids_r, pos_r = orig_prepare(runner, seqs)
            ctx = get_context()
            W = ctx.block_tables.size(1)
            assert torch.equal(seen["input_ids"][:B], ids_r)
            assert torch.equal(seen["positions"][:B], pos_r)
            assert torch.equal(seen["slot_mapping"][:B], ctx.slot_mapping)
            assert bool((seen["slot_mapping"][B:] == -1).all())
            assert torch.equal(seen["context_lens"][:B], ctx.context_lens)
            assert bool((seen["context_lens"][B:] == 0).all())
            assert torch.equal(seen["block_tables"][:B, :W], ctx.block_tables)
            reset_context()
        assert len(replays) == 4
    finally:
        install.uninstall()
