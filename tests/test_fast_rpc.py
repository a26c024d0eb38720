"""SURVEY §8f-2, RPC half (byox_vllm_b200/fast_rpc.py): the compact per-step record that replaces
the reference's pickled `[method, list[Sequence], is_prefill]` (engine/model_runner.py:85-100).
CPU only: records <-> sequence views, the wire format through a real shared-memory block between
two processes with the reference's event protocol, block sizing at the BASELINE configurations,
and the fallbacks."""
import multiprocessing as mp
import pickle
import random
import sys
from multiprocessing.shared_memory import SharedMemory
from pathlib import Path

import numpy as np
import pytest

from byox_vllm_b200 import fast_rpc as rpc

BLOCK = 256


class Seq:
    """The fields of the reference's Sequence (engine/sequence.py:14-77) that travel or are read
    by the metadata builders."""

    def __init__(self, token_ids, block_table, num_cached_tokens=0, block_size=BLOCK):
        self.token_ids = list(token_ids)
        self.last_token = self.token_ids[-1]
        self.num_tokens = len(self.token_ids)
        self.num_cached_tokens = num_cached_tokens
        self.block_table = list(block_table)
        self.block_size = block_size
        self.temperature = 0.6

    def __len__(self):
        return self.num_tokens

    def __getitem__(self, key):
        return self.token_ids[key]

    @property
    def num_blocks(self):
        return (self.num_tokens + self.block_size - 1) // self.block_size

    @property
    def num_cached_blocks(self):
        return self.num_cached_tokens // self.block_size

    @property
    def last_block_num_tokens(self):
        return self.num_tokens - (self.num_blocks - 1) * self.block_size


def make_seqs(rng, n, max_len, cached=False, cls=Seq):
    out = []
    for _ in range(n):
        L = rng.randint(1, max_len)
        nb = (L + BLOCK - 1) // BLOCK
        s = cls([rng.randrange(150000) for _ in range(L)], [rng.randrange(1 << 20) for _ in range(nb)])
        if cached and L > BLOCK:
            s.num_cached_tokens = rng.randint(0, (L - 1) // BLOCK) * BLOCK
        out.append(s)
    return out


def same_view(view, seq, prefill):
    assert len(view) == len(seq) == view.num_tokens
    assert view.block_table == seq.block_table
    assert view.num_blocks == seq.num_blocks
    assert view.last_block_num_tokens == seq.last_block_num_tokens
    assert view.num_cached_tokens == (seq.num_cached_tokens if prefill else 0)
    if prefill:
        assert view.num_cached_blocks == seq.num_cached_blocks
        assert view[view.num_cached_tokens:] == seq[seq.num_cached_tokens:]
        with pytest.raises(IndexError):
            view[0:3]  # a worker never holds the cached part of a prompt
    else:
        assert view.last_token == seq.last_token


def test_records_round_trip():
    rng = random.Random(0)
    buf = np.zeros(1 << 20, dtype=np.int32)
    for B in (1, 2, 17, 256):
        seqs = make_seqs(rng, B, 5000)
        n = rpc.pack_decode(seqs, buf)
        assert n == 3 * B + sum(len(s.block_table) for s in seqs)
        for v, s in zip(rpc.unpack_decode(buf[:n].copy(), B, BLOCK), seqs):
            same_view(v, s, prefill=False)
    for B in (1, 5, 40):
        seqs = make_seqs(rng, B, 3000, cached=True)
        n, n_ids = rpc.pack_prefill(seqs, buf)
        assert n_ids == sum(s.num_tokens - s.num_cached_tokens for s in seqs)
        for v, s in zip(rpc.unpack_prefill(buf[:n].copy(), B, n_ids, BLOCK), seqs):
            same_view(v, s, prefill=True)


def test_decode_record_is_the_device_stager_layout():
    """A worker hands the decode record to fast_meta's DecodeMetaStager without unpacking it: the
    two layouts must be the same bytes."""
    from byox_vllm_b200 import fast_meta as fm

    rng = random.Random(1)
    seqs = make_seqs(rng, 9, 4000)
    buf = np.zeros(4096, dtype=np.int32)
    n = rpc.pack_decode(seqs, buf)
    rec, _ = fm.pack_decode_record([s.last_token for s in seqs], [s.num_tokens for s in seqs],
                                   [s.block_table for s in seqs])
    assert buf[:n].tolist() == rec.tolist()


def test_write_and_read_call_kinds():
    rng = random.Random(2)
    block = bytearray(rpc.block_bytes(64, 4096, BLOCK, 16384))
    mv = memoryview(block)
    seqs = make_seqs(rng, 12, 4096)
    used = rpc.write_call(mv, "run", (seqs, False))
    assert used == rpc.HEADER_BYTES + 4 * (3 * 12 + sum(len(s.block_table) for s in seqs))
    name, args = rpc.read_call(mv, BLOCK)
    assert name == "_b200_run_decode" and args[1] == 12
    for v, s in zip(rpc.unpack_decode(args[0], 12, BLOCK), seqs):
        same_view(v, s, prefill=False)
    # the record handed out is a private copy: rank 0 overwrites the block on the next step
    first = args[0].copy()
    rpc.write_call(mv, "run", (make_seqs(rng, 3, 100), False))
    assert (args[0] == first).all()

    pseqs = make_seqs(rng, 7, 2000, cached=True)
    rpc.write_call(mv, "run", (pseqs, True))
    name, args = rpc.read_call(mv, BLOCK)
    assert name == "_b200_run_prefill" and args[1] == 7
    for v, s in zip(rpc.unpack_prefill(*args, BLOCK), pseqs):
        same_view(v, s, prefill=True)

    # everything else keeps the reference's pickle (model_runner.py:93-100): exit, odd arguments
    for call in (("exit", ()), ("run", ([], False)), ("run", ([1, 2], False)),
                 ("warmup", (3, {"a": 1}))):
        rpc.write_call(mv, *call)
        name, args = rpc.read_call(mv, BLOCK)
        assert name == call[0] and tuple(args) == tuple(call[1])
    rpc.write_call(mv, "run", (seqs[:3], "yes"))  # is_prefill is not a bool: not ours to compact
    name, args = rpc.read_call(mv, BLOCK)
    assert name == "run" and args[1] == "yes"
    assert [s.token_ids for s in args[0]] == [s.token_ids for s in seqs[:3]]


def test_block_sized_from_engine_limits_and_fallbacks():
    """The block is sized from max_num_seqs / max_model_len / block size / max_num_batched_tokens:
    the largest decode and prefill steps of BASELINE configs 3 and 4 fit, where the reference's
    pickle of the same step does not fit its 1 MiB constant."""
    rng = random.Random(3)
    for num_seqs, ctx, max_len, batched in ((128, 4096, 4096 + 512, 16384),      # config 3
                                            (256, 8192, 8192 + 64, 32768)):      # config 4
        size = rpc.block_bytes(num_seqs, max_len, BLOCK, batched)
        block = memoryview(bytearray(size))
        seqs = [Seq([rng.randrange(150000) for _ in range(ctx)],
                    [rng.randrange(1 << 20) for _ in range((ctx + BLOCK - 1) // BLOCK)])
                for _ in range(num_seqs)]
        used = rpc.write_call(block, "run", (seqs, False))
        assert rpc.read_call(block, BLOCK)[0] == "_b200_run_decode"
        reference_pickle = len(pickle.dumps(["run", seqs, False]))
        assert reference_pickle > (1 << 20) > used
        assert used <= 4 * num_seqs * (3 + (max_len + BLOCK - 1) // BLOCK) + rpc.HEADER_BYTES
        # the largest prefill step: max_num_batched_tokens uncached ids over whole sequences
        n_pre = batched // ctx
        used = rpc.write_call(block, "run", (seqs[:n_pre], True))
        assert rpc.read_call(block, BLOCK)[0] == "_b200_run_prefill"
        assert used < size
    # a step beyond the sized block falls back to the pickle, and that fails loudly when it does
    # not fit either (never a silent truncation)
    small = memoryview(bytearray(rpc.block_bytes(2, 512, BLOCK, 512)))
    few = make_seqs(rng, 2, 300)
    rpc.write_call(small, "run", (few, False))
    assert rpc.read_call(small, BLOCK)[0] == "_b200_run_decode"
    with pytest.raises(OverflowError):
        rpc.write_call(small, "run", (seqs[:64], True))  # 64 x 8192 uncached token ids


def _worker(name, size, event, n_calls, q):
    """Rank > 0: ModelRunner._loop's read side (model_runner.py:78-91) on the bound drop-in."""
    shm = SharedMemory(name=name)
    rpc.mark_attached(shm.buf, 1)  # what the bound SharedMemory subclass does on attach

    class Runner:
        world_size, rank, block_size = 2, 1, BLOCK

    r = Runner()
    r.shm, r.event = shm, event
    out = []
    for _ in range(n_calls):
        name_, args = rpc._read_shm(r)
        if name_ == "_b200_run_decode":
            views = rpc.unpack_decode(args[0], args[1], BLOCK)
            out.append(("d", [(v.last_token, v.num_tokens, v.block_table) for v in views]))
        elif name_ == "_b200_run_prefill":
            views = rpc.unpack_prefill(*args, BLOCK)
            out.append(("p", [(v.num_tokens, v.num_cached_tokens, v.block_table,
                               v[v.num_cached_tokens:]) for v in views]))
        else:
            out.append((name_, list(args)))
        q.put(len(out))  # acknowledge: rank 0 must not overwrite the block before it was read
    q.put(out)
    shm.close()


def test_two_processes_through_shared_memory():
    """Rank 0's _write_shm and a spawned worker's _read_shm over a real SharedMemory block and
    Event, the reference's protocol (model_runner.py:44-49,85-100)."""
    ctx = mp.get_context("spawn")
    size = rpc.block_bytes(32, 4096, BLOCK, 8192)
    shm = SharedMemory(create=True, size=size)
    try:
        event, q = ctx.Event(), ctx.Queue()
        rng = random.Random(4)
        steps = [("run", (make_seqs(rng, 5, 3000, cached=True), True)),
                 ("run", (make_seqs(rng, 5, 3000), False)),
                 ("run", (make_seqs(rng, 32, 4096), False)),
                 ("exit", ())]
        p = ctx.Process(target=_worker, args=(shm.name, size, event, len(steps), q))
        p.start()

        class Runner:
            world_size, rank = 2, 0

        r = Runner()
        r.shm, r.event = shm, [event]
        before = rpc.stats()
        for i, (name, args) in enumerate(steps):
            rpc._write_shm(r, name, *args)
            assert q.get(timeout=120) == i + 1
        got = q.get(timeout=120)
        p.join(timeout=60)
        assert p.exitcode == 0
        after = rpc.stats()
        assert after["steps"] - before["steps"] == 4 and after["bytes"] > before["bytes"]
        want = [("p", [(s.num_tokens, s.num_cached_tokens, s.block_table,
                        s[s.num_cached_tokens:]) for s in steps[0][1][0]]),
                ("d", [(s.last_token, s.num_tokens, s.block_table) for s in steps[1][1][0]]),
                ("d", [(s.last_token, s.num_tokens, s.block_table) for s in steps[2][1][0]]),
                ("exit", [])]
        assert got == want
    finally:
        shm.close()
        shm.unlink()


def test_mixed_installs_fail_loudly():
    """A rank that did not bind fast_rpc cannot read the block and the reference's pickles cannot
    be read here: a reader says what it found instead of unpickling garbage, and rank 0 refuses to
    send its first call until every worker has raised its flag (instead of hanging in NCCL later)."""
    rng = random.Random(6)
    block = memoryview(bytearray(rpc.block_bytes(8, 1024, BLOCK, 1024)))
    # the reference's format (model_runner.py:93-98): length word, then the pickle
    data = pickle.dumps(["run", [1, 2], False])
    block[0:4] = len(data).to_bytes(4, "little")
    block[4:4 + len(data)] = data
    with pytest.raises(RuntimeError, match="not the fast_rpc magic"):
        rpc.read_call(block, BLOCK)
    rpc.write_call(block, "run", (make_seqs(rng, 3, 500), False))
    assert int.from_bytes(block[0:4], "little") == rpc.MAGIC
    assert rpc.read_call(block, BLOCK)[0] == "_b200_run_decode"
    # an unbound reference reader would take MAGIC for the pickle length: far beyond any block
    assert rpc.MAGIC > (1 << 28)

    class Runner:
        world_size, rank = 4, 0

    r = Runner()

    class Ev:
        def __init__(self):
            self.n = 0

        def set(self):
            self.n += 1

    r.shm = type("Shm", (), {"buf": block})()
    r.event = [Ev(), Ev(), Ev()]
    rpc.mark_attached(block, 1)
    rpc.mark_attached(block, 3)
    assert rpc.missing_ranks(block, 4) == [2]
    old = rpc.ATTACH_TIMEOUT_S
    rpc.ATTACH_TIMEOUT_S = 0.05
    try:
        with pytest.raises(RuntimeError, match=r"rank\(s\) \[2\] did not attach"):
            rpc._write_shm(r, "run", make_seqs(rng, 3, 500), False)
        assert all(e.n == 0 for e in r.event)  # nothing was announced to the workers
        rpc.mark_attached(block, 2)
        rpc._write_shm(r, "run", make_seqs(rng, 3, 500), False)
        assert all(e.n == 1 for e in r.event)
        # the flags survive the header rewrite of every call
        assert rpc.missing_ranks(block, 4) == []
    finally:
        rpc.ATTACH_TIMEOUT_S = old


def test_reference_sequence_objects_and_bound_model_runner():
    """With the reference's own Sequence class (baseline/_ref): the records carry exactly what its
    builders read, and install(fast_rpc=True) rebinds the transport of the reference's ModelRunner
    (and uninstall restores it)."""
    ref = Path(__file__).resolve().parents[1] / "baseline" / "_ref"
    if not (ref / "nanovllm").is_dir():
        pytest.skip("baseline/_ref (the installed reference) is not present")
    if str(ref) not in sys.path:
        sys.path.insert(0, str(ref))
    from byox_vllm_b200 import install

    install.install(fuse_qkv=False, fast_meta=False, fast_rpc=True)
    try:
        import nanovllm.engine.model_runner as mr
        from nanovllm.engine.sequence import Sequence

        assert mr.ModelRunner._write_shm is rpc._write_shm
        assert mr.ModelRunner._read_shm is rpc._read_shm and mr.ModelRunner.run is rpc._run
        assert mr.SharedMemory is not SharedMemory and issubclass(mr.SharedMemory, SharedMemory)
        rng = random.Random(5)
        seqs = []
        for _ in range(20):
            s = Sequence([rng.randrange(150000) for _ in range(rng.randint(1, 3000))], BLOCK)
            s.block_table = [rng.randrange(1 << 20) for _ in range(s.num_blocks)]
            if s.num_tokens > BLOCK:
                s.num_cached_tokens = rng.randint(0, (s.num_tokens - 1) // BLOCK) * BLOCK
            seqs.append(s)
        block = memoryview(bytearray(rpc.block_bytes(32, 4096, BLOCK, 65536)))
        rpc.write_call(block, "run", (seqs, True))
        name, args = rpc.read_call(block, BLOCK)
        for v, s in zip(rpc.unpack_prefill(*args, BLOCK), seqs):
            same_view(v, s, prefill=True)
        for s in seqs:
            s.append_token(rng.randrange(150000))
            if s.num_blocks > len(s.block_table):
                s.block_table.append(rng.randrange(1 << 20))
        rpc.write_call(block, "run", (seqs, False))
        name, args = rpc.read_call(block, BLOCK)
        for v, s in zip(rpc.unpack_decode(args[0], args[1], BLOCK), seqs):
            same_view(v, s, prefill=False)
    finally:
        install.uninstall()
    import nanovllm.engine.model_runner as mr

    assert mr.ModelRunner._write_shm is not rpc._write_shm and mr.SharedMemory is SharedMemory
