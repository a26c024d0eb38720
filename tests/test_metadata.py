"""Host logic: the metadata builders reproduce the reference's expectations
(reference tests/test_prepare_funcs.py:17-206, block_size 16) and scale to block_size 256."""
import torch

from tools import metadata_ref as md
from byox_vllm_b200.context import get_context, reset_context, set_context


def S(n, table, cached=0):
    return md.SeqState(num_tokens=n, block_table=list(table), num_cached_tokens=cached)


def test_prefill_single_sequence():  # test_prepare_funcs.py:17-36
    m = md.prepare_prefill([S(5, [0])], 16)
    assert m["cu_seqlens_q"].tolist() == [0, 5] and m["cu_seqlens_k"].tolist() == [0, 5]
    assert m["max_seqlen_q"] == 5 and m["max_seqlen_k"] == 5
    assert m["slot_mapping"].tolist() == [0, 1, 2, 3, 4]
    assert m["context_lens"] is None and m["block_tables"] is None
    assert m["positions"].tolist() == [0, 1, 2, 3, 4]


def test_prefill_multiple_sequences():  # :39-59
    m = md.prepare_prefill([S(3, [0]), S(5, [1])], 16)
    assert m["cu_seqlens_q"].tolist() == [0, 3, 8] and m["cu_seqlens_k"].tolist() == [0, 3, 8]
    assert m["max_seqlen_q"] == 5 and m["max_seqlen_k"] == 5
    assert m["slot_mapping"].tolist() == [0, 1, 2, 16, 17, 18, 19, 20]
    assert m["block_tables"] is None
    assert m["positions"].tolist() == [0, 1, 2, 0, 1, 2, 3, 4]


def test_prefill_with_cached_tokens():  # :62-80 (num_cached_blocks = 3 // 16 = 0)
    m = md.prepare_prefill([S(6, [0, 1], cached=3)], 16)
    assert m["cu_seqlens_q"].tolist() == [0, 3] and m["cu_seqlens_k"].tolist() == [0, 6]
    assert m["max_seqlen_q"] == 3 and m["max_seqlen_k"] == 6
    assert m["slot_mapping"].tolist() == [0, 1, 2, 3, 4, 5]
    assert m["block_tables"].tolist() == [[0, 1]]
    assert m["positions"].tolist() == [3, 4, 5]


def test_prefill_multiple_with_cached_tokens():  # :146-166
    m = md.prepare_prefill([S(5, [0], cached=2), S(6, [1], cached=1)], 16)
    assert m["cu_seqlens_q"].tolist() == [0, 3, 8] and m["cu_seqlens_k"].tolist() == [0, 5, 11]
    assert m["max_seqlen_q"] == 5 and m["max_seqlen_k"] == 6
    assert m["slot_mapping"].tolist() == [0, 1, 2, 3, 4, 16, 17, 18, 19, 20, 21]
    assert m["positions"].tolist() == [2, 3, 4, 1, 2, 3, 4, 5]


def test_decode_cases():  # :83-143, :169-206 (the mock's -1 last entries give negative slots)
    m = md.prepare_decode([S(5, [0, -1, -1, -1])], 16)
    assert m["slot_mapping"].tolist() == [-12] and m["context_lens"].tolist() == [5]
    assert m["block_tables"].tolist() == [[0, -1, -1, -1]] and m["positions"].tolist() == [4]
    m = md.prepare_decode([S(3, [0, -1, -1, -1]), S(4, [1, -1, -1, -1])], 16)
    assert m["slot_mapping"].tolist() == [-14, -13] and m["context_lens"].tolist() == [3, 4]
    m = md.prepare_decode([S(4, [0, -1, -1, -1], 2), S(5, [1, -1, -1, -1], 3)], 16)
    assert m["slot_mapping"].tolist() == [-13, -12] and m["positions"].tolist() == [3, 4]
    assert m["cu_seqlens_q"] is None and m["max_seqlen_q"] == 0


def test_block_size_256_and_padding():
    seqs = [S(257, [7, 3]), S(256, [5]), S(1000, [9, 8, 2, 1])]
    m = md.prepare_decode(seqs, 256)
    assert m["slot_mapping"].tolist() == [3 * 256 + 0, 5 * 256 + 255, 1 * 256 + 231]
    assert m["block_tables"].tolist() == [[7, 3, -1, -1], [5, -1, -1, -1], [9, 8, 2, 1]]
    assert m["block_tables"].dtype == torch.int32 and m["context_lens"].dtype == torch.int32
    p = md.prepare_prefill([S(600, [4, 6, 0], cached=512), S(300, [1, 2])], 256)
    assert p["cu_seqlens_q"].tolist() == [0, 88, 388] and p["cu_seqlens_k"].tolist() == [0, 600, 900]
    assert p["slot_mapping"].numel() == 388  # == total_q (reference attention.py:46)
    assert p["slot_mapping"][:88].tolist() == list(range(0, 88))  # block 0, third page
    assert p["block_tables"].tolist() == [[4, 6, 0], [1, 2, -1]]
    # warm-up sequences (no block table) produce no slots (model_runner.py:179-180)
    w = md.prepare_prefill([S(4096, []), S(4096, [])], 256)
    assert w["slot_mapping"].numel() == 0 and w["block_tables"] is None


def test_context_contract_roundtrip():
    reset_context()
    c = get_context()
    assert c.is_prefill is False and c.slot_mapping is None and c.max_seqlen_q == 0
    m = md.prepare_decode([S(5, [0])], 16)
    set_context(**md.context_kwargs(m))
    c = get_context()
    assert not c.is_prefill and c.context_lens.tolist() == [5] and c.cu_seqlens_q is None
    reset_context()
    assert get_context().context_lens is None
    fields = [f for f in get_context().__dataclass_fields__]
    assert fields == ["is_prefill", "cu_seqlens_q", "cu_seqlens_k", "max_seqlen_q", "max_seqlen_k",
                      "slot_mapping", "context_lens", "block_tables"]  # context.py:6-15 order
