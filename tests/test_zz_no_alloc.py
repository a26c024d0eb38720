"""b200attn_decode_static_scratch: the split decode for callers that cannot allocate.

Who needs it: an engine that gave all free HBM to its KV cache (a later torch.empty / cudaMalloc for
the split workspace fails), a plugin running inside somebody else's stream capture, a cgo / JNI
binding that owns no device allocator. They pass inputs and the output only; partial results and
counters live in an array built into libb200attn.so (8 MiB + 4096 counters per device).

The GPU test calls the new entry point through the C-ABI twice in a row, on different batches, and
compares each result with b200attn_decode (caller workspace + registered counters) on the same
inputs, at the tolerance of the other decode tests (gpu_util.ATOL / RTOL). The second call finds the
first call's partials in the shared array; nothing of them may show in its output.

This file sorts last on purpose: under `pytest -x` a failure here hides no other test.
"""
import pytest
import torch

from gpu_util import assert_close, to_dev
from tools import synth


def _static_decode(lib, d, scale, splits, stream=0):
    """Raw C-ABI call: device pointers and sizes only, no workspace argument."""
    from byox_vllm_b200 import _lib

    m = d["meta"]
    q, kc, vc = d["q"], d["k_cache"], d["v_cache"]
    bt, lens = m["block_tables"], m["context_lens"]
    B, Hq, D = q.shape
    num_blocks, block_size, Hkv, _ = kc.shape
    out = torch.full_like(q, float("nan"))  # every row must be written by the call
    rc = lib.b200attn_decode_static_scratch(
        q.data_ptr(), kc.data_ptr(), vc.data_ptr(), bt.data_ptr(), bt.stride(0), bt.size(1),
        lens.data_ptr(), out.data_ptr(), B, Hq, Hkv, D, block_size, num_blocks, float(scale),
        splits, _lib.dtype_code(q.dtype), stream)
    _lib.check("b200attn_decode_static_scratch", rc)
    return out


@pytest.mark.gpu
def test_static_scratch_decode_matches_decode_twice_in_a_row(cuda_dev):
    from byox_vllm_b200 import _lib
    from byox_vllm_b200.attention import paged_decode

    lib = _lib.load()
    # call 1: ragged rows incl. a padding row (ctx 0), 16/8 heads, every row cut in 4
    # call 2: another batch, GQA group 8, rows cut in 8: overwrites part of call 1's partials only
    shapes = [([300, 1500, 64, 2049, 1], 16, 8, 4, 71), ([4096, 777], 8, 1, 8, 72)]
    cases = []
    for ctx, Hq, Hkv, splits, seed in shapes:
        c = synth.decode_case(len(ctx), ctx, Hq, Hkv, seed=seed)
        d = to_dev(c, cuda_dev)
        cases.append((c, d, splits))
    cases[0][1]["meta"]["context_lens"][4] = 0
    want = []
    for c, d, splits in cases:  # the existing entry point, torch-allocated workspace
        m = d["meta"]
        want.append(paged_decode(d["q"], d["k_cache"], d["v_cache"], m["context_lens"],
                                 m["block_tables"], c.scale, num_splits=splits))
    torch.cuda.synchronize()

    # first call on the device resolves the array's address; after it, no allocation may happen
    got0 = _static_decode(lib, cases[0][1], cases[0][0].scale, cases[0][2])
    free_before = torch.cuda.mem_get_info()[0]
    got1 = _static_decode(lib, cases[1][1], cases[1][0].scale, cases[1][2])
    got0_again = _static_decode(lib, cases[0][1], cases[0][0].scale, cases[0][2])
    torch.cuda.synchronize()
    assert torch.cuda.mem_get_info()[0] >= free_before, "device memory was allocated by the call"

    for name, got, ref in (("call 1", got0, want[0]), ("call 2", got1, want[1]),
                           ("call 3 (call 1's batch again)", got0_again, want[0])):
        assert torch.isfinite(got.float()).all(), f"{name}: rows left unwritten"
        assert_close(got, ref, f"{name}: static scratch vs b200attn_decode")
    assert torch.count_nonzero(got0[4]) == 0  # padding row gives zeros, as b200attn_decode does

    # num_splits == 1 touches no scratch at all and agrees too
    one = _static_decode(lib, cases[0][1], cases[0][0].scale, 1)
    torch.cuda.synchronize()
    assert_close(one, want[0], "static scratch, one split")


def test_static_scratch_rejects_sizes_beyond_the_array(built_lib):
    """Too many partials or too many counters: B200ATTN_E_WORKSPACE with the sizes in the message,
    before anything touches the GPU (so this runs without one; pointers are never dereferenced)."""
    from byox_vllm_b200 import _lib

    lib = _lib.load()

    def call(batch, Hq, Hkv, splits, head_dim=128):
        return lib.b200attn_decode_static_scratch(
            256, 256, 256, 256, 16, 16, 256, 256, batch, Hq, Hkv, head_dim, 256, 64, 0.088,
            splits, _lib.DTYPE_BF16, 0)

    need = lib.b200attn_decode_workspace_bytes(64, 32, 128, 8, 0)
    assert need > (8 << 20)
    assert call(64, 32, 8, 8) == _lib.E_WORKSPACE          # partials: 64 x 32 x 8 slots > 8 MiB
    assert b"built-in" in lib.b200attn_last_error() and str(need).encode() in lib.b200attn_last_error()
    assert call(4095, 1, 1, 2) == _lib.E_WORKSPACE          # 4095 rows x 2 splits: same limit
    assert call(0, 16, 8, 4) == 0                           # empty batch: nothing to do
    assert call(4, 16, 8, 0) == _lib.E_BADARG               # num_splits < 1
    assert call(4, 16, 8, 4, head_dim=64) == _lib.E_UNSUPPORTED
