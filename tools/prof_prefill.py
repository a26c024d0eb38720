"""Three launches of the tcgen05 prefill kernel per shape, for ncu (1x16384 32/8, then the reference
bench's 24-prompt prefill step 16/8):

    ncu --set full --clock-control none --import-source on -k regex:prefill_tc_kernel -s 2 -c 1 \
        -o gpurun_out/prof_prefill_1x16384 python tools/prof_prefill.py
    ncu ... -s 5 -c 1 -o gpurun_out/prof_prefill_bench24 python tools/prof_prefill.py
"""
import itertools
import random
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
import torch  # noqa: E402

from byox_vllm_b200.attention import prefill_varlen  # noqa: E402

rng = random.Random(0)
SHAPES = [([16384], 32, 8), ([rng.randint(100, 1024) for _ in range(24)], 16, 8)]
with torch.inference_mode():
    for lens, Hq, Hkv in SHAPES:
        total = sum(lens)
        g = torch.Generator(device="cuda").manual_seed(1)
        q = torch.empty(total, Hq, 128, dtype=torch.bfloat16, device="cuda").normal_(generator=g)
        k = torch.empty(total, Hkv, 128, dtype=torch.bfloat16, device="cuda").normal_(generator=g)
        v = torch.empty(total, Hkv, 128, dtype=torch.bfloat16, device="cuda").normal_(generator=g)
        cu = torch.tensor([0] + list(itertools.accumulate(lens)), dtype=torch.int32, device="cuda")
        o = torch.empty_like(q)
        for _ in range(3):
            prefill_varlen(q, k, v, cu, cu, max(lens), max(lens), 128 ** -0.5, out=o)
        torch.cuda.synchronize()
