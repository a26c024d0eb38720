"""Step-by-step statistics of two tools/engine_logits.py recordings (max / rms difference, rms of the
reference logits): the numbers behind the tolerances of tests/test_engine_parity.py."""
import json
import sys

import torch

a, b = torch.load(sys.argv[1]), torch.load(sys.argv[2])
rows = []
for i, ((p_r, n_r, l_r), (p_o, n_o, l_o)) in enumerate(zip(a["steps"], b["steps"])):
    x, y = l_r.float(), l_o.float()
    d = (x - y).abs()
    rows.append(dict(step=i, prefill=p_r, tokens=n_r, same_batch=(p_r, n_r) == (p_o, n_o),
                     max_abs=d.max().item(), rms_diff=d.pow(2).mean().sqrt().item(),
                     rms_ref=x.pow(2).mean().sqrt().item(), max_ref=x.abs().max().item(),
                     excess_2e2=(d - (2e-2 + 1e-2 * x.abs())).max().item()))
print(json.dumps(dict(a=a["info"], b=b["info"], worst_max_abs=max(r["max_abs"] for r in rows),
                      worst_rms_ratio=max(r["rms_diff"] / r["rms_ref"] for r in rows),
                      steps=rows)))
