#!/usr/bin/env python
"""Per-kernel counts of the Blackwell-specific SASS opcodes in libb200attn.so (cuobjdump -sass).

    python tools/sass_summary.py [--out profiles/r02_sass_opcodes.txt]

The mnemonics are the ones /opt/skills/guides/B200_PROFILING.md lists as proof of each hardware
path: UTCHMMA (tcgen05.mma), LDTM / STTM (tcgen05.ld / .st on TMEM), UTCBAR (tcgen05.commit),
UTMALDG (cp.async.bulk.tensor loads through a tensor map), UBLKCP (1-D cp.async.bulk), HMMA
(mma.sync), LDSM (ldmatrix), SYNCS (mbarrier). tests/test_abi.py asserts the non-zero ones.
"""
from __future__ import annotations

import argparse
import re
import subprocess
import sys
from collections import Counter, OrderedDict
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
OPCODES = ("UTCHMMA", "LDTM", "STTM", "UTCBAR", "UTMALDG", "UBLKCP", "HMMA", "LDSM", "MOVM",
           "SYNCS", "MUFU.EX2", "ACQBULK", "USETMAXREG")


def kernel_opcodes(lib: Path) -> "OrderedDict[str, Counter]":
    sass = subprocess.run(["cuobjdump", "-sass", str(lib)], capture_output=True, text=True,
                          check=True).stdout
    out: "OrderedDict[str, Counter]" = OrderedDict()
    cur = None
    for line in sass.splitlines():
        m = re.search(r"Function : (\S+)", line)
        if m:
            cur = out.setdefault(m.group(1), Counter())
            continue
        if cur is None:
            continue
        m = re.match(r"\s+/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", line)
        if not m:
            continue
        op = m.group(1)
        cur["_instructions"] += 1
        for name in OPCODES:
            if op == name or op.startswith(name + ".") or op.startswith(name + "_"):
                cur[name] += 1
    return out


def demangle(names):
    r = subprocess.run(["c++filt"], input="\n".join(names), capture_output=True, text=True)
    return r.stdout.splitlines() if r.returncode == 0 else list(names)


def short(name: str) -> str:
    name = name.replace("b200::(anonymous namespace)::", "").replace("b200::", "")
    name = re.sub(r"^void ", "", name)
    return re.sub(r"\(.*$", "", name)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--lib", default=str(ROOT / "byox_vllm_b200" / "libb200attn.so"))
    ap.add_argument("--out", default=None)
    args = ap.parse_args()
    table = kernel_opcodes(Path(args.lib))
    names = demangle(table.keys())
    lines = [f"# cuobjdump -sass {Path(args.lib).name}: opcode counts per kernel (sm_100a)",
             "# " + " ".join(f"{o:>9s}" for o in ("instrs",) + OPCODES) + "  kernel"]
    for (_, c), n in zip(table.items(), names):
        lines.append("  " + " ".join(f"{c.get(o, 0):9d}" for o in ("_instructions",) + OPCODES) +
                     "  " + short(n))
    text = "\n".join(lines) + "\n"
    if args.out:
        Path(args.out).write_text(text)
    sys.stdout.write(text)


if __name__ == "__main__":
    main()
