"""The C-ABI library loads and exports every symbol include/b200attn.h declares (no GPU)."""
import ctypes
import re
import subprocess

from conftest import ROOT


def _declared():
    text = (ROOT / "include" / "b200attn.h").read_text()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    decls = {}
    for m in re.finditer(r"\b(b200attn_\w+)\s*\(([^;{]*?)\)\s*;", text, flags=re.S):
        args = m.group(2).strip()
        n = 0 if args in ("", "void") else args.count(",") + 1
        decls[m.group(1)] = n
    return decls


def test_header_declares_expected_entry_points():
    d = _declared()
    for name in ("b200attn_store_kvcache", "b200attn_decode", "b200attn_store_decode",
                 "b200attn_prefill_varlen", "b200attn_last_error", "b200attn_abi_version",
                 "b200attn_decode_pick_splits", "b200attn_decode_workspace_bytes"):
        assert name in d


def test_library_exports_every_declared_symbol(built_lib):
    lib = ctypes.CDLL(str(built_lib))
    for name in _declared():
        assert hasattr(lib, name), f"{name} declared in b200attn.h but not exported"
    out = subprocess.run(["nm", "-D", "--defined-only", str(built_lib)], capture_output=True,
                         text=True).stdout
    exported = set(re.findall(r"\bT (b200attn_\w+)", out))
    assert exported == set(_declared()), "exported C symbols and header declarations differ"


def test_python_prototypes_match_header(built_lib):
    from byox_vllm_b200 import _lib

    d = _declared()
    assert set(_lib.PROTOTYPES) == set(d)
    for name, (_, argtypes) in _lib.PROTOTYPES.items():
        assert len(argtypes) == d[name], f"{name}: {len(argtypes)} ctypes args vs {d[name]} in header"
    lib = _lib.load()
    assert lib.b200attn_abi_version() == _lib.ABI_VERSION


def test_no_torch_types_in_abi():
    text = (ROOT / "include" / "b200attn.h").read_text()
    code = re.sub(r"/\*.*?\*/", "", text, flags=re.S)  # comments may mention the Python side
    assert "torch" not in code.lower()
    assert "at::" not in code and "c10::" not in code and "#include <torch" not in text


def test_sass_is_sm100a_only(built_lib):
    out = subprocess.run(["cuobjdump", "--list-elf", str(built_lib)], capture_output=True,
                         text=True).stdout
    archs = set(re.findall(r"sm_(\d+a?)", out))
    assert archs == {"100a"}, archs


def test_sass_uses_the_blackwell_paths(built_lib):
    """tcgen05 / TMEM / TMA opcodes are present in the kernels DESIGN.md says use them
    (tools/sass_summary.py; the committed table is profiles/r02_sass_opcodes.txt)."""
    from tools import sass_summary

    table = sass_summary.kernel_opcodes(built_lib)
    names = sass_summary.demangle(table.keys())
    by_kernel = {}
    for (_, counts), n in zip(table.items(), names):
        by_kernel.setdefault(sass_summary.short(n).split("<")[0], []).append(counts)
    for c in by_kernel["prefill_tc_kernel"]:
        for op in ("UTCHMMA", "LDTM", "STTM", "UTCBAR", "UTMALDG"):
            assert c[op] > 0, f"prefill_tc_kernel without {op}"
    for c in by_kernel["decode_stream_kernel"]:
        assert c["UTMALDG"] > 0 and c["HMMA"] > 0 and c["LDSM"] > 0
    for c in by_kernel["qk_norm_rope_bulk_kernel"]:
        assert c["UBLKCP"] > 0
