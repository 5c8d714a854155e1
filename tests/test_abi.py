"""CPU: the C-ABI library loads and exports every symbol include/q4k_b200.h declares (no compute)."""
import ctypes
import os
import re

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "ggml-neon-opt_b200", "lib", "libq4k_b200.so")


def declared():
    src = open(os.path.join(ROOT, "include", "q4k_b200.h")).read()
    return sorted(set(re.findall(r"\b(q[46]k_[a-z0-9_A-Z]+)\s*\(", src)))


def test_exports_match_header():
    assert os.path.exists(LIB), "build with __graft_entry__.build()"
    lib = ctypes.CDLL(LIB)
    names = declared()
    assert len(names) >= 12
    for n in names:
        assert hasattr(lib, n), n


def test_pure_host_entries():
    lib = ctypes.CDLL(LIB)
    lib.q4k_mul_mat_workspace_bytes.restype = ctypes.c_size_t
    lib.q4k_mul_mat_workspace_bytes.argtypes = [ctypes.c_int64, ctypes.c_int64]
    assert lib.q4k_abi_version() == 1
    assert lib.q4k_mul_mat_workspace_bytes(2048, 1) == 8 * 292
    assert lib.q4k_mul_mat_workspace_bytes(100, 1) == 0
    lib.q4k_build_info.restype = ctypes.c_char_p
    assert b"sm_100a" in lib.q4k_build_info()


def test_product_path_never_imports_oracle():
    pkg = os.path.join(ROOT, "ggml-neon-opt_b200")
    for d, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".cpp", ".c")):
                txt = open(os.path.join(d, f)).read()
                assert "oracle" not in txt.replace("the oracle", ""), os.path.join(d, f)


def test_q6k_host_entries():
    lib = ctypes.CDLL(LIB)
    lib.q6k_mul_mat_workspace_bytes.restype = ctypes.c_size_t
    lib.q6k_mul_mat_workspace_bytes.argtypes = [ctypes.c_int64, ctypes.c_int64]
    assert lib.q6k_mul_mat_workspace_bytes(5632, 2) == 2 * 22 * 292
    assert lib.q6k_mul_mat_workspace_bytes(100, 1) == 0
    assert "q6k_mul_mat" in declared() and "q6k_debug_partials" in declared()
