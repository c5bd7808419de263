"""CPU-side checks of the drop-in boundary: the library loads, exports exactly what
include/tfp_store.h declares, and refuses to work without a GPU (no CPU fallback)."""
import ctypes
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    text = open(os.path.join(ROOT, "include", "tfp_store.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(tfp_[a-z0-9_]+)\s*\(", text)))


def test_header_declares_the_store_api():
    syms = declared_symbols()
    for name in ("tfp_store_open", "tfp_store_close", "tfp_store_set", "tfp_store_delete", "tfp_store_get",
                 "tfp_store_find", "tfp_store_find_batch", "tfp_last_error"):
        assert name in syms


def test_library_exports_every_declared_symbol(tfp):
    L = ctypes.CDLL(tfp.lib_path())
    for name in declared_symbols():
        assert hasattr(L, name), f"{name} is declared in include/tfp_store.h but not exported"


def test_binding_covers_every_declared_symbol(tfp):
    from localai_tfp_b200 import _lib
    assert sorted(_lib.PROTOTYPES) == declared_symbols()


def test_header_cites_the_reference():
    text = open(os.path.join(ROOT, "include", "tfp_store.h")).read()
    for cite in ("store.go:106-198", "store.go:200-264", "store.go:266-311", "store.go:477-511"):
        assert cite in text


def test_product_never_touches_the_oracle():
    pkg = os.path.join(ROOT, "localai_tfp_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".hpp", ".h", ".inl")):
                src = open(os.path.join(dirpath, f), errors="replace").read()
                assert "oracle" not in src.lower(), f"{f} mentions the oracle"


@pytest.mark.parametrize("compiler,std,ext", [("gcc", "-std=c99", ".c"), ("g++", "-std=c++17", ".cpp")])
def test_header_is_plain_c(tmp_path, compiler, std, ext):
    """The boundary is a C ABI: the header must compile as strict C99 (what cgo's C compiler sees) and as C++,
    with nothing but the standard headers."""
    import shutil
    import subprocess
    if not shutil.which(compiler):
        pytest.skip(compiler + " is not installed")
    src = tmp_path / ("use_header" + ext)
    src.write_text('#include "tfp_store.h"\n'
                   "int main(void) { tfp_cfg cfg; tfp_stats st; (void)cfg; (void)st; return tfp_version() == 0; }\n")
    r = subprocess.run([compiler, std, "-Wall", "-Wextra", "-pedantic", "-Werror", "-fsyntax-only",
                        "-I", os.path.join(ROOT, "include"), str(src)], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr


def test_open_fails_loudly_without_a_gpu(tfp):
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    with pytest.raises(RuntimeError) as e:
        tfp.Store()
    assert "no CPU fallback" in str(e.value)
