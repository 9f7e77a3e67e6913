"""CPU suite: the C-ABI library builds for sm_100a, loads, and exports every symbol that
include/docagents_b200.h declares.  No compute calls here (no GPU in the build container)."""
import ctypes as C
import os
import re
import subprocess

import pytest

from doc_agents_b200 import _lib as L

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "docagents_b200.h")


def declared_symbols():
    text = open(HEADER).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(da_[a-z0-9_]+)\s*\(", text)))


def test_header_declares_the_path_entry_points():
    syms = declared_symbols()
    for must in ("da_store_create", "da_store_destroy", "da_store_save_chunks", "da_store_save_embeddings",
                 "da_store_topk", "da_store_search", "da_last_error"):
        assert must in syms


def test_library_exports_every_declared_symbol(built_lib):
    lib = C.CDLL(built_lib)
    for s in declared_symbols():
        assert hasattr(lib, s), f"{s} is declared in the header but not exported"
    out = subprocess.run(["nm", "-D", "--defined-only", built_lib], capture_output=True, text=True, check=True).stdout
    exported = set(re.findall(r" T (da_[a-z0-9_]+)", out))
    assert exported == set(declared_symbols()), exported ^ set(declared_symbols())
    assert set(L.SIGNATURES) == set(declared_symbols())       # the ctypes twin binds exactly the header


def test_header_compiles_as_c_and_struct_sizes_match(built_lib, tmp_path):
    """The header is plain C (cgo includes it as C) and the ctypes structs have the C layout."""
    src = tmp_path / "sz.c"
    src.write_text('#include <stdio.h>\n#include "docagents_b200.h"\n'
                   'int main(void){printf("%zu %zu %zu %zu %zu\\n", sizeof(da_config), sizeof(da_search_params),'
                   'sizeof(da_ivf_params), sizeof(da_synth_params), sizeof(da_stats));return 0;}\n')
    exe = tmp_path / "sz"
    subprocess.run(["/usr/bin/gcc", "-std=c99", "-Wall", "-Werror", "-I", os.path.dirname(HEADER), str(src), "-o", str(exe)],
                   check=True)
    sizes = [int(x) for x in subprocess.run([str(exe)], capture_output=True, text=True, check=True).stdout.split()]
    assert sizes == [C.sizeof(L.da_config), C.sizeof(L.da_search_params), C.sizeof(L.da_ivf_params),
                     C.sizeof(L.da_synth_params), C.sizeof(L.da_stats)]


def test_library_is_sm100a_and_has_no_torch_dependency(built_lib):
    needed = subprocess.run(["readelf", "-d", built_lib], capture_output=True, text=True, check=True).stdout
    assert "libnccl" in needed                                  # cudart is linked statically
    assert "torch" not in needed and "libc10" not in needed
    sass = subprocess.run(["cuobjdump", "-lelf", built_lib], capture_output=True, text=True)
    if sass.returncode == 0:
        assert "sm_100a" in sass.stdout


def test_abi_version_and_error_channel(built_lib):
    lib = L.load()
    assert lib.da_abi_version() == 1
    # argument errors are reported without touching a device
    assert lib.da_store_create(None, None) == L.DA_ERR_INVALID
    assert "NULL" in L.last_error()
    cfg = L.da_config()
    cfg.struct_size = 3
    h = C.c_void_p()
    assert lib.da_store_create(C.byref(cfg), C.byref(h)) == L.DA_ERR_INVALID
    cfg.struct_size = C.sizeof(L.da_config)
    cfg.dim = 3070
    cfg.n_devices = 1
    cfg.capacity_rows = 10
    assert lib.da_store_create(C.byref(cfg), C.byref(h)) == L.DA_ERR_DIM
    assert lib.da_store_save_embeddings(None, None, None, 0, 0) == L.DA_ERR_INVALID


def test_no_silent_fallback_without_a_gpu(built_lib):
    """Without a CUDA device the product path must fail loudly (no CPU fallback, north_star)."""
    import doc_agents_b200
    try:
        import torch
        has_gpu = torch.cuda.is_available()
    except Exception:
        has_gpu = os.path.exists("/dev/nvidia0")
    if has_gpu:
        pytest.skip("a GPU is present; the loud-failure path is exercised in the CPU container")
    with pytest.raises(doc_agents_b200.StoreError) as ei:
        doc_agents_b200.GpuStore(dim=3072, capacity_rows=16)
    assert ei.value.code == L.DA_ERR_CUDA
    assert "no CPU fallback" in ei.value.message


def test_product_never_touches_the_oracle():
    """oracle/ is test infrastructure: nothing under doc_agents_b200/ may import or link it."""
    pkg = os.path.join(ROOT, "doc_agents_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".h", ".cuh", ".cpp")):
                text = open(os.path.join(dirpath, f), errors="replace").read()
                assert not re.search(r"^\s*(import|from)\s+oracle\b", text, flags=re.M), f
                assert "liboracle" not in text and '#include "oracle' not in text, f


def test_library_synthetic_queries_and_ids_equal_the_oracle(built_lib):
    """bench.py takes its queries and ids from the product library (it may not touch oracle/ outside the
    cpu_baseline leg); they must be the oracle's, bit for bit.  Host-only code: runs without a GPU."""
    import numpy as np
    import oracle
    from doc_agents_b200 import synth
    for kw in ({}, {"n_clusters": 7, "cmul": 3, "nmul": 2, "seed_query": 99}):
        so, sp = oracle.Synth(**kw), synth.Synth(**kw)
        for dim in (3072, 1536, 100, 4):
            for qi, cl in ((0, 0), (5, 3), (123456, 6)):
                a, b = oracle.gen_query(so, qi, cl, dim), synth.gen_query(sp, qi, cl, dim)
                assert np.array_equal(a.view(np.uint32), b.view(np.uint32))
        assert np.array_equal(oracle.chunk_ids(so, 12345, 50), synth.chunk_ids(sp, 12345, 50))
        assert np.array_equal(oracle.doc_ids(so, 0, 50), synth.doc_ids(sp, 0, 50))


def test_python_flag_constants_match_the_header():
    """The ctypes twin restates the header's flag and error constants by hand: keep them equal."""
    import re
    from doc_agents_b200 import _lib as L
    text = open(HEADER).read()
    seen = 0
    for name, value in re.findall(r"#define\s+(DA_(?:SEARCH|STORE|IVF|ERR)_[A-Z0-9_]+)\s+\(?(-?0x[0-9a-fA-F]+|-?\d+)u?\)?", text):
        if hasattr(L, name):
            assert getattr(L, name) == int(value, 0), name
            seen += 1
    assert seen >= 9
    for name, value in re.findall(r"\b(DA_(?:OK|ERR_[A-Z_]+))\s*=\s*(-?\d+)", text):      # the status enum
        assert getattr(L, name) == int(value), name
    for name in ("DA_STORE_SPLIT16", "DA_STORE_Q8_SIDECAR", "DA_SEARCH_NO_TENSOR", "DA_SEARCH_NO_Q8", "DA_SEARCH_NO_COMPACT"):
        assert hasattr(L, name) and re.search(r"#define\s+%s\b" % name, text), name
