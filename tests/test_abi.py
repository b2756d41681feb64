"""CPU-only: the C-ABI library loads and exports every symbol include/mrb.h declares; no compute calls."""
import os
import re
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    src = open(os.path.join(ROOT, "include", "mrb.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(mrb_[a-z0-9_]+)\s*\(", src)))


def test_header_declares_the_path():
    syms = declared_symbols()
    for need in ("mrb_fast_perim", "mrb_sqdist", "mrb_sqrt", "mrb_dfield", "mrb_create_bands", "mrb_find_indexes",
                 "mrb_index_map", "mrb_prime_qp", "mrb_run_qp", "mrb_optimize", "mrb_mcubes", "mrb_write_ply",
                 "mrb_reconstruct"):
        assert need in syms


def test_library_exports_every_declared_symbol():
    from mesh_reconstruction_b200 import _lib
    assert os.path.exists(_lib.SO_PATH), "libmrb.so not built: run __graft_entry__.build()"
    out = subprocess.check_output(["nm", "-D", "--defined-only", _lib.SO_PATH]).decode()
    exported = set(re.findall(r" T (mrb_[a-z0-9_]+)", out))
    missing = [s for s in declared_symbols() if s not in exported]
    assert not missing, missing
    lib = _lib.load()                      # binds every symbol with ctypes
    assert set(_lib.SYMBOLS) == set(declared_symbols())
    assert lib.mrb_version().decode().endswith("sm_100a")


def test_no_cpu_fallback_without_gpu():
    """Without a GPU the context creation must fail loudly (there is no CPU path in the product)."""
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    import mesh_reconstruction_b200 as mrb
    with pytest.raises(mrb.MrbError):
        mrb.Context(0)


def test_product_does_not_reference_the_oracle():
    """The oracle is test infrastructure: no product file may call (mrbo_*), import or load it."""
    pkg = os.path.join(ROOT, "mesh_reconstruction_b200")
    bad = []
    for dirpath, _, files in os.walk(pkg):
        if "_build" in dirpath:
            continue
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".cpp")):
                txt = open(os.path.join(dirpath, f), errors="ignore").read()
                if "mrbo_" in txt or "libmrb_oracle" in txt or re.search(r"^\s*(from|import)\s+oracle", txt, re.M):
                    bad.append(f)
    assert not bad, bad


def test_sweep_kernels_keep_the_brick_in_registers():
    """ptxas must not put any part of a brick update on the stack (it did once at 512 threads: 3.0 -> 4.3 ms)."""
    log = os.path.join(ROOT, "mesh_reconstruction_b200", "_build", "mrb_bricks.ptxas.log")
    if not os.path.exists(log):
        pytest.skip("no ptxas log (library built elsewhere)")
    text = open(log).read()
    blocks = re.split(r"ptxas info\s*: Compiling entry function ", text)
    seen = 0
    for b in blocks:
        m = re.match(r"'(\S+)'", b)
        if not m or not any(k in m.group(1) for k in ("k_qp_bricks_tiledILi480ELi1ELb0", "k_qp_bricks_shardedILi256", "11k_qp_bricksILi256ELi2ELb0")):
            continue
        seen += 1
        assert re.search(r"\b0 bytes stack frame, 0 bytes spill stores, 0 bytes spill loads", b), m.group(1)
    assert seen >= 3
