"""CPU tests: the C-ABI library loads, exports every symbol include/lpgpu.h declares, refuses to
run without a GPU (no CPU fallback), and its host-side helpers agree bit for bit with the oracle."""
import ctypes as C
import os
import re

import numpy as np
import pytest

import helpers as H
import oracle_lib as O

import linearprogramming_b200 as lp
from linearprogramming_b200 import capi

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared_symbols():
    src = open(os.path.join(ROOT, "include", "lpgpu.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(lpgpu_[a-z_0-9]+)\s*\(", src)))


def test_library_exports_every_declared_symbol():
    L = capi.lib()
    names = _declared_symbols()
    assert len(names) >= 25
    for n in names:
        assert hasattr(L, n), f"{n} declared in include/lpgpu.h but not exported"
        assert n in capi.SIGNATURES, f"{n} has no ctypes signature in capi.py"
    assert sorted(capi.SIGNATURES) == names


def test_struct_layouts_match_header():
    L = capi.lib()
    assert C.sizeof(capi.Opts) == L.lpgpu_abi_sizeof(0) == 40 and C.sizeof(capi.Result) == L.lpgpu_abi_sizeof(1) == 48
    o = capi.default_opts()
    assert o.eps == 1e-9 and o.method == lp.PLAIN and o.check_every == 64 and o.max_iter == 2**31 - 1


def test_no_cpu_fallback_without_device():
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    with pytest.raises(lp.LpgpuError):
        lp.Context(0)


def test_product_does_not_import_oracle():
    pkg = os.path.join(ROOT, "linearprogramming_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".c", ".cpp")):
                text = open(os.path.join(dirpath, f), errors="replace").read()
                assert "oracle_lib" not in text and "lp_oracle" not in text and "liblp_oracle" not in text, f


@pytest.mark.parametrize("seed", range(8))
def test_build_tableau_matches_oracle_bitwise(seed):
    rng = np.random.default_rng(100 + seed)
    m, ns = int(rng.integers(2, 20)), int(rng.integers(2, 20))
    A, b, c, rel = H.random_mixed_lp(rng, m, ns)
    To, bo, ao = O.build_tableau(A, b, c, rel, R=2)
    Tp, bp, ap = lp.build_tableau(A, b, c, rel, R=2)
    assert ao == ap and np.array_equal(bo, bp)
    assert To.shape == Tp.shape and np.array_equal(To, Tp) and np.array_equal(np.signbit(To), np.signbit(Tp))


def test_build_tableau_plain_rejects_artificials():
    A = np.ones((2, 2)); b = np.array([1.0, 1.0]); c = np.ones(2)
    with pytest.raises(ValueError):
        lp.build_tableau(A, b, c, np.array([2, -2], np.int32), R=1)
    T, basis, na = lp.build_tableau(A, b, c, np.array([-2, -2], np.int32), R=1)
    assert na == 0 and basis.tolist() == [3, 4] and T.shape == (3, 5)


@pytest.mark.parametrize("m,ns", [(3, 2), (32, 32), (17, 40)])
def test_synth_host_matches_oracle_bitwise(m, ns):
    To, bo = O.synth_feasible(1234, 9, m, ns)
    Tp, bp = lp.synth_host(1234, 9, m, ns)
    assert np.array_equal(To, Tp) and np.array_equal(bo, bp)


def test_header_constants_match_the_ctypes_mirror():
    """Every #define of include/lpgpu.h that capi.py restates (statuses, methods, error codes, option bits) has the same value."""
    import re
    hdr = open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "include", "lpgpu.h")).read()
    defs = {k: int(v, 0) for k, v in re.findall(r"#define\s+LPGPU_(\w+)\s+\(?(-?(?:0x[0-9a-fA-F]+|\d+))\)?", hdr)}
    for name in ("OK", "EINVAL", "ENODEV", "ECUDA", "ENOMEM", "ENCCL", "OPTIMAL", "UNBOUNDED", "INFEASIBLE", "ITER_LIMIT",
                 "RUNNING", "PLAIN", "BIGM", "TWOPHASE", "DUAL"):
        assert defs[name] == getattr(capi, name), name
    # every option bit of lpgpu_opts.variant is a #define of the header; capi.py reads them from there
    for cname, hname in (("BLAND", "VARIANT_BLAND"), ("DEVEX", "VARIANT_DEVEX"), ("RANK1", "VARIANT_RANK1"), ("PIPELINED", "VARIANT_PIPELINED"),
                         ("INPLACE", "VARIANT_INPLACE"), ("DMMA_PASS", "VARIANT_DMMA_PASS"), ("DFMA_PASS", "VARIANT_DFMA_PASS"), ("STEP_PERSISTENT", "VARIANT_STEP_PERSISTENT"), ("STEP_CL8", "VARIANT_STEP_CL8"),
                         ("NO_PDL", "VARIANT_NO_PDL"), ("NO_ALTERNATE", "VARIANT_NO_ALTERNATE"), ("STATIC", "VARIANT_UPD_STATIC"),
                         ("DIRECT", "VARIANT_UPD_DIRECT"), ("KEEP_FINAL", "VARIANT_BATCH_KEEP_FINAL"), ("FORCE_SMEM", "VARIANT_BATCH_SMEM")):
        assert defs[hname] == getattr(capi, cname), cname
    bits = [defs[h] for h in defs if h.startswith("VARIANT_") and not h.endswith("_MASK")]
    assert len(set(bits)) == len(bits), "two option bits share a value"
    assert capi.BLOCK_K(5) & defs["VARIANT_BLOCK_K_MASK"] == 5 << 16
    # no magic variant numbers left in the product's host code
    for f in ("dense.cu", "batched.cu", "shard.cu"):
        text = open(os.path.join(ROOT, "linearprogramming_b200", "csrc", f)).read()
        assert not re.search(r"variant\s*&\s*\(?0x", text), f
    assert defs["NCCL_ID_BYTES"] == capi.NCCL_ID_BYTES and defs["IPC_HANDLE_BYTES"] == capi.IPC_HANDLE_BYTES
