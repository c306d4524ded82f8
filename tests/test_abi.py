"""The C-ABI shared library loads on a CPU-only box and exports every symbol include/fhe_b200.h
declares; host-only entry points work, device entry points fail loudly (no CPU fallback)."""
import ctypes
import os
import re
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "fhe_b200.h")


def declared_symbols():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(fhe_[a-z0-9_]+)\s*\(", src)))


def test_header_declares_and_library_exports_same_symbols(pkg):
    names = declared_symbols()
    assert len(names) >= 30
    lib = ctypes.CDLL(pkg.LIB_PATH)
    for n in names:
        assert hasattr(lib, n), f"{n} declared in include/fhe_b200.h but not exported"
    assert sorted(pkg.SIGNATURES) == names, "ctypes binding and header disagree"
    out = subprocess.run(["nm", "-D", "--defined-only", pkg.LIB_PATH], capture_output=True, text=True).stdout
    exported = set(re.findall(r" T (fhe_[a-z0-9_]+)$", out, flags=re.M))
    assert exported == set(names), f"undeclared exports: {exported ^ set(names)}"


def test_header_compiles_as_c(tmp_path):
    c = tmp_path / "t.c"
    c.write_text('#include "fhe_b200.h"\n#include "pim/common.h"\n#include "pim/kernel.h"\nint main(void){return FHE_OK;}\n')
    subprocess.check_call(["gcc", "-std=c99", "-Wall", "-Werror", "-I", os.path.join(ROOT, "include"), "-c", str(c),
                           "-o", str(tmp_path / "t.o")])


def test_gen_params_is_host_only_and_matches_reference_values(pkg):
    q, psi = pkg.gen_params(60, 1 << 14, 2)
    assert q.tolist() == [1152921504606748673, 1152921504606683137]
    assert psi.tolist() == [62213374832584, 212089012217363]
    assert pkg._native.lib.fhe_root_of_unity(8, 113) == 18
    with pytest.raises(pkg.FheError):
        pkg.gen_params(61, 1 << 14, 2)
    with pytest.raises(pkg.FheError):
        pkg.gen_params(60, 3000, 2)


def test_gen_params_matches_oracle(pkg, oracle):
    for nbits, n, t in [(14, 4, 3), (28, 8, 2), (50, 1 << 12, 3), (60, 1 << 16, 5), (59, 1 << 15, 4)]:
        q, psi = pkg.gen_params(nbits, n, t)
        qo, po = oracle.gen_params(nbits, n, t)
        assert np.array_equal(q, qo) and np.array_equal(psi, po)


def test_no_cpu_fallback_without_a_gpu(pkg):
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    with pytest.raises(pkg.FheError) as e:
        pkg.Backend(0)
    assert e.value.code == pkg._native.FHE_ENODEV
    assert "no CPU fallback" in str(e.value)


def test_product_does_not_link_the_oracle(pkg):
    out = subprocess.run(["ldd", pkg.LIB_PATH], capture_output=True, text=True).stdout
    assert "oracle" not in out and "openfhe_ref" not in out
    banned = re.compile(r"fhe_oracle|libfhe_oracle|openfhe_ref|\borc_[a-z_]+\s*\(|from oracle|import oracle|oracle_py")
    for root, _, files in os.walk(os.path.join(ROOT, "upmem--openfhe_b200")):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".cpp")):
                assert not banned.search(open(os.path.join(root, f)).read()), f
    sym = subprocess.run(["nm", "-D", pkg.LIB_PATH], capture_output=True, text=True).stdout
    assert " orc_" not in sym and " ref_" not in sym


def test_cpp_surface_builds_and_fails_loudly_without_gpu():
    """include/pim/*.h compile as C++17 against the C ABI; with no GPU the surface throws (no fallback)."""
    import torch
    subprocess.check_call(["make", "-C", ROOT, "host"], stdout=subprocess.DEVNULL)
    exe = os.path.join(ROOT, "tests", "cpp", "bin", "pim_surface_test")
    assert os.path.exists(exe)
    if torch.cuda.is_available():
        pytest.skip("a GPU is present: the binary is run by the gpu tests")
    r = subprocess.run([exe], capture_output=True, text=True, timeout=120)
    assert r.returncode != 0 and "no CPU fallback" in r.stdout


def test_hps_scale_tables_equal_the_reference():
    """Host-side table recipe (no GPU needed): equals the tables the reference's BigInteger code produced
    (tests/golden/ext_vectors.json, written by tests/golden/make_golden_ext.py from the compiled reference)."""
    import importlib
    import json
    import numpy as np
    pkg = importlib.import_module("upmem--openfhe_b200")
    with open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "ext_vectors.json")) as f:
        cases = json.load(f)["bfv"]
    for c in cases:
        sq = c["sizeQ"]
        frac, modr = pkg.hps_scale_tables(c["q"][:sq], c["q"][sq:], c["t"])
        assert [float(v).hex() for v in frac] == c["frac_hex"]
        assert modr.reshape(-1).tolist() == c["modr"]
