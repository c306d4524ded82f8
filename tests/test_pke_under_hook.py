"""The reference's own scheme layer (src/pke: BFVrns with all four multiplication techniques, BGVrns, CKKSrns; key
generation, encryption, EvalAdd / EvalSub / EvalMult with relinearisation, ModReduce / Rescale, EvalRotate, decryption)
on top of the hooked DCRTPoly: oracle/_ref/pke-driver-hooked (oracle/pke_driver.cpp, `make -C oracle pke`) is run twice
with the reference's deterministic PRNG (-DFIXED_SEED, one thread) — once with FHE_B200_ATTACH=1, every DCRTPoly on the
GPU, once without, the reference's CPU statements — and every ciphertext digest and every decrypted value must agree.
Only where the binary was built (the container that has /root/reference); it travels to the GPU box."""
import os
import re
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
DRIVER = os.path.join(ROOT, "oracle", "_ref", "pke-driver-hooked")
pytestmark = pytest.mark.skipif(not os.path.exists(DRIVER), reason="oracle/_ref/pke-driver-hooked not built here")


def run(attach, ring_dim, scenario=""):
    env = dict(os.environ, OMP_NUM_THREADS="1", FHE_B200_ATTACH="1" if attach else "0")
    args = [DRIVER, str(ring_dim)] + ([scenario] if scenario else [])
    r = subprocess.run(args, env=env, capture_output=True, text=True, timeout=900)
    assert r.returncode == 0, r.stderr[-2000:]
    launches = int(re.search(r"total kernel launches (\d+)", r.stderr).group(1))
    return r.stdout.splitlines(), launches, r.stderr


def test_cpu_statements_of_the_hooked_build_decrypt_correctly():
    """Without the handle the hooked build is the reference's CPU path: the schemes decrypt what they should."""
    out, launches, _ = run(False, 2048, "bfv-behz")
    assert launches == 0
    lines = dict(l.split(" pt ", 1) for l in out if " pt " in l)
    assert lines["bfv-behz.dec.add"].split() == [str(a + b) for a, b in zip([1, 2, 3, 4, 5, 6, 7, 8, 9, 10, 11, 12], [3, 2, 1, 4, 5, 6, 7, 8, 9, 10, 11, 12])]
    assert lines["bfv-behz.dec.mul"].split()[:4] == ["3", "4", "3", "16"]


@pytest.mark.gpu
@pytest.mark.parametrize("ring_dim", [4096, 16384])
def test_schemes_on_the_gpu_equal_the_cpu_path(ring_dim):
    cpu, l0, _ = run(False, ring_dim)
    gpu, l1, err = run(True, ring_dim)
    assert l0 == 0 and l1 > 1000, f"the attached run must launch kernels ({l1})"
    assert len(cpu) == len(gpu) > 90
    for a, b in zip(cpu, gpu):
        assert a == b, f"CPU path: {a}\nGPU path: {b}\n{err[-1500:]}"


@pytest.mark.gpu
@pytest.mark.parametrize("example", ["simple-ckks-bootstrapping", "simple-integers"])
def test_reference_examples_unmodified(example):
    """The reference's own src/pke/examples/<example>.cpp, source untouched (`make -C oracle example-<example>`), with
    every DCRTPoly on the GPU (FHE_B200_ATTACH=1) prints exactly what it prints on the CPU path — CKKS bootstrapping
    included (linear transforms with hoisted rotations, Chebyshev evaluation, rescaling, level reduction)."""
    binary = os.path.join(ROOT, "oracle", "_ref", "example-" + example)
    if not os.path.exists(binary):
        pytest.skip("example binary not built here")
    outs = []
    for attach in ("0", "1"):
        env = dict(os.environ, OMP_NUM_THREADS="1", FHE_B200_ATTACH=attach, FHE_B200_ATTACH_REPORT="1")
        r = subprocess.run([binary], env=env, capture_output=True, text=True, timeout=900)
        assert r.returncode == 0, r.stderr[-2000:]
        outs.append(r)
    assert outs[0].stdout == outs[1].stdout and len(outs[0].stdout) > 50
    m = re.search(r"fhe_b200: (\d+) kernel launches", outs[1].stderr)
    assert m and int(m.group(1)) > 20, outs[1].stderr[-500:]
    assert "kernel launches" not in outs[0].stderr
