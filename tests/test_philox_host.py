"""The PRODUCT'S dropout generator on the host: fleetx_b200/csrc/philox.cuh is `__host__ __device__`, so the
very source the kernels compile is built here into a small host library (nvcc, no GPU code, no GPU needed) and
compared with the oracle's numpy restatement and with Random123's published known-answer vectors. The GPU tests
compare the masks the kernels write with the oracle bit for bit; this closes the same loop on a CPU-only box."""
import ctypes as C
import os
import shutil
import subprocess

import numpy as np
import pytest

from oracle import philox

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SRC = r"""
#include <stdint.h>
#include "philox.cuh"
extern "C" void words(uint64_t seed, uint64_t offset, uint64_t group, uint32_t* out) {
  const fx::Philox8 r = fx::philox4x32_10(seed, offset, group);
  for (int i = 0; i < 4; ++i) out[i] = r.lanes[i];
}
extern "C" uint32_t threshold(float p) { return fx::dropout_threshold(p); }
extern "C" void keep8(uint64_t seed, uint64_t offset, uint64_t group, float p, uint8_t* out) {
  const fx::Philox8 r = fx::philox4x32_10(seed, offset, group);
  const uint32_t t = fx::dropout_threshold(p);
  for (int j = 0; j < 8; ++j) out[j] = fx::philox_keep(r, j, t);
}
// the packed two-lanes-at-once form the attention kernels use: bits 15 / 31 of the result
extern "C" void keep8_packed(uint64_t seed, uint64_t offset, uint64_t group, float p, uint8_t* out) {
  const fx::Philox8 r = fx::philox4x32_10(seed, offset, group);
  const uint32_t t = fx::dropout_threshold(p), t2 = t | (t << 16);
  for (int i = 0; i < 4; ++i) {
    const uint32_t f = fx::keep_flags2(r.lanes[i], t2);
    out[2 * i] = (f >> 15) & 1u;
    out[2 * i + 1] = (f >> 31) & 1u;
  }
}
extern "C" void round_keys(uint64_t seed, uint32_t* out) {
  const fx::PhiloxKeys k = fx::philox_round_keys(seed);
  for (int i = 0; i < 20; ++i) out[i] = k.rk[i];
}
"""


@pytest.fixture(scope="module")
def host_lib(tmp_path_factory):
    nvcc = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    if not os.path.exists(nvcc):
        pytest.skip("nvcc not found")
    d = tmp_path_factory.mktemp("philox_host")
    src = d / "philox_host.cu"
    src.write_text(SRC)
    so = d / "libphilox_host.so"
    subprocess.run([nvcc, "-std=c++17", "-O2", "-shared", "-Xcompiler", "-fPIC", "-I", os.path.join(ROOT, "fleetx_b200", "csrc"),
                    "-o", str(so), str(src)], check=True, capture_output=True)
    lib = C.CDLL(str(so))
    lib.words.argtypes = [C.c_uint64, C.c_uint64, C.c_uint64, C.POINTER(C.c_uint32)]
    lib.threshold.argtypes, lib.threshold.restype = [C.c_float], C.c_uint32
    lib.keep8.argtypes = lib.keep8_packed.argtypes = [C.c_uint64, C.c_uint64, C.c_uint64, C.c_float, C.POINTER(C.c_uint8)]
    lib.round_keys.argtypes = [C.c_uint64, C.POINTER(C.c_uint32)]
    return lib


def _words(lib, seed, offset, group):
    out = (C.c_uint32 * 4)()
    lib.words(seed, offset, group, out)
    return list(out)


def test_product_philox_matches_random123_known_answers(host_lib):
    # Random123 kat_vectors, philox4x32 10 rounds: counter words (c0..c3), key (k0, k1) -> output.
    # This library maps (group, offset, seed) to counter = (group lo, group hi, offset lo, offset hi), key = seed.
    kat = [((0, 0, 0, 0), (0, 0), (0x6627e8d5, 0xe169c58d, 0xbc57ac4c, 0x9b00dbd8)),
           ((0xffffffff,) * 4, (0xffffffff, 0xffffffff), (0x408f276d, 0x41c83b0e, 0xa20bc7c6, 0x6d5451fd)),
           ((0x243f6a88, 0x85a308d3, 0x13198a2e, 0x03707344), (0xa4093822, 0x299f31d0),
            (0xd16cfe09, 0x94fdcceb, 0x5001e420, 0x24126ea1))]
    for (c0, c1, c2, c3), (k0, k1), want in kat:
        got = _words(host_lib, k0 | (k1 << 32), c2 | (c3 << 32), c0 | (c1 << 32))
        assert got == list(want)


def test_product_philox_and_keep_rules_match_the_oracle(host_lib):
    rng = np.random.default_rng(7)
    for _ in range(200):
        seed = int(rng.integers(0, 2 ** 63))
        offset = int(rng.integers(0, 2 ** 40))
        group = int(rng.integers(0, 2 ** 45))
        want = philox.philox4x32_10(seed, offset, np.array([group], dtype=np.uint64))[0]
        assert _words(host_lib, seed, offset, group) == [int(w) for w in want]
    for p in (0.0, 0.1, 0.25, 0.5, 0.9, 0.99999):
        assert host_lib.threshold(p) == philox.dropout_threshold(p)
    out, outp = (C.c_uint8 * 8)(), (C.c_uint8 * 8)()
    for p in (0.1, 0.25, 0.5):
        want = philox.dropout_mask(1147, 3, 8 * 64, p).reshape(64, 8)
        for g in range(64):
            host_lib.keep8(1147, 3, g, p, out)
            host_lib.keep8_packed(1147, 3, g, p, outp)
            assert list(out) == want[g].tolist() == list(outp)


def test_precomputed_round_keys_are_the_key_schedule(host_lib):
    out = (C.c_uint32 * 20)()
    seed = 0x0123456789ABCDEF
    host_lib.round_keys(seed, out)
    k0, k1 = seed & 0xFFFFFFFF, seed >> 32
    for r in range(10):
        assert (out[2 * r], out[2 * r + 1]) == ((k0 + r * 0x9E3779B9) & 0xFFFFFFFF, (k1 + r * 0xBB67AE85) & 0xFFFFFFFF)
