"""Device headers compiled for the HOST (g++, a dozen stand-in definitions) so that kernel arithmetic which does not need a GPU is checked
on every CPU run: the Philox4x32-10 generator of ``csrc/common/philox.cuh`` against the published known-answer vectors and against the
numpy oracle the GPU dropout tests use (``ops.rng.reference_keep_mask``).  If the header and the oracle ever drift apart, this fails here
instead of on the GPU box."""
import ctypes
import shutil
import subprocess
from pathlib import Path

import numpy as np
import pytest

ROOT = Path(__file__).resolve().parent.parent

HARNESS = r"""
#include <cstddef>
#include <cstdint>
#define __device__
#define __forceinline__ inline
struct uint2 { uint32_t x, y; };
struct uint4 { uint32_t x, y, z, w; };
static inline uint2 make_uint2(uint32_t x, uint32_t y) { return uint2{x, y}; }
static inline uint4 make_uint4(uint32_t x, uint32_t y, uint32_t z, uint32_t w) { return uint4{x, y, z, w}; }
static inline uint32_t __umulhi(uint32_t a, uint32_t b) { return (uint32_t)(((uint64_t)a * (uint64_t)b) >> 32); }
#include "common/philox.cuh"

extern "C" void philox(const uint32_t* ctr, const uint32_t* key, uint32_t* out) {
  uint4 r = hyp::philox4x32_10(make_uint4(ctr[0], ctr[1], ctr[2], ctr[3]), make_uint2(key[0], key[1]));
  out[0] = r.x; out[1] = r.y; out[2] = r.z; out[3] = r.w;
}
// keep[rows, cols] as bytes, built exactly the way the kernels do: one dropout_keep32 per 32 columns
extern "C" void keep_mask(const unsigned long long* state, uint32_t salt, uint32_t plane, uint32_t thresh, int rows, int cols,
                          uint8_t* keep) {
  const uint2 key = hyp::dropout_key(state);
  for (int r = 0; r < rows; ++r)
    for (int c0 = 0; c0 < cols; c0 += 32) {
      const uint32_t m = hyp::dropout_keep32(key, salt, (uint32_t)c0, (uint32_t)r, plane, thresh);
      for (int i = 0; i < 32 && c0 + i < cols; ++i) keep[(std::size_t)r * cols + c0 + i] = (m >> i) & 1u;
    }
}
extern "C" uint32_t plane_2d() { return hyp::kDropoutPlane2D; }
"""


@pytest.fixture(scope="module")
def host_lib(tmp_path_factory):
    gxx = shutil.which("g++")
    if gxx is None:
        pytest.skip("no g++ on this machine")
    d = tmp_path_factory.mktemp("hostlib")
    (d / "harness.cpp").write_text(HARNESS)
    so = d / "libphilox_host.so"
    subprocess.run([gxx, "-O2", "-std=c++17", "-shared", "-fPIC", "-I", str(ROOT / "csrc"), str(d / "harness.cpp"), "-o", str(so)], check=True)
    return ctypes.CDLL(str(so))


def _philox(lib, ctr, key):
    c, k, o = (ctypes.c_uint32 * 4)(*ctr), (ctypes.c_uint32 * 2)(*key), (ctypes.c_uint32 * 4)()
    lib.philox(c, k, o)
    return list(o)


def test_philox_known_answer_vectors(host_lib):
    # Random123 kat_vectors, philox4x32 with 10 rounds
    assert _philox(host_lib, [0, 0, 0, 0], [0, 0]) == [0x6627E8D5, 0xE169C58D, 0xBC57AC4C, 0x9B00DBD8]
    assert _philox(host_lib, [0xFFFFFFFF] * 4, [0xFFFFFFFF] * 2) == [0x408F276D, 0x41C83B0E, 0xA20BC7C6, 0x6D5451FD]
    assert _philox(host_lib, [0x243F6A88, 0x85A308D3, 0x13198A2E, 0x03707344], [0xA4093822, 0x299F31D0]) == \
        [0xD16CFE09, 0x94FDCCEB, 0x5001E420, 0x24126EA1]


@pytest.mark.parametrize("rows,cols,p,plane", [(7, 64, 0.1, None), (5, 100, 0.25, None), (3, 33, 0.5, 11), (2, 16, 0.9, 0)])
def test_device_header_matches_the_numpy_oracle(host_lib, rows, cols, p, plane):
    from hyperion_b200.ops.rng import reference_keep_mask

    host_lib.plane_2d.restype = ctypes.c_uint32
    plane = host_lib.plane_2d() if plane is None else plane
    assert host_lib.plane_2d() == 0x80000000
    seed, step, salt = 0x1234_5678_9ABC, 41, 7           # the key is the low 32 bits of seed and step
    want, scale = reference_keep_mask(seed, step, salt, plane, rows, cols, p)
    thresh = min(255, max(1, int(p * 256.0 + 0.5)))
    assert abs(scale - 256.0 / (256 - thresh)) < 1e-12
    state = (ctypes.c_ulonglong * 2)(seed, step)
    got = np.zeros((rows, cols), dtype=np.uint8)
    host_lib.keep_mask(state, ctypes.c_uint32(salt), ctypes.c_uint32(plane), ctypes.c_uint32(thresh), rows, cols,
                       got.ctypes.data_as(ctypes.POINTER(ctypes.c_uint8)))
    assert np.array_equal(got.astype(bool), want)
    # and it is a dropout mask: keep rate within 5 sigma of 1 - thresh/256 on a larger draw
    big, _ = reference_keep_mask(seed, step, salt + 1, plane, 256, 512, p)
    q = 1.0 - thresh / 256.0
    assert abs(big.mean() - q) < 5 * np.sqrt(q * (1 - q) / big.size) + 1e-9


def test_masks_differ_across_step_salt_and_plane(host_lib):
    from hyperion_b200.ops.rng import reference_keep_mask

    base, _ = reference_keep_mask(3, 0, 0, 0x80000000, 32, 128, 0.5)
    for kw in (dict(step=1), dict(salt=1), dict(plane=5), dict(seed=4)):
        a = dict(seed=3, step=0, salt=0, plane=0x80000000)
        a.update(kw)
        other, _ = reference_keep_mask(a["seed"], a["step"], a["salt"], a["plane"], 32, 128, 0.5)
        assert 0.35 < (other != base).mean() < 0.65, kw
