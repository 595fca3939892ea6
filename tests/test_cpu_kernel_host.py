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


# ---------------------------------------------------------------------------------------------------------------------------------
# csrc/common/tiling.cuh: the tile schedules of the persistent GEMM / convolution kernels, same source as the device code
TILING_HARNESS = r"""
#include "common/tiling.cuh"
using namespace hyp::tiling;
extern "C" void gemm_tile(int tile, int tiles_m, int tiles_n, int nkb, int splits, int* out) {
  const TileCoord c = decode_gemm_tile<8>(tile, tiles_m, tiles_n, nkb, splits);
  out[0] = c.m_blk; out[1] = c.n_blk; out[2] = c.kb0; out[3] = c.kb1;
}
extern "C" void conv_tile(int t, int tiles_k, int tiles_w, int tiles_h, int* out) {
  const ConvTile c = decode_conv_tile(t, tiles_k, tiles_w, tiles_h);
  out[0] = c.kb; out[1] = c.pw; out[2] = c.ph; out[3] = c.pn;
}
extern "C" void wgrad_item(int t, int splits, int cblocks, int taps, int total_pb, int pb_per_split, int* out) {
  const WgradItem w = decode_wgrad_item(t, splits, cblocks, taps, total_pb, pb_per_split);
  out[0] = w.kb; out[1] = w.tap; out[2] = w.cb; out[3] = w.pb0; out[4] = w.pb1;
}
"""


@pytest.fixture(scope="module")
def tiling_lib(tmp_path_factory):
    gxx = shutil.which("g++")
    if gxx is None:
        pytest.skip("no g++ on this machine")
    d = tmp_path_factory.mktemp("tilinglib")
    (d / "harness.cpp").write_text(TILING_HARNESS)
    so = d / "libtiling_host.so"
    subprocess.run([gxx, "-O2", "-std=c++17", "-shared", "-fPIC", "-I", str(ROOT / "csrc"), str(d / "harness.cpp"), "-o", str(so)], check=True)
    return ctypes.CDLL(str(so))


@pytest.mark.parametrize("tiles_m,tiles_n,nkb,splits", [(1, 1, 1, 1), (8, 4, 16, 1), (13, 5, 7, 1), (3, 197, 12, 1), (2, 2, 64, 4),
                                                          (1, 3, 10, 3), (17, 1, 5, 8), (32, 32, 32, 1)])
def test_gemm_schedule_covers_every_tile_and_k_block_once(tiling_lib, tiles_m, tiles_n, nkb, splits):
    out = (ctypes.c_int * 4)()
    seen = np.zeros((tiles_m, tiles_n, nkb), dtype=np.int32)
    order = []
    for t in range(tiles_m * tiles_n * splits):
        tiling_lib.gemm_tile(t, tiles_m, tiles_n, nkb, splits, out)
        m, n, k0, k1 = list(out)
        assert 0 <= m < tiles_m and 0 <= n < tiles_n and 0 <= k0 and k1 <= nkb
        if k1 > k0:                                   # a trailing split may be empty when splits does not divide nkb
            seen[m, n, k0:k1] += 1
        order.append((m, n))
    assert (seen == 1).all()
    # L2 grouping: within the first split, any 8 consecutive tiles of a full group share one n block (one B panel)
    if tiles_m >= 8 and splits == 1:
        first_group = order[:8 * tiles_n]
        for i in range(0, len(first_group), 8):
            assert len({n for _, n in first_group[i:i + 8]}) == 1 and sorted(m for m, _ in first_group[i:i + 8]) == list(range(8))


@pytest.mark.parametrize("tiles_k,tiles_w,tiles_h,tiles_n", [(1, 1, 1, 1), (4, 7, 7, 2), (2, 3, 5, 4), (16, 1, 2, 1)])
def test_conv_forward_schedule(tiling_lib, tiles_k, tiles_w, tiles_h, tiles_n):
    out = (ctypes.c_int * 4)()
    seen = set()
    prev = None
    for t in range(tiles_k * tiles_w * tiles_h * tiles_n):
        tiling_lib.conv_tile(t, tiles_k, tiles_w, tiles_h, out)
        kb, pw, ph, pn = list(out)
        assert 0 <= kb < tiles_k and 0 <= pw < tiles_w and 0 <= ph < tiles_h and 0 <= pn < tiles_n
        seen.add((kb, pw, ph, pn))
        if prev is not None and kb != 0:              # the output-channel blocks of one patch are consecutive (input patch stays in L2)
            assert prev[1:] == (pw, ph, pn) and prev[0] == kb - 1
        prev = (kb, pw, ph, pn)
    assert len(seen) == tiles_k * tiles_w * tiles_h * tiles_n


@pytest.mark.parametrize("kblocks,taps,cblocks,total_pb,splits", [(1, 1, 1, 1, 1), (2, 9, 2, 49, 4), (4, 9, 1, 10, 3), (1, 1, 4, 7, 7), (2, 49, 1, 5, 8)])
def test_conv_wgrad_schedule(tiling_lib, kblocks, taps, cblocks, total_pb, splits):
    per = (total_pb + splits - 1) // splits
    out = (ctypes.c_int * 5)()
    seen = np.zeros((kblocks, taps, cblocks, total_pb), dtype=np.int32)
    for t in range(kblocks * taps * cblocks * splits):
        tiling_lib.wgrad_item(t, splits, cblocks, taps, total_pb, per, out)
        kb, tap, cb, pb0, pb1 = list(out)
        assert 0 <= kb < kblocks and 0 <= tap < taps and 0 <= cb < cblocks and pb1 <= total_pb
        if pb1 > pb0:
            seen[kb, tap, cb, pb0:pb1] += 1
    assert (seen == 1).all()                          # every (filter block, tap, channel block) sees every patch exactly once


# ---------------------------------------------------------------------------------------------------------------------------------
# csrc/common/desc.cuh: tcgen05 instruction / shared-memory descriptors, same source as the device code
DESC_HARNESS = r"""
#include "common/desc.cuh"
extern "C" unsigned idesc(unsigned fmt, unsigned m, unsigned n, int a_mn, int b_mn) { return hyp::make_idesc(fmt, m, n, a_mn != 0, b_mn != 0); }
extern "C" unsigned long long smem_desc(unsigned addr, unsigned lbo, unsigned sbo) { return hyp::make_smem_desc_sw128(addr, lbo, sbo); }
#ifdef WITH_CUTLASS
#include <cute/arch/mma_sm100_desc.hpp>
extern "C" unsigned cutlass_idesc(unsigned fmt, unsigned m, unsigned n, int a_mn, int b_mn) {
  cute::UMMA::InstrDescriptor d = {};
  d.a_format_ = fmt; d.b_format_ = fmt; d.c_format_ = 1;          // F32 accumulator
  d.a_major_ = a_mn ? 1 : 0; d.b_major_ = b_mn ? 1 : 0;
  d.n_dim_ = n >> 3; d.m_dim_ = m >> 4;
  return d.desc_;
}
extern "C" unsigned long long cutlass_smem_desc(unsigned addr, unsigned lbo, unsigned sbo) {
  cute::UMMA::SmemDescriptor d;
  d.desc_ = 0;
  d.start_address_ = (addr & 0x3FFFFu) >> 4; d.leading_byte_offset_ = lbo >> 4; d.stride_byte_offset_ = sbo >> 4;
  d.version_ = 1; d.layout_type_ = 2;                              // Blackwell, SWIZZLE_128B
  return d.desc_;
}
extern "C" int has_cutlass() { return 1; }
#else
extern "C" int has_cutlass() { return 0; }
#endif
"""


def _cutlass_include():
    try:
        import flashinfer
    except Exception:  # noqa: BLE001
        return None
    inc = Path(flashinfer.__file__).parent / "data/cutlass/include"
    return inc if (inc / "cute/arch/mma_sm100_desc.hpp").exists() else None


@pytest.fixture(scope="module")
def desc_lib(tmp_path_factory):
    gxx = shutil.which("g++")
    if gxx is None:
        pytest.skip("no g++ on this machine")
    d = tmp_path_factory.mktemp("desclib")
    (d / "harness.cpp").write_text(DESC_HARNESS)
    base = [gxx, "-O1", "-std=c++17", "-shared", "-fPIC", "-I", str(ROOT / "csrc"), str(d / "harness.cpp")]
    inc = _cutlass_include()
    if inc is not None:
        so = d / "libdesc_cutlass.so"
        r = subprocess.run(base + ["-DWITH_CUTLASS", "-I", str(inc), "-I", "/usr/local/cuda/include", "-o", str(so)], capture_output=True, text=True)
        if r.returncode == 0:
            return ctypes.CDLL(str(so))
    so = d / "libdesc.so"
    subprocess.run(base + ["-o", str(so)], check=True)
    return ctypes.CDLL(str(so))


def test_instruction_descriptor_fields(desc_lib):
    desc_lib.idesc.restype = ctypes.c_uint32
    combos = [(fmt, m, n, a, b) for fmt in (0, 1, 2) for m in (64, 128, 256) for n in (8, 64, 128, 256) for a in (0, 1) for b in (0, 1)]
    for fmt, m, n, a, b in combos:
        v = desc_lib.idesc(fmt, m, n, a, b)
        assert (v >> 4) & 0x3 == 1 and v & 0xF == 0                                    # fp32 accumulator, dense, no saturate
        assert (v >> 7) & 0x7 == fmt and (v >> 10) & 0x7 == fmt                        # A / B format: 0 f16, 1 bf16, 2 tf32
        assert (v >> 13) & 0x3 == 0                                                    # no negation
        assert (v >> 15) & 1 == a and (v >> 16) & 1 == b                               # 0 = K-major, 1 = MN-major
        assert (v >> 17) & 0x3F == n >> 3 and (v >> 23) & 1 == 0
        assert (v >> 24) & 0x1F == m >> 4 and v >> 29 == 0
    if desc_lib.has_cutlass():
        desc_lib.cutlass_idesc.restype = ctypes.c_uint32
        for c in combos:
            assert desc_lib.idesc(*c) == desc_lib.cutlass_idesc(*c), c


def test_shared_memory_descriptor_fields(desc_lib):
    desc_lib.smem_desc.restype = ctypes.c_uint64
    cases = [(0x0, 16, 1024), (0x400, 16, 1024), (0x23C00, 0, 1024), (0x3FFF0, 8192, 16), (0x8000, 2048, 512)]
    for addr, lbo, sbo in cases:
        v = desc_lib.smem_desc(addr, lbo, sbo)
        assert v & 0x3FFF == addr >> 4 and (v >> 14) & 0x3 == 0
        assert (v >> 16) & 0x3FFF == lbo >> 4 and (v >> 32) & 0x3FFF == sbo >> 4
        assert (v >> 46) & 0x3 == 1                                                     # descriptor version: Blackwell
        assert (v >> 48) & 0x1FFF == 0 and v >> 61 == 2                                 # base offset 0, SWIZZLE_128B
    if desc_lib.has_cutlass():
        desc_lib.cutlass_smem_desc.restype = ctypes.c_uint64
        for c in cases:
            assert desc_lib.smem_desc(*c) == desc_lib.cutlass_smem_desc(*c), c
    # advancing the start address by k * 32 bytes (one UMMA_K step of 16-bit elements inside a 128-byte swizzle row) only touches the address field
    a, b = desc_lib.smem_desc(0x400, 16, 1024), desc_lib.smem_desc(0x400 + 64, 16, 1024)
    assert b - a == 64 >> 4
