// tc_tile.cuh -- what the tcgen05 tile kernels (tc_kernels.cu, tc_fwd_sym.cu) share: tile geometry,
// shared-memory descriptor helpers and the host-side tensor-map / launch-attribute caches.
#pragma once
#include <cuda.h>

#include "common.cuh"
#include "tc_ptx.cuh"

namespace ntxent_b200 {
namespace tc {

constexpr int kTile = 128;                  // rows per row tile = columns per column tile
constexpr int kAtomBytes = kTile * 128;     // one 128B-swizzle atom: [128 rows x 64 bf16]
constexpr float kFixedMaxLimit = 50.0f;     // fixed-max logsumexp while log2(e)/tau <= this
// of every 8 exponentials in the fixed-max paths, this many are evaluated by ex2_poly
#ifndef NTX_FWD_POLY
#define NTX_FWD_POLY 3
#endif
#ifndef NTX_BWD_POLY
#define NTX_BWD_POLY 0
#endif
constexpr int kFwdPoly = NTX_FWD_POLY, kBwdPoly = NTX_BWD_POLY;
// the packed (f32x2) forward loops: of every 4 consecutive PAIRS of exponentials, this many take
// the polynomial.  Measured at 2N = 32768 x 128 (profiles/r2_fwd_poly_sweep.txt): 1 of 4 is the
// fastest in both forward kernels.
#ifndef NTX_FWD_POLY_PAIRS
#define NTX_FWD_POLY_PAIRS 1
#endif
constexpr int kFwdPolyPairs = NTX_FWD_POLY_PAIRS;
// the same choice for the backward's packed weight loop (0: every exponential on the MUFU)
#ifndef NTX_BWD_POLY_PAIRS
#define NTX_BWD_POLY_PAIRS 0
#endif
constexpr int kBwdPolyPairs = NTX_BWD_POLY_PAIRS;

// TF32 = false: bf16 operands, .kind::f16 (K = 16 per MMA).  TF32 = true: fp32 operands read as
// tf32, .kind::tf32 (K = 8 per MMA).  Either way one MMA consumes 32 bytes of every operand row
// and a 128-byte swizzle row holds four such slices, so the descriptor arithmetic is shared.
template <int D, bool TF32 = false>
struct Cfg {
    static_assert(D % 64 == 0 && D >= 64 && D <= (TF32 ? 128 : 256),
                  "tcgen05 tiles cover D in {64,128,192,256} (bf16) or {64,128} (tf32)");
    static constexpr int kElemBytes = TF32 ? 4 : 2;
    static constexpr int kAtoms = D * kElemBytes / 128;
    static constexpr int kAtomCols = 128 / kElemBytes;  // elements per 128-byte swizzle row
    static constexpr int kTileBytes = kAtoms * kAtomBytes;
    static constexpr int kKSteps = D * kElemBytes / 32;  // MMAs per S tile
    static constexpr int kKPerMma = 32 / kElemBytes;
    // forward: two CTAs per SM when the tiles allow it
    static constexpr int kFwdStages = (kTileBytes <= 16 * 1024) ? 3 : 2;
    static constexpr int kFwdSmem = (1 + kFwdStages) * kTileBytes + 2048;
    static constexpr int kFwdCtasPerSm = (kFwdSmem <= 112 * 1024) ? 2 : 1;
};

#ifdef __CUDACC__
// K-major operand (rows = M or N index, 128 B of K per row, 8-row groups 1024 B apart).
__device__ __forceinline__ uint64_t desc_kmajor(uint32_t tile_addr, int kstep) {
    // kstep counts 32-byte K slices: 4 per swizzle atom
    const uint32_t addr = tile_addr + (kstep >> 2) * kAtomBytes + (kstep & 3) * 32;
    return ptx::smem_desc_sw128(addr, 16, 1024);
}
// MN-major view of the same tile (K index = tile row): 16 (bf16) or 8 (tf32) K rows per MMA; the
// next 128 bytes of N live one atom (16 KB) further: LBO = atom stride, SBO = 8-row group stride.
template <bool TF32 = false>
__device__ __forceinline__ uint64_t desc_mnmajor(uint32_t tile_addr, int kstep) {
    return ptx::smem_desc_sw128(tile_addr + kstep * (TF32 ? 1024 : 2048), kAtomBytes, 1024);
}
__device__ __forceinline__ uint8_t* align1024(uint8_t* p) {
    return reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(p) + 1023) & ~uintptr_t(1023));
}
#endif

// ---- host side (defined in tc_kernels.cu) ----------------------------------------
// [rows, cols] row-major (bf16 or fp32), box = [box_rows x 128 bytes], 128-byte swizzle, zero fill
// past the edge; the last few (pointer, shape) pairs are remembered.
cudaError_t cached_tmap_2d(const void* base, bool tf32, int64_t rows, int64_t cols, int box_rows,
                           CUtensorMap* out);
// cudaFuncAttributeMaxDynamicSharedMemorySize is per function and per device: set it the first
// time an instantiation is launched on a device, not on every launch.
cudaError_t set_smem_once_impl(const void* kern, int smem, unsigned long long& done_mask);
template <typename K>
inline cudaError_t set_smem_once(K kern, int smem, unsigned long long& done_mask) {
    return set_smem_once_impl(reinterpret_cast<const void*>(kern), smem, done_mask);
}

}  // namespace tc
}  // namespace ntxent_b200
