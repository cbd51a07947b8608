// Coalesced movement of one warp's 32 rows between a row-major [rows][256] global tensor (operand dtype) and the warp's
// own rows of an x tile in shared memory (4 K chunks of [128 rows][64 channels], 128-byte rows, 128B swizzle = the UMMA
// K-major operand layout).
//
// The fused epilogues own one accumulator row per lane (tcgen05.ld 32x32b).  Accessing global memory in that layout makes
// every 16-byte access of a warp touch 32 different lines; measured in resqkv_fused.cu, those row-per-lane loads and stores
// were the epilogue's bottleneck (~1.5-3 cycles per lane-line in the LSU, 24 000 cycles for 128 x 256 values).  Staged
// through the x tile, a warp instruction moves 4 rows x 128 contiguous bytes instead.
#pragma once
#include <stdint.h>

namespace mb {

// 16-byte unit `u` (0..7) of row `r` inside one 16 KB K chunk
__device__ __forceinline__ uint32_t xtile_off(int r, int u) {
  return static_cast<uint32_t>(r * 128 + ((u ^ (r & 7)) << 4));
}

// global rows [row0 + r0, +32) x channels [128 half, +128)  ->  x tile rows [r0, +32), chunks 2 half, 2 half + 1
// NC: read through the non-coherent path (data not written by this kernel)
template <typename T, bool NC = true>
__device__ __forceinline__ void warp_rows_g2s(uint8_t* x_tile, const T* src, long row0, long rows, int r0, int half, int lane) {
#pragma unroll
  for (int c = 0; c < 2; ++c) {
    uint4 v[8];
#pragma unroll
    for (int it = 0; it < 8; ++it) {
      const int r = r0 + it * 4 + (lane >> 3), u = lane & 7;
      v[it] = make_uint4(0, 0, 0, 0);
      if (row0 + r < rows) {
        const uint4* p = reinterpret_cast<const uint4*>(src + (row0 + r) * 256 + half * 128 + c * 64) + u;
        v[it] = NC ? __ldg(p) : *p;
      }
    }
#pragma unroll
    for (int it = 0; it < 8; ++it) {
      const int r = r0 + it * 4 + (lane >> 3), u = lane & 7;
      *reinterpret_cast<uint4*>(x_tile + (half * 2 + c) * 16384 + xtile_off(r, u)) = v[it];
    }
  }
}

// the inverse: x tile rows [r0, +32) of this half -> global
template <typename T>
__device__ __forceinline__ void warp_rows_s2g(const uint8_t* x_tile, T* dst, long row0, long rows, int r0, int half, int lane) {
#pragma unroll
  for (int c = 0; c < 2; ++c) {
#pragma unroll
    for (int it = 0; it < 8; ++it) {
      const int r = r0 + it * 4 + (lane >> 3), u = lane & 7;
      const uint4 v = *reinterpret_cast<const uint4*>(x_tile + (half * 2 + c) * 16384 + xtile_off(r, u));
      if (row0 + r < rows) *(reinterpret_cast<uint4*>(dst + (row0 + r) * 256 + half * 128 + c * 64) + u) = v;
    }
  }
}

}  // namespace mb
