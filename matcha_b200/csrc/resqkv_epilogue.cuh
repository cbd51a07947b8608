// Epilogue warps of resqkv_fused.cu (E0: res_conv tail + GroupNorm/Mish + LayerNorm1 -> res, x tile; E1: qkv store).
// Arithmetic follows the unfused epilogue (epilogue3.cuh, feature class kFC_GNLN) step by step.
#pragma once
#include "conv.cuh"
#include "epilogue3.cuh"
#include "ptx.cuh"
#include "rowio.cuh"

namespace mb {

// PAIR (CTA pairs, see resqkv_fused.cu): tiles are numbered 2 u + rank and the barriers the issuer waits on (x_full, t_free)
// live in the pair's leader CTA.
template <typename T, bool PAIR>
__device__ __forceinline__ void rq_epilogue(const RQGeom& g, const float* par, uint8_t* x_tile, float2* xch, uint64_t* t0_full,
                                            uint64_t* x_full, uint64_t* t1_full, uint64_t* t_free, uint32_t tmem_base,
                                            int warp, int lane, int nt_cta, int u0, int ustep, int rank) {
  auto arrive_mma = [](uint64_t* bar) { if constexpr (PAIR) mbar_arrive_leader(bar); else mbar_arrive(bar); };
  const int lane_grp = warp & 3, half = (warp - 2) >> 2;
  const int r = lane_grp * 32 + lane;                      // row of the tile held by this lane
  const uint32_t tlane = tmem_base + (static_cast<uint32_t>(lane_grp * 32) << 16);
  auto lds8 = [&](int arr, int col, float* f) {   // 8 consecutive parameters from shared memory (broadcast reads)
    const float4 a = *reinterpret_cast<const float4*>(par + arr * 256 + col), c = *reinterpret_cast<const float4*>(par + arr * 256 + col + 4);
    f[0] = a.x; f[1] = a.y; f[2] = a.z; f[3] = a.w; f[4] = c.x; f[5] = c.y; f[6] = c.z; f[7] = c.w;
  };
  const long long pt0 = clock64();
  const bool probe = g.dbg_out && warp == 2 && lane == 0;
  long long* po = g.dbg_out ? g.dbg_out + blockIdx.x * 16 : nullptr;
  for (int n = 0; n < nt_cta; ++n) {
    const int tile = PAIR ? 2 * (u0 + n * ustep) + rank : u0 + n * ustep;   // a pair's odd tile may lie past the end: all padding
    const long row = static_cast<long>(tile) * 128 + r;
    const bool ok = row < g.rows;
    // ---- E0.  The loops below are deliberately ROLLED: fully unrolled, this epilogue was 4 600 straight-line instructions
    // executed once per tile, and the warps ran at 4.8 cycles per instruction on instruction fetch (ncu source view: samples
    // spread evenly over every line).  One 32-channel GroupNorm group per iteration; the rounded values go to HBM (res) and,
    // un-normalised, into the x tile, where the LayerNorm pass picks them up again (same thread, same addresses).
    // Global traffic is staged through the x tile (rowio.cuh): the block2 conv output arrives there with coalesced loads,
    // the rounded result leaves from there with coalesced stores, and the LayerNorm pass then normalises it in place.
    const long row0 = static_cast<long>(tile) * 128;
    const int r0 = lane_grp * 32;
    warp_rows_g2s<T>(x_tile, reinterpret_cast<const T*>(g.gn_raw), row0, g.rows, r0, half, lane);
    const float m = ok ? __ldg(g.mask + row) : 0.f;
    const int b = ok ? static_cast<int>(row / g.L) : 0;
    const float* st = g.gn_stats + (b * 8 + half * 4) * 2;
    // this row's four GroupNorm groups: finished sums, or the fixed-order sum of the producing conv's per-32-row partial
    // sums (it then ran without its ordered-sum tail).  Fetched here so the loads fly while the accumulator is awaited.
    float gs[4], gq[4];
    if (g.gn_part) {
      const float2* pp = g.gn_part + static_cast<long>(b * 8 + half * 4) * g.gn_slots;
#pragma unroll
      for (int c = 0; c < 4; ++c) { gs[c] = 0.f; gq[c] = 0.f; }
      for (int i = 0; i < g.gn_slots; ++i) {
#pragma unroll
        for (int c = 0; c < 4; ++c) { const float2 v = __ldcg(pp + c * g.gn_slots + i); gs[c] += v.x; gq[c] += v.y; }
      }
    } else {
#pragma unroll
      for (int c = 0; c < 4; ++c) { gs[c] = __ldg(st + 2 * c); gq[c] = __ldg(st + 2 * c + 1); }
    }
    __syncwarp();
    if (probe && n == 0) po[0] = clock64() - pt0;
    mbar_wait(t0_full, n & 1);
    tc_fence_after();
    if (probe && n == 0) po[1] = clock64() - pt0;
    float sm = 0.f, sq = 0.f;
#pragma unroll 1
    for (int cb = 0; cb < 4; ++cb) {
      const int n0 = half * 128 + cb * 32;
      float acc[32];
      tmem_ld_32x32(tlane + n0, acc);
      const float gsum = cb == 0 ? gs[0] : cb == 1 ? gs[1] : cb == 2 ? gs[2] : gs[3];   // (rolled loop: no dynamic register index)
      const float gsq = cb == 0 ? gq[0] : cb == 1 ? gq[1] : cb == 2 ? gq[2] : gq[3];
      const float gmean = gsum * g.gn_inv_count;
      const float grstd = rsqrtf(fmaxf(gsq * g.gn_inv_count - gmean * gmean, 0.f) + 1e-5f);
      uint4* xp[4];
      uint4 cur[4];
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        const int i = cb * 4 + q;                            // 16-byte unit of this half-row
        xp[q] = reinterpret_cast<uint4*>(x_tile + (half * 2 + (i >> 3)) * 16384 + xtile_off(r, i & 7));
        cur[q] = *xp[q];
      }
      tmem_ld_wait();
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        float bv[8], pa[8], pb[8], rv[8], v[8];
        lds8(0, n0 + 8 * q, bv);
        lds8(1, n0 + 8 * q, pa);
        lds8(2, n0 + 8 * q, pb);
        unpack8<T>(cur[q], rv);
#pragma unroll
        for (int ch = 0; ch < 8; ++ch) {
          pa[ch] *= grstd;
          pb[ch] -= gmean * pa[ch];
          v[ch] = acc[8 * q + ch] + bv[ch];
          v[ch] += mish_fast(fmaf(rv[ch], pa[ch], pb[ch])) * m;
          sm += v[ch];
          sq = fmaf(v[ch], v[ch], sq);
        }
        *xp[q] = pack8<T>(v);
      }
    }
    __syncwarp();
    warp_rows_s2g<T>(x_tile, reinterpret_cast<T*>(g.res_out), row0, g.rows, r0, half, lane);
    __syncwarp();
    if (probe && n == 0) po[2] = clock64() - pt0;
    // LayerNorm1 row statistics: the other half of the row belongs to the partner warp (same TMEM lane quarter)
    xch[half * 128 + r] = make_float2(sm, sq);
    named_bar_sync(1 + lane_grp, 64);
    const float2 s0 = xch[r], s1 = xch[128 + r];
    named_bar_sync(1 + lane_grp, 64);                       // both warps have read before the next tile overwrites
    const float mean = (s0.x + s1.x) * (1.f / 256.f);
    const float rstd = rsqrtf(fmaxf((s0.y + s1.y) * (1.f / 256.f) - mean * mean, 0.f) + 1e-5f);
#pragma unroll 2
    for (int i = 0; i < 16; ++i) {
      float ga[8], be[8], xv[8];
      lds8(3, half * 128 + 8 * i, ga);
      lds8(4, half * 128 + 8 * i, be);
      uint32_t off = static_cast<uint32_t>(r * 128 + (i & 7) * 16);
      off ^= ((off >> 7) & 7u) << 4;
      uint4* xp = reinterpret_cast<uint4*>(x_tile + (half * 2 + (i >> 3)) * 16384 + off);
      unpack8<T>(*xp, xv);
#pragma unroll
      for (int ch = 0; ch < 8; ++ch) xv[ch] = fmaf((xv[ch] - mean) * rstd, ga[ch], be[ch]);
      *xp = pack8<T>(xv);
    }
    tc_fence_before();
    fence_proxy_async();          // generic-proxy stores -> visible to the tensor core's async-proxy reads
    __syncwarp();
    if (lane == 0) arrive_mma(x_full);
    if (probe && n == 0) po[3] = clock64() - pt0;
    // ---- E1: q | k | v columns [192 half, +192) of this lane's row
    // q | k | v columns [192 half, +192) of this warp's 32 rows, in two rounds of 96 columns staged through a warp-private
    // slice of the x tile (free once M1 has completed; pitch 208 B keeps the row-per-lane writes conflict-free)
    mbar_wait(t1_full, n & 1);
    tc_fence_after();
    if (probe && n == 0) po[4] = clock64() - pt0;
    uint8_t* stg = x_tile + (warp - 2) * 8192;
#pragma unroll 1
    for (int rd = 0; rd < 2; ++rd) {
#pragma unroll 1
      for (int cb = 0; cb < 3; ++cb) {
        float acc[32];
        tmem_ld_32x32(tlane + half * 192 + rd * 96 + cb * 32, acc);
        tmem_ld_wait();
#pragma unroll
        for (int q = 0; q < 4; ++q) *reinterpret_cast<uint4*>(stg + lane * 208 + (cb * 4 + q) * 16) = pack8<T>(acc + 8 * q);
      }
      __syncwarp();
#pragma unroll
      for (int it = 0; it < 12; ++it) {
        const int idx = it * 32 + lane, rr = idx / 12, u = idx % 12;
        const uint4 v = *reinterpret_cast<const uint4*>(stg + rr * 208 + u * 16);
        const long grow = row0 + r0 + rr;
        if (grow < g.rows) *(reinterpret_cast<uint4*>(reinterpret_cast<T*>(g.qkv_out) + grow * 384 + half * 192 + rd * 96) + u) = v;
      }
      __syncwarp();
    }
    tc_fence_before();
    __syncwarp();
    if (lane == 0) arrive_mma(t_free);
    if (n + 1 < nt_cta) named_bar_sync(5, 256);   // every warp's staging slice is drained before the next tile's rows land in the x tile
    if (probe && n == 0) po[5] = clock64() - pt0;
  }
  if (probe) po[6] = clock64() - pt0;
}

}  // namespace mb
