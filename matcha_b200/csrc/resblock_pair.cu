// One HiFi-GAN ResBlock1 step as ONE persistent tcgen05 kernel (narrow stages, C = 64 / 32 channels):
//
//     x <- x + c2( lrelu( c1( lrelu(x) ) ) )            matcha/hifigan/models.py:90-97
//
// Run as two conv-GEMM launches the step moves five activation-sized streams through HBM (read lrelu(x), write the
// intermediate, read it back, read the residual, write the result) and stages 3-4 sit on the HBM roofline.  Here the
// intermediate never leaves the SM: per tile of R = 128*MB rows
//     TMA   x halo tile (R + (K-1)*dil rows)                       -> shared memory ring
//     MMA1  c1 over the halo tile (taps = row-shifted descriptors) -> TMEM accumulator ring 1
//     E1    + bias1, LeakyReLU, zero outside [0, T), half          -> shared memory "mid" tile in the UMMA K-major
//                                                                     swizzled layout (the writer swizzles by hand)
//     MMA2  c2 over the mid tile                                   -> TMEM accumulator ring 2
//     E2    the usual fused epilogue (epilogue3.cuh): + bias2 + residual (+ branch sum, scale) -> global
// Tiles advance by S = R - (K-1) output rows: c1 is recomputed on the K-1 halo rows (2-4 %), which is what makes the
// tiles independent.  Two streams per step instead of five.
//
// Warp roles (448 threads): 0 TMA producer, 1 MMA issuer (both convs, software-pipelined M1(n+1) before M2(n)),
// 2-5 E1, 6-13 E2.  Producer and issuer are converged warps with elected single-thread instructions.
#include "conv.cuh"
#include "epilogue3.cuh"
#include "launch.h"
#include "ptx.cuh"

namespace mb {

namespace {
constexpr int kPairBarBytes = 512;
// developer probe (PairGeom::dbg_out, MB200_PAIR_DBG): cycles a role spends inside one barrier wait
#define PW(stmt, acc) do { if (g.dbg_out) { const long long pw_a = clock64(); stmt; acc += clock64() - pw_a; } else { stmt; } } while (0)
}

// PAIR: two CTAs of a cluster as one tcgen05 cta_group::2 unit (conv_gemm2.cu): each CTA runs its own tile, every weight
// tile is split between the two CTAs (C / 2 output channels each), the even CTA issues all MMAs with M = 256.
template <int C, int MB, typename T, bool PAIR>
__global__ void __launch_bounds__(448, 1)
resblock_pair_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmA8,
                     const __grid_constant__ CUtensorMap tmW1, const __grid_constant__ CUtensorMap tmW2,
                     const __grid_constant__ PairGeom g, const __grid_constant__ EpiParams e) {
  constexpr int ROWB = C * 2;             // bytes per activation row = one swizzle span (128 B / 64 B)
  constexpr int W_TILE = (PAIR ? C / 2 : C) * ROWB;   // one tap of one conv: the output channels THIS CTA holds
  constexpr int W_STAGE = 2 * W_TILE;     // streamed weights: two taps per ring stage
  constexpr int SLOT = MB * C;            // TMEM columns of one accumulator buffer
  constexpr int R = 128 * MB;
  constexpr int EG = 2;
  constexpr uint32_t SWZ = ROWB == 128 ? 7u : 3u;
  static_assert(4 * SLOT <= 512, "two double-buffered accumulator rings must fit TMEM");
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);

  const int K = g.K, p = (K - 1) / 2;
  const int n_wst = (K + 1) / 2;          // ring stages per conv when the weights are streamed
  uint8_t* a_ring = smem;
  uint8_t* mid = a_ring + g.a_stages * g.a_bytes;
  uint8_t* w_base = mid + g.mid_stages * g.mid_bytes;
  const int w_bytes = g.w_resident ? 2 * K * W_TILE : g.w_stages * W_STAGE;
  uint8_t* tail = w_base + w_bytes;
  uint64_t* bars = reinterpret_cast<uint64_t*>(tail);
  uint64_t* a_full = bars;            // [4]
  uint64_t* a_empty = a_full + 4;     // [4]
  uint64_t* w_full = a_empty + 4;     // [8]
  uint64_t* w_empty = w_full + 8;     // [8]
  uint64_t* t1_full = w_empty + 8;    // [2]
  uint64_t* t1_empty = t1_full + 2;   // [2]
  uint64_t* t2_full = t1_empty + 2;   // [2]
  uint64_t* t2_empty = t2_full + 2;   // [2]
  uint64_t* mid_full = t2_empty + 2;  // [2]
  uint64_t* mid_empty = mid_full + 2; // [2]
  uint64_t* wres = mid_empty + 2;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(wres + 1);
  uint8_t* stage_base = tail + kPairBarBytes;

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int rank = PAIR ? static_cast<int>(cluster_ctarank()) : 0;
  // work units: tiles, or pairs of tiles (2u, 2u + 1); the odd CTA of the last pair may get a tile past the end (a ghost:
  // its loads are out of range = zero fill, it stores nothing)
  const int units = PAIR ? (g.total_tiles + 1) / 2 : g.total_tiles;
  const int u0 = PAIR ? static_cast<int>(blockIdx.x >> 1) : static_cast<int>(blockIdx.x);
  const int ustep = PAIR ? static_cast<int>(gridDim.x >> 1) : static_cast<int>(gridDim.x);
  const int nt_cta = units > u0 ? (units - 1 - u0) / ustep + 1 : 0;
  auto tile_of = [&](int n) { const int u = u0 + n * ustep; return PAIR ? 2 * u + rank : u; };
  auto arrive_mma = [](uint64_t* bar) { if constexpr (PAIR) mbar_arrive_leader(bar); else mbar_arrive(bar); };   // barriers the issuer waits on
  auto expect = [&](uint64_t* bar, uint32_t bytes) { if constexpr (PAIR) { if (rank == 0) mbar_expect_tx(bar, 2 * bytes); } else mbar_expect_tx(bar, bytes); };
  auto load2 = [](void* dst, const CUtensorMap* m, uint64_t* bar, int c0, int c1) {
    if constexpr (PAIR) tma_load_2d_pair(dst, m, bar, c0, c1); else tma_load_2d(dst, m, bar, c0, c1);
  };
  auto load3 = [](void* dst, const CUtensorMap* m, uint64_t* bar, int c0, int c1, int c2) {
    if constexpr (PAIR) tma_load_3d_pair(dst, m, bar, c0, c1, c2); else tma_load_3d(dst, m, bar, c0, c1, c2);
  };
  const int wrow0 = PAIR ? rank * (C / 2) : 0;   // first output channel of this CTA's weight rows

  if (threadIdx.x == 0) {
    tma_prefetch_desc(&tmA); tma_prefetch_desc(&tmA8); tma_prefetch_desc(&tmW1); tma_prefetch_desc(&tmW2);
    for (int s = 0; s < 4; ++s) { mbar_init(&a_full[s], 1); mbar_init(&a_empty[s], 1); }
    for (int s = 0; s < 8; ++s) { mbar_init(&w_full[s], 1); mbar_init(&w_empty[s], 1); }
    for (int s = 0; s < 2; ++s) {
      mbar_init(&t1_full[s], 1); mbar_init(&t1_empty[s], PAIR ? 8 : 4);
      mbar_init(&t2_full[s], 1); mbar_init(&t2_empty[s], (PAIR ? 8 : 4) * EG);
    }
    for (int s = 0; s < 2; ++s) { mbar_init(&mid_full[s], PAIR ? 8 : 4); mbar_init(&mid_empty[s], 1); }
    mbar_init(wres, 1);
    fence_barrier_init();
  }
  if (warp == 1) { if constexpr (PAIR) tmem_alloc_pair<512>(tmem_slot); else tmem_alloc<512>(tmem_slot); }
  tc_fence_before();
  if constexpr (PAIR) cluster_sync_all(); else __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    // ================================================================= TMA producer
    int as = 0, ws = 0;
    uint32_t aph = 0, wph = 0;
    long long pw[2] = {0, 0};
    const long long pt0 = clock64();
    auto load_A = [&](int n) {
      const int tile = tile_of(n);
      const int b = tile / g.tiles_per_b, t0 = (tile % g.tiles_per_b) * g.S;
      const int row0 = t0 - p - p * g.dil;
      PW(mbar_wait(&a_empty[as], aph ^ 1), pw[0]);
      if (elect_one()) {
        uint8_t* dst = a_ring + as * g.a_bytes;
        expect(&a_full[as], (R + 8 * g.extra) * ROWB);
#pragma unroll
        for (int i = 0; i < MB; ++i) load3(dst + i * 128 * ROWB, &tmA, &a_full[as], 0, row0 + 128 * i, b);
        for (int x = 0; x < g.extra; ++x) load3(dst + (R + 8 * x) * ROWB, &tmA8, &a_full[as], 0, row0 + R + 8 * x, b);
      }
      __syncwarp();
      if (++as == g.a_stages) { as = 0; aph ^= 1; }
    };
    auto load_W = [&](const CUtensorMap* tm) {
      for (int st = 0; st < n_wst; ++st) {
        const int nt = min(2, K - 2 * st);
        PW(mbar_wait(&w_empty[ws], wph ^ 1), pw[1]);
        if (elect_one()) {
          expect(&w_full[ws], nt * W_TILE);
          for (int j = 0; j < nt; ++j) load2(w_base + ws * W_STAGE + j * W_TILE, tm, &w_full[ws], (2 * st + j) * C, wrow0);
        }
        __syncwarp();
        if (++ws == g.w_stages) { ws = 0; wph ^= 1; }
      }
    };
    if (g.w_resident) {
      if (elect_one()) {
        expect(wres, static_cast<uint32_t>(w_bytes));
        for (int j = 0; j < K; ++j) load2(w_base + j * W_TILE, &tmW1, wres, j * C, wrow0);
        for (int j = 0; j < K; ++j) load2(w_base + (K + j) * W_TILE, &tmW2, wres, j * C, wrow0);
      }
      __syncwarp();
      for (int n = 0; n < nt_cta; ++n) load_A(n);
    } else if (nt_cta > 0) {
      // same order as the issuer consumes: M1(0) | M1(n+1), M2(n) ...; the x tile of M1(n+2) is requested as soon as
      // its ring stage can be free (M1(n) done), one whole c2 ahead of its use
      load_A(0);
      if (nt_cta > 1) load_A(1);
      load_W(&tmW1);
      for (int n = 0; n < nt_cta; ++n) {
        if (n + 1 < nt_cta) load_W(&tmW1);
        if (n + 2 < nt_cta) load_A(n + 2);
        load_W(&tmW2);
      }
    }
    if (g.dbg_out && lane == 0) {
      long long* o = g.dbg_out + blockIdx.x * 16;
      o[11] = clock64() - pt0; o[12] = pw[0]; o[13] = pw[1]; o[14] = nt_cta;
    }
  } else if (warp == 1) {
    // ================================================================= MMA issuer (pair: the leader's only)
    if (rank == 0) {
    constexpr uint32_t idesc = umma_idesc_f16(std::is_same<T, __nv_bfloat16>::value ? 1 : 0, PAIR ? 256 : 128, C);
    auto umma_f16_lohi = [](uint32_t d, uint32_t a, uint32_t b, uint32_t hi, uint32_t id, uint32_t acc) {
      if constexpr (PAIR) mb::umma_f16_lohi_pair(d, a, b, hi, id, acc); else mb::umma_f16_lohi(d, a, b, hi, id, acc);
    };
    auto umma_commit = [](uint64_t* bar) { if constexpr (PAIR) mb::umma_commit_pair(bar); else mb::umma_commit(bar); };
    constexpr uint32_t W16 = W_TILE >> 4, BLK16 = (128 * ROWB) >> 4, ROW16 = ROWB >> 4, WST16 = W_STAGE >> 4;
    const uint32_t desc_hi = static_cast<uint32_t>(umma_smem_desc<ROWB>(0) >> 32);
    const uint32_t a_lo0 = static_cast<uint32_t>(umma_smem_desc<ROWB>(smem_u32(a_ring)));
    const uint32_t mid_lo = static_cast<uint32_t>(umma_smem_desc<ROWB>(smem_u32(mid)));
    const uint32_t w_lo0 = static_cast<uint32_t>(umma_smem_desc<ROWB>(smem_u32(w_base)));
    const uint32_t a_stage16 = static_cast<uint32_t>(g.a_bytes) >> 4;
    const int resident = g.w_resident;
    if (resident) { mbar_wait(wres, 0); tc_fence_after(); }
    int as = 0, ws = 0;
    uint32_t aph = 0, wph = 0;
    long long pw[5] = {0, 0, 0, 0, 0};
    const long long pt0 = clock64();
    // all taps of one conv for the MB row blocks of a tile; operand rows advance by `row_step` rows per tap
    auto issue_conv = [&](const uint32_t a_lo, const int row_step, const uint32_t dtm, const int which) {
      // one tap for all MB row blocks; consecutive MMAs go to different accumulators (no back-to-back dependency)
      auto tap_mmas = [&](const int tap, const uint32_t w) {
        const uint32_t a = a_lo + static_cast<uint32_t>(tap * row_step) * ROW16;
#pragma unroll
        for (int k = 0; k < C / 16; ++k) {
#pragma unroll
          for (int i = 0; i < MB; ++i)
            umma_f16_lohi(dtm + i * C, a + i * BLK16 + 2 * k, w + 2 * k, desc_hi, idesc, (tap | k) != 0 ? 1u : 0u);
        }
      };
      if (resident) {   // nothing to wait for: the whole conv is one elected issue block
        tc_fence_after();
        if (elect_one()) {
          const uint32_t w_lo = w_lo0 + static_cast<uint32_t>(which * K) * W16;
#pragma unroll 1
          for (int tap = 0; tap < K; ++tap) tap_mmas(tap, w_lo + tap * W16);
        }
        __syncwarp();
        return;
      }
#pragma unroll 1
      for (int st = 0; st < n_wst; ++st) {
        PW(mbar_wait(&w_full[ws], wph), pw[2]);
        const uint32_t w_lo = w_lo0 + ws * WST16;
        tc_fence_after();
        if (elect_one()) {
          tap_mmas(2 * st, w_lo);
          if (2 * st + 1 < K) tap_mmas(2 * st + 1, w_lo + W16);
          umma_commit(&w_empty[ws]);
        }
        __syncwarp();
        if (++ws == g.w_stages) { ws = 0; wph ^= 1; }
      }
    };
    auto M1 = [&](int n) {
      const int slot = n & 1;
      PW(mbar_wait(&t1_empty[slot], ((n >> 1) & 1) ^ 1), pw[0]);
      PW(mbar_wait(&a_full[as], aph), pw[1]);
      issue_conv(a_lo0 + as * a_stage16, g.dil, tmem_base + slot * SLOT, 0);
      if (elect_one()) { umma_commit(&a_empty[as]); umma_commit(&t1_full[slot]); }
      __syncwarp();
      if (++as == g.a_stages) { as = 0; aph ^= 1; }
    };
    auto M2 = [&](int n) {
      const int slot = n & 1;
      PW(mbar_wait(&t2_empty[slot], ((n >> 1) & 1) ^ 1), pw[3]);
      // mid tiles: ring of g.mid_stages (2 when shared memory allows: E1 of the next tile then overlaps this c2)
      const int ms = g.mid_stages == 2 ? slot : 0;
      const uint32_t mpar = g.mid_stages == 2 ? ((n >> 1) & 1) : (n & 1);
      PW(mbar_wait(&mid_full[ms], mpar), pw[4]);
      issue_conv(mid_lo + ms * (static_cast<uint32_t>(g.mid_bytes) >> 4), 1, tmem_base + (2 + slot) * SLOT, 1);
      if (elect_one()) { umma_commit(&mid_empty[ms]); umma_commit(&t2_full[slot]); }
      __syncwarp();
    };
    if (nt_cta > 0) M1(0);
    for (int n = 0; n < nt_cta; ++n) {
      if (n + 1 < nt_cta) M1(n + 1);
      M2(n);
    }
    if (g.dbg_out && lane == 0) {
      long long* o = g.dbg_out + blockIdx.x * 16;
      o[0] = clock64() - pt0;
      for (int i = 0; i < 5; ++i) o[1 + i] = pw[i];
    }
    }
  } else if (warp < 6) {
    // ================================================================= E1: accumulator 1 -> mid tile (A operand of c2)
    const int lane_grp = warp & 3;
    const float slope = g.slope;
    long long pw[2] = {0, 0};
    const long long pt0 = clock64();
    constexpr int ITEMS = MB * (C / 32);
    for (int n = 0; n < nt_cta; ++n) {
      const int tile = tile_of(n);
      const int t0 = (tile % g.tiles_per_b) * g.S, m0 = t0 - p;
      const int slot = n & 1;
      const int ms = g.mid_stages == 2 ? slot : 0;
      uint8_t* mid_t = mid + ms * g.mid_bytes;
      const uint32_t tbase = tmem_base + (static_cast<uint32_t>(lane_grp * 32) << 16) + slot * SLOT;
      PW(mbar_wait(&t1_full[slot], (n >> 1) & 1), pw[0]);
      tc_fence_after();
      // accumulator loads run one item ahead of the arithmetic: item+1 is in flight while item is converted and stored
      float acc[2][32];
      tmem_ld_32x32(tbase, acc[0]);
      // the mid tile is free once the c2 that read it has completed
      PW(mbar_wait(&mid_empty[ms], (g.mid_stages == 2 ? ((n >> 1) & 1) : (n & 1)) ^ 1), pw[1]);
#pragma unroll
      for (int item = 0; item < ITEMS; ++item) {
        const int i = item / (C / 32), c = item % (C / 32);
        const int r = i * 128 + lane_grp * 32 + lane;   // row of the mid tile held by this lane
        const int m = m0 + r;
        const bool live = (m >= 0) && (m < g.T);        // c2 zero-pads its INPUT: rows outside the utterance are 0
        tmem_ld_wait();
        if (item + 1 < ITEMS) tmem_ld_32x32(tbase + ((item + 1) / (C / 32)) * C + ((item + 1) % (C / 32)) * 32, acc[(item + 1) & 1]);
        const float* a = acc[item & 1];
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          float bv[8], v[8];
          ldg8f(g.bias1 + c * 32 + 8 * q, bv);
#pragma unroll
          for (int ch = 0; ch < 8; ++ch) v[ch] = a[8 * q + ch] + bv[ch];
          uint4 o;
          if constexpr (std::is_same<T, __half>::value) {   // same arithmetic as the unfused fp16 epilogue
            o = pack8<__half>(v);
            __half2* h = reinterpret_cast<__half2*>(&o);
            const __half2 sl = __float2half2_rn(slope);
#pragma unroll
            for (int j = 0; j < 4; ++j) h[j] = __hmax2(h[j], __hmul2(h[j], sl));
          } else {
#pragma unroll
            for (int ch = 0; ch < 8; ++ch) v[ch] = fmaxf(v[ch], v[ch] * slope);
            o = pack8<T>(v);
          }
          if (!live) o = make_uint4(0, 0, 0, 0);
          uint32_t off = static_cast<uint32_t>(r * ROWB + (c * 4 + q) * 16);
          off ^= ((off >> 7) & SWZ) << 4;
          *reinterpret_cast<uint4*>(mid_t + off) = o;
        }
      }
      tc_fence_before();
      fence_proxy_async();     // generic-proxy stores -> visible to the tensor core's async-proxy reads
      __syncwarp();
      if (lane == 0) { arrive_mma(&t1_empty[slot]); arrive_mma(&mid_full[ms]); }
    }
    if (g.dbg_out && warp == 2 && lane == 0) {
      long long* o = g.dbg_out + blockIdx.x * 16;
      o[6] = clock64() - pt0; o[7] = pw[0]; o[8] = pw[1];
    }
  } else {
    // ================================================================= E2: accumulator 2 -> global (fused epilogue)
    const int grp = (warp - 6) >> 2;
    const int lane_grp = warp & 3;
    uint8_t* stg = stage_base + (warp - 6) * kStageBytes;
    long long pw = 0;
    const long long pt0 = clock64();
    for (int n = 0; n < nt_cta; ++n) {
      const int tile = tile_of(n);
      const int b = tile / g.tiles_per_b, t0 = (tile % g.tiles_per_b) * g.S;
      const int slot = n & 1;
      const int t_lim = min(g.T, t0 + g.S);     // rows past S belong to the next tile
      const uint32_t taddr = tmem_base + (static_cast<uint32_t>(lane_grp * 32) << 16) + (2 + slot) * SLOT;
      if (PAIR && tile >= g.total_tiles) {
        mbar_wait(&t2_full[slot], (n >> 1) & 1);   // ghost tile: consume the phase, store nothing
      } else {
        epilogue_tile3<C, MB, EG, T, kFC_VOC>(e, b, t0, t_lim, 0, 0, taddr, stg, nullptr, lane, lane_grp, grp, &t2_full[slot],
                                     (n >> 1) & 1, g.dbg_out ? &pw : nullptr);
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) arrive_mma(&t2_empty[slot]);
    }
    if (g.dbg_out && warp == 6 && lane == 0) {
      long long* o = g.dbg_out + blockIdx.x * 16;
      o[9] = clock64() - pt0; o[10] = pw;
    }
  }
  tc_fence_before();
  if constexpr (PAIR) cluster_sync_all(); else __syncthreads();
  if (warp == 1) { if constexpr (PAIR) tmem_dealloc_pair<512>(tmem_base); else tmem_dealloc<512>(tmem_base); }
}

template <int C, int MB, typename T, bool PAIR>
static cudaError_t launch_pair_inst(const PairLaunch& L, cudaStream_t st) {
  auto kern = resblock_pair_kernel<C, MB, T, PAIR>;
  static PerDeviceOnce configured;
  if (configured.needed()) {
    cudaError_t err = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
    if (err != cudaSuccess) return err;
    configured.mark();
  }
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3(L.grid);
  cfg.blockDim = dim3(448);
  cfg.dynamicSmemBytes = L.smem_bytes;
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = 2; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr;
  cfg.numAttrs = PAIR ? 1 : 0;
  return cudaLaunchKernelEx(&cfg, kern, L.tmA, L.tmA8, L.tmW1, L.tmW2, L.g, L.e);
}

template <typename T>
cudaError_t launch_resblock_pair(const PairLaunch& L, cudaStream_t st) {
  if (L.C == 64 && L.mb == 2) return L.cta_pair ? launch_pair_inst<64, 2, T, true>(L, st) : launch_pair_inst<64, 2, T, false>(L, st);
  if (L.C == 32 && L.mb == 4) return L.cta_pair ? launch_pair_inst<32, 4, T, true>(L, st) : launch_pair_inst<32, 4, T, false>(L, st);
  return cudaErrorInvalidConfiguration;
}
template cudaError_t launch_resblock_pair<__nv_bfloat16>(const PairLaunch&, cudaStream_t);
template cudaError_t launch_resblock_pair<__half>(const PairLaunch&, cudaStream_t);

}  // namespace mb
