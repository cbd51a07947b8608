// The transformer block's feed-forward (matcha/models/components/transformer.py:131-134 with SnakeBeta :64-80, residual
// and mask of :311-316 / decoder.py:398-404) as ONE persistent tcgen05 kernel:
//
//     out = ( SnakeBeta(x W1^T + b1) W2^T + b2 + resid ) * mask            x = LayerNorm3(hidden) [rows][256]
//
// Run as two conv-GEMM launches the 1024-wide hidden activation makes a round trip through HBM (59 MB written and read
// back per T-level block at cfg2) and each launch has only one or two tiles per CTA, i.e. no steady state.  Here the
// hidden activation never leaves the SM, and one tile is a 16-step software pipeline: per tile of 128 rows
//     TMA   x tile [128][256] (4 K chunks)                               -> shared memory (kept for the whole tile)
//     for j in 0..7 (chunks of 128 hidden channels):
//       MMA1  x . W1[128j..128j+127]^T            (N = 128, K = 256)      -> TMEM accumulator ring 1 (2 x 128 columns)
//       E1    + b1, SnakeBeta, -> operand dtype                           -> shared memory "mid" chunk in the UMMA K-major
//                                                                            128B-swizzled layout (swizzled by hand)
//       MMA2  mid_j . W2[:, 128j..128j+127]^T     (N = 256, K = 128)      -> TMEM accumulator 2 (256 columns), accumulated over j
//     E2    + b2 + resid, * mask                                          -> global
// The issuer interleaves M1(j+1) before M2(j), so E1(j) overlaps M1(j+1) and M2(j-1).  Weights stream through a ring of
// three 32 KB stages (W1 chunk = 2 stages of two [128 x 64] boxes, W2 chunk = 2 stages of one [256 x 64] box) in exactly
// the order the issuer consumes them.  Same arithmetic and roundings as the two-launch form: results are bit-identical.
//
// Warp roles (320 threads): 0 TMA producer, 1 MMA issuer (converged warps, elected single-thread instructions),
// 2-9 E1 / E2 (warp % 4 = TMEM lane quarter, (warp - 2) / 4 = column half).
#include "conv.cuh"
#include "epilogue3.cuh"
#include "launch.h"
#include "ptx.cuh"
#include "rowio.cuh"

namespace mb {

namespace {
constexpr int kFFXBytes = 4 * 16384;      // x tile: 4 K chunks of [128 rows][64 ch]
constexpr int kFFMidBytes = 2 * 16384;    // one mid chunk: 2 K chunks of [128 rows][64 ch]
constexpr int kFFStage = 32768;           // one weight ring stage
constexpr int kFFStages = 3;
constexpr int kFFBarBytes = 512;
// developer probe (FFGeom::dbg_out): cycles a role spends inside one barrier wait
#define FW(stmt, acc) do { if (g.dbg_out) { const long long fw_a = clock64(); stmt; acc += clock64() - fw_a; } else { stmt; } } while (0)
}  // namespace

int ff_fused_smem_bytes() { return 1024 + kFFXBytes + 2 * kFFMidBytes + kFFStages * kFFStage + kFFBarBytes; }

// PAIR: two CTAs of a cluster run as one tcgen05 cta_group::2 unit (see conv_gemm2.cu): each keeps its own x tile, mid
// chunks and accumulators, but every weight box is split between the two CTAs' shared memories (W1: 64 of the chunk's 128
// hidden channels each, W2 / Wo: 128 of the 256 output channels each), so a CTA streams half of the 1 MB of weights a tile
// needs from L2.  The even CTA issues all MMAs (M = 256); barriers the issuer waits on collect arrivals from both CTAs.
template <typename T, bool PAIR>
__global__ void __launch_bounds__(320, 1)
ff_fused_kernel(const __grid_constant__ CUtensorMap tmX, const __grid_constant__ CUtensorMap tmW1,
                const __grid_constant__ CUtensorMap tmW2, const __grid_constant__ CUtensorMap tmWo,
                const __grid_constant__ FFGeom g) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  uint8_t* x_tile = smem;
  uint8_t* mid = x_tile + kFFXBytes;
  uint8_t* w_ring = mid + 2 * kFFMidBytes;
  constexpr int STAGE = PAIR ? kFFStage / 2 : kFFStage;   // this CTA's share of one weight ring stage
  constexpr uint32_t WBOX16 = (STAGE / 2) >> 4;            // second W1 box of a stage, in descriptor units
  uint64_t* bars = reinterpret_cast<uint64_t*>(w_ring + kFFStages * kFFStage);
  uint64_t* x_full = bars;             // [1]
  uint64_t* x_empty = bars + 1;        // [1]
  uint64_t* w_full = bars + 2;         // [3]
  uint64_t* w_empty = bars + 5;        // [3]
  uint64_t* t1_full = bars + 8;        // [2]
  uint64_t* t1_empty = bars + 10;      // [2]
  uint64_t* mid_full = bars + 12;      // [2]
  uint64_t* mid_empty = bars + 14;     // [2]
  uint64_t* t2_full = bars + 16;       // [1]
  uint64_t* t2_empty = bars + 17;      // [1]
  uint64_t* att_full = bars + 18;      // [1] pre-stage: attention tile landed in mid[0]
  uint64_t* att_empty = bars + 19;     // [1] mid[0] no longer read by this tile's M2 (chunk 6)
  uint64_t* t0_full = bars + 20;       // [1] pre-stage accumulator (TMEM columns 0..255) complete
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 21);
  const bool pre = g.pre != 0;         // tmX then maps the attention output [rows][128] instead of the LayerNorm output

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int rank = PAIR ? static_cast<int>(cluster_ctarank()) : 0;
  // work units: tiles, or pairs of tiles (2u, 2u + 1); the odd CTA of the last pair may get a tile past the end, which is
  // all padding (TMA zero fill, every global access is bounded by g.rows)
  const int units = PAIR ? (g.total_tiles + 1) / 2 : g.total_tiles;
  const int u0 = PAIR ? static_cast<int>(blockIdx.x >> 1) : static_cast<int>(blockIdx.x);
  const int ustep = PAIR ? static_cast<int>(gridDim.x >> 1) : static_cast<int>(gridDim.x);
  const int nt_cta = units > u0 ? (units - 1 - u0) / ustep + 1 : 0;
  auto tile_of = [&](int n) { const int u = u0 + n * ustep; return PAIR ? 2 * u + rank : u; };
  auto arrive_mma = [](uint64_t* bar) { if constexpr (PAIR) mbar_arrive_leader(bar); else mbar_arrive(bar); };   // barriers the issuer waits on

  if (threadIdx.x == 0) {
    tma_prefetch_desc(&tmX); tma_prefetch_desc(&tmW1); tma_prefetch_desc(&tmW2);
    tma_prefetch_desc(&tmWo);
    constexpr int EW = PAIR ? 16 : 8;   // epilogue warps that arrive on an issuer-side barrier
    mbar_init(x_full, pre ? EW : 1); mbar_init(x_empty, 1);
    mbar_init(att_full, 1); mbar_init(att_empty, 1); mbar_init(t0_full, 1);
    for (int s = 0; s < kFFStages; ++s) { mbar_init(&w_full[s], 1); mbar_init(&w_empty[s], 1); }
    for (int s = 0; s < 2; ++s) {
      mbar_init(&t1_full[s], 1); mbar_init(&t1_empty[s], EW);
      mbar_init(&mid_full[s], EW); mbar_init(&mid_empty[s], 1);
    }
    mbar_init(t2_full, 1); mbar_init(t2_empty, EW);
    fence_barrier_init();
  }
  if (warp == 1) { if constexpr (PAIR) tmem_alloc_pair<512>(tmem_slot); else tmem_alloc<512>(tmem_slot); }
  tc_fence_before();
  if constexpr (PAIR) cluster_sync_all(); else __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  // loads: in a pair the bytes of both CTAs are counted on the leader's barrier, which only the leader arms
  auto expect = [&](uint64_t* bar, uint32_t bytes) { if constexpr (PAIR) { if (rank == 0) mbar_expect_tx(bar, 2 * bytes); } else mbar_expect_tx(bar, bytes); };
  auto load2 = [](void* dst, const CUtensorMap* m, uint64_t* bar, int c0, int c1) {
    if constexpr (PAIR) tma_load_2d_pair(dst, m, bar, c0, c1); else tma_load_2d(dst, m, bar, c0, c1);
  };
  auto load3 = [](void* dst, const CUtensorMap* m, uint64_t* bar, int c0, int c1, int c2) {
    if constexpr (PAIR) tma_load_3d_pair(dst, m, bar, c0, c1, c2); else tma_load_3d(dst, m, bar, c0, c1, c2);
  };

  if (warp == 0) {
    // ================================================================= TMA producer
    int ws = 0;
    uint32_t wph = 0;
    long long pw[2] = {0, 0};
    const long long pt0 = clock64();
    auto load_X = [&](int n) {
      const int tile = tile_of(n);
      FW(mbar_wait(x_empty, (n & 1) ^ 1), pw[0]);
      if (elect_one()) {
        expect(x_full, kFFXBytes);
#pragma unroll
        for (int kc = 0; kc < 4; ++kc) load3(x_tile + kc * 16384, &tmX, x_full, kc * 64, tile * 128, 0);
      }
      __syncwarp();
    };
    auto load_W1 = [&](int j) {   // hidden channels [128 j, +128): two stages of two [128 n][64 k] boxes
      for (int s = 0; s < 2; ++s) {
        FW(mbar_wait(&w_empty[ws], wph ^ 1), pw[1]);
        if (elect_one()) {
          expect(&w_full[ws], STAGE);        // pair: this CTA's 64 of the chunk's 128 hidden channels
          const int n0 = j * 128 + (PAIR ? rank * 64 : 0);
          load2(w_ring + ws * STAGE, &tmW1, &w_full[ws], (2 * s) * 64, n0);
          load2(w_ring + ws * STAGE + STAGE / 2, &tmW1, &w_full[ws], (2 * s + 1) * 64, n0);
        }
        __syncwarp();
        if (++ws == kFFStages) { ws = 0; wph ^= 1; }
      }
    };
    auto load_W2 = [&](int j) {   // K slice [128 j, +128) of the down projection: two stages of one [256 n][64 k] box
      for (int s = 0; s < 2; ++s) {
        FW(mbar_wait(&w_empty[ws], wph ^ 1), pw[1]);
        if (elect_one()) {
          expect(&w_full[ws], STAGE);        // pair: this CTA's 128 of the 256 output channels
          load2(w_ring + ws * STAGE, &tmW2, &w_full[ws], j * 128 + s * 64, PAIR ? rank * 128 : 0);
        }
        __syncwarp();
        if (++ws == kFFStages) { ws = 0; wph ^= 1; }
      }
    };
    auto load_pre = [&](int n) {   // attention tile [128][128] -> mid[0] (two K chunks), Wo [256][128] -> two ring stages
      const int tile = tile_of(n);
      FW(mbar_wait(att_empty, (n & 1) ^ 1), pw[0]);
      if (elect_one()) {
        expect(att_full, kFFMidBytes);
        load3(mid, &tmX, att_full, 0, tile * 128, 0);
        load3(mid + 16384, &tmX, att_full, 64, tile * 128, 0);
      }
      __syncwarp();
      for (int s = 0; s < 2; ++s) {
        FW(mbar_wait(&w_empty[ws], wph ^ 1), pw[1]);
        if (elect_one()) {
          expect(&w_full[ws], STAGE);
          load2(w_ring + ws * STAGE, &tmWo, &w_full[ws], s * 64, PAIR ? rank * 128 : 0);
        }
        __syncwarp();
        if (++ws == kFFStages) { ws = 0; wph ^= 1; }
      }
    };
    if (nt_cta > 0 && !pre) load_X(0);
    for (int n = 0; n < nt_cta; ++n) {
      if (pre) load_pre(n);
      load_W1(0);
      for (int j = 1; j < 8; ++j) { load_W1(j); load_W2(j - 1); if (j == 7 && n + 1 < nt_cta && !pre) load_X(n + 1); }
      load_W2(7);
    }
    if (g.dbg_out && lane == 0) { long long* o = g.dbg_out + blockIdx.x * 16; o[0] = clock64() - pt0; o[1] = pw[0]; o[2] = pw[1]; }
  } else if (warp == 1) {
    // ================================================================= MMA issuer (pair: the leader's only)
    if (rank == 0) {
    constexpr int fmt = std::is_same<T, __nv_bfloat16>::value ? 1 : 0;
    constexpr int MM = PAIR ? 256 : 128;
    constexpr uint32_t idesc1 = umma_idesc_f16(fmt, MM, 128), idesc2 = umma_idesc_f16(fmt, MM, 256);
    auto umma_f16_lohi = [](uint32_t d, uint32_t a, uint32_t b, uint32_t hi, uint32_t id, uint32_t acc) {
      if constexpr (PAIR) mb::umma_f16_lohi_pair(d, a, b, hi, id, acc); else mb::umma_f16_lohi(d, a, b, hi, id, acc);
    };
    auto umma_commit = [](uint64_t* bar) { if constexpr (PAIR) mb::umma_commit_pair(bar); else mb::umma_commit(bar); };
    const uint32_t desc_hi = static_cast<uint32_t>(umma_smem_desc<128>(0) >> 32);
    const uint32_t x_lo = static_cast<uint32_t>(umma_smem_desc<128>(smem_u32(x_tile)));
    const uint32_t mid_lo = static_cast<uint32_t>(umma_smem_desc<128>(smem_u32(mid)));
    const uint32_t w_lo0 = static_cast<uint32_t>(umma_smem_desc<128>(smem_u32(w_ring)));
    int ws = 0;
    uint32_t wph = 0;
    long long iw[5] = {0, 0, 0, 0, 0};
    const long long it0 = clock64();
    uint32_t pe[2] = {0, 0};   // uses of each acc1 slot so far (parity): the pre-stage adds one use per tile
    for (int n = 0; n < nt_cta; ++n) {
      if (pre) {                // ---- M0: TMEM[0..255] = att . Wo^T   (N = 256, K = 128)
        FW(mbar_wait(att_full, n & 1), iw[0]);
        FW(mbar_wait(&t1_empty[0], pe[0] ^ 1), iw[1]); pe[0] ^= 1;
        FW(mbar_wait(&t1_empty[1], pe[1] ^ 1), iw[1]); pe[1] ^= 1;
        for (int s = 0; s < 2; ++s) {
          FW(mbar_wait(&w_full[ws], wph), iw[2]);
          tc_fence_after();
          if (elect_one()) {
            const uint32_t w = w_lo0 + ws * (STAGE >> 4), a = mid_lo + s * 1024u;
#pragma unroll
            for (int k = 0; k < 4; ++k) umma_f16_lohi(tmem_base, a + 2 * k, w + 2 * k, desc_hi, idesc2, (s | k) != 0 ? 1u : 0u);
            umma_commit(&w_empty[ws]);
            if (s == 1) umma_commit(t0_full);
          }
          __syncwarp();
          if (++ws == kFFStages) { ws = 0; wph ^= 1; }
        }
      }
      FW(mbar_wait(x_full, n & 1), iw[0]);
#pragma unroll 1
      for (int j = 0; j <= 8; ++j) {
        if (j < 8) {            // ---- M1(j): acc1[slot] = x . W1_j^T
          const int jj = n * 8 + j, slot = jj & 1;
          FW(mbar_wait(&t1_empty[slot], pe[slot] ^ 1), iw[1]); pe[slot] ^= 1;
          const uint32_t dtm = tmem_base + slot * 128;
          for (int s = 0; s < 2; ++s) {
            FW(mbar_wait(&w_full[ws], wph), iw[2]);
            tc_fence_after();
            if (elect_one()) {
              const uint32_t w = w_lo0 + ws * (STAGE >> 4);
#pragma unroll
              for (int c = 0; c < 2; ++c) {
                const uint32_t a = x_lo + (2 * s + c) * 1024u, wc = w + c * WBOX16;
#pragma unroll
                for (int k = 0; k < 4; ++k) umma_f16_lohi(dtm, a + 2 * k, wc + 2 * k, desc_hi, idesc1, (s | c | k) != 0 ? 1u : 0u);
              }
              umma_commit(&w_empty[ws]);
              if (s == 1) { umma_commit(&t1_full[slot]); if (j == 7) umma_commit(x_empty); }
            }
            __syncwarp();
            if (++ws == kFFStages) { ws = 0; wph ^= 1; }
          }
        }
        if (j >= 1) {           // ---- M2(j-1): acc2 += mid_{j-1} . W2[:, slice j-1]^T
          const int jm = j - 1, jj = n * 8 + jm, ms = jj & 1;
          if (jm == 0) FW(mbar_wait(t2_empty, (n & 1) ^ 1), iw[4]);
          FW(mbar_wait(&mid_full[ms], (jj >> 1) & 1), iw[3]);
          const uint32_t dtm = tmem_base + 256;
          for (int s = 0; s < 2; ++s) {
            FW(mbar_wait(&w_full[ws], wph), iw[2]);
            tc_fence_after();
            if (elect_one()) {
              const uint32_t w = w_lo0 + ws * (STAGE >> 4);
              const uint32_t a = mid_lo + ms * (kFFMidBytes >> 4) + s * 1024u;
#pragma unroll
              for (int k = 0; k < 4; ++k) umma_f16_lohi(dtm, a + 2 * k, w + 2 * k, desc_hi, idesc2, (jm | s | k) != 0 ? 1u : 0u);
              umma_commit(&w_empty[ws]);
              if (s == 1) { umma_commit(&mid_empty[ms]); if (jm == 6) umma_commit(att_empty); if (jm == 7) umma_commit(t2_full); }
            }
            __syncwarp();
            if (++ws == kFFStages) { ws = 0; wph ^= 1; }
          }
        }
      }
    }
    if (g.dbg_out && lane == 0) { long long* o = g.dbg_out + blockIdx.x * 16; o[3] = clock64() - it0; for (int i = 0; i < 5; ++i) o[4 + i] = iw[i]; }
    }
  } else {
    // ================================================================= E1 (SnakeBeta -> mid) and E2 (-> global)
    const int lane_grp = warp & 3, half = (warp - 2) >> 2;
    const int r = lane_grp * 32 + lane;                     // row of the tile held by this lane
    const uint32_t tlane = tmem_base + (static_cast<uint32_t>(lane_grp * 32) << 16);
    long long ew[3] = {0, 0, 0};
    const long long et0 = clock64();
    // Warm L1 with the per-channel parameters (13 KB) while the first x tile and weights are in flight: every later
    // parameter load is then an L1 hit instead of a dependent chain of cold L2 round trips (measured: E1 was 4 000
    // cycles per chunk against 2 048 cycles of MMA work).
    {
      const int tid = threadIdx.x - 64;   // 0..255
      for (int i = tid; i < 3 * 1024 / 32 + 256 / 32; i += 256) {
        const float* base = i < 32 ? g.bias1 : (i < 64 ? g.snake_a : (i < 96 ? g.snake_invb : g.bias2));
        asm volatile("prefetch.global.L1 [%0];" ::"l"(base + (i & 31) * 32));
      }
      if (pre && tid < 24) {
        const float* base = tid < 8 ? g.bias_o : (tid < 16 ? g.ln_g : g.ln_b);
        asm volatile("prefetch.global.L1 [%0];" ::"l"(base + (tid & 7) * 32));
      }
    }
    // Pair mode has 48 KB of shared memory to spare (half-size weight stages): the 3 x 1024 per-channel SnakeBeta parameters
    // live there for the whole kernel and E1 reads them with broadcast LDS.128 (6 per 8 channels) instead of three warp
    // shuffles per element (192 per chunk and lane: the shuffle pipe, not the MMAs, set E1's pace).
    float* ptab = reinterpret_cast<float*>(w_ring + kFFStages * STAGE);
    if constexpr (PAIR) {
      const int tid = threadIdx.x - 64;
      for (int i = tid; i < 3 * 1024; i += 256) ptab[i] = __ldg((i < 1024 ? g.bias1 : i < 2048 ? g.snake_a : g.snake_invb) + (i & 1023));
      named_bar_sync(5, 256);
    }
    float pc[6], pn[6];
    auto ldpar = [&](int col, float (&pp)[6]) {   // lane l: channels col + l and col + 32 + l of the three arrays
      pp[0] = __ldg(g.bias1 + col + lane); pp[1] = __ldg(g.bias1 + col + 32 + lane);
      pp[2] = __ldg(g.snake_a + col + lane); pp[3] = __ldg(g.snake_a + col + 32 + lane);
      pp[4] = __ldg(g.snake_invb + col + lane); pp[5] = __ldg(g.snake_invb + col + 32 + lane);
    };
    for (int n = 0; n < nt_cta; ++n) {
      const int tile = tile_of(n);
      if (pre) {
        // ---- E0: hid = att Wo^T + bo + res -> global; x = LayerNorm(hid) -> x tile (A operand of M1).  Same arithmetic as
        // the unfused out-projection epilogue: row sums from the fp32 values, normalisation of the ROUNDED stored values.
        const long row = static_cast<long>(tile) * 128 + r;
        const bool ok = row < g.rows;
        // Rolled loops on purpose: as straight-line code this once-per-tile stage is bound by instruction fetch
        // (resqkv_fused.cu's epilogue, measured).  The rounded hidden values go to HBM and, un-normalised, into the x tile,
        // where the LayerNorm pass normalises them in place (same thread, same addresses).
        // Global traffic is staged through the x tile (rowio.cuh): coalesced loads of the residual rows, coalesced stores
        // of the hidden rows; row-per-lane global accesses touch 32 lines per instruction and were the bottleneck.
        const long row0 = static_cast<long>(tile) * 128;
        const int r0 = lane_grp * 32;
        FW(mbar_wait(x_empty, (n & 1) ^ 1), ew[1]);          // the previous tile's M1s no longer read the x tile
        warp_rows_g2s<T>(x_tile, reinterpret_cast<const T*>(g.res_in), row0, g.rows, r0, half, lane);
        __syncwarp();
        FW(mbar_wait(t0_full, n & 1), ew[0]);
        tc_fence_after();
        float sm = 0.f, sq = 0.f;
#pragma unroll 1
        for (int cb = 0; cb < 4; ++cb) {
          const int n0 = half * 128 + cb * 32;
          float acc[32];
          tmem_ld_32x32(tlane + n0, acc);
          uint4* xp[4];
          uint4 cur[4];
#pragma unroll
          for (int q = 0; q < 4; ++q) {
            const int i = cb * 4 + q;
            xp[q] = reinterpret_cast<uint4*>(x_tile + (half * 2 + (i >> 3)) * 16384 + xtile_off(r, i & 7));
            cur[q] = *xp[q];
          }
          tmem_ld_wait();
#pragma unroll
          for (int q = 0; q < 4; ++q) {
            float bv[8], rv[8], v[8];
            ldg8f(g.bias_o + n0 + 8 * q, bv);
            unpack8<T>(cur[q], rv);
#pragma unroll
            for (int ch = 0; ch < 8; ++ch) { v[ch] = acc[8 * q + ch] + bv[ch]; v[ch] += rv[ch]; sm += v[ch]; sq = fmaf(v[ch], v[ch], sq); }
            *xp[q] = pack8<T>(v);
          }
        }
        __syncwarp();
        warp_rows_s2g<T>(x_tile, reinterpret_cast<T*>(g.hid_out), row0, g.rows, r0, half, lane);
        __syncwarp();
        tc_fence_before();
        __syncwarp();
        if (lane == 0) { arrive_mma(&t1_empty[0]); arrive_mma(&t1_empty[1]); }   // TMEM columns 0..255 are free again
        // row statistics: the other half of the row belongs to the partner warp (same TMEM lane quarter); mid[0] (the
        // attention tile, dead once M0 completed) carries the exchange
        float2* xch = reinterpret_cast<float2*>(mid);
        xch[half * 128 + r] = make_float2(sm, sq);
        named_bar_sync(1 + lane_grp, 64);
        const float2 s0 = xch[r], s1 = xch[128 + r];
        named_bar_sync(1 + lane_grp, 64);                     // both warps have read before E1 overwrites mid[0]
        const float mean = (s0.x + s1.x) * (1.f / 256.f);
        const float rstd = rsqrtf(fmaxf((s0.y + s1.y) * (1.f / 256.f) - mean * mean, 0.f) + 1e-5f);
#pragma unroll 2
        for (int i = 0; i < 16; ++i) {
          float ga[8], be[8], xv[8];
          ldg8f(g.ln_g + half * 128 + 8 * i, ga);
          ldg8f(g.ln_b + half * 128 + 8 * i, be);
          uint32_t off = static_cast<uint32_t>(r * 128 + (i & 7) * 16);
          off ^= ((off >> 7) & 7u) << 4;
          uint4* xp = reinterpret_cast<uint4*>(x_tile + (half * 2 + (i >> 3)) * 16384 + off);
          unpack8<T>(*xp, xv);
#pragma unroll
          for (int ch = 0; ch < 8; ++ch) xv[ch] = fmaf((xv[ch] - mean) * rstd, ga[ch], be[ch]);
          *xp = pack8<T>(xv);
        }
        fence_proxy_async();
        __syncwarp();
        if (lane == 0) arrive_mma(x_full);
      }
#pragma unroll 1
      for (int j = 0; j < 8; ++j) {
        const int jj = n * 8 + j, slot = jj & 1;
        uint8_t* mid_c = mid + slot * kFFMidBytes + half * 16384;
        const int nb = j * 128 + half * 64;                 // first hidden channel of this lane's 64 columns
        // Per-channel parameters of this chunk (bias, exp(alpha), 1/(exp(beta)+eps) of 64 hidden channels): lane l holds
        // channels l and 32 + l, fetched with six coalesced loads ONE CHUNK AHEAD, and every use is a warp shuffle.  As
        // 48 broadcast LDG.128 per chunk issued one 8-channel group ahead, these loads were the kernel's dominant stall
        // (long scoreboard: half of all warp-stall samples, profiles/r01_ncu_full_ff_fused.txt) - with ~226 KB of the SM
        // carved out as shared memory the L1 does not hold even 13 KB of parameters.
        if constexpr (!PAIR) {
          if (j == 0 && n == 0) ldpar(nb, pc);
          ldpar(((j + 1) & 7) * 128 + half * 64, pn);
        }
        FW(mbar_wait(&t1_full[slot], (jj >> 1) & 1), ew[0]);
        tc_fence_after();
        float acc[2][32];
        tmem_ld_32x32(tlane + slot * 128 + half * 64, acc[0]);
        tmem_ld_32x32(tlane + slot * 128 + half * 64 + 32, acc[1]);
        FW(mbar_wait(&mid_empty[slot], ((jj >> 1) & 1) ^ 1), ew[1]);   // the M2 that read this mid buffer two chunks ago is done
        tmem_ld_wait();
#pragma unroll
        for (int q = 0; q < 8; ++q) {
          float v[8];
          const float* a = acc[q >> 2] + 8 * (q & 3);
          if constexpr (PAIR) {
            const float4* tb = reinterpret_cast<const float4*>(ptab + nb + 8 * q);
            const float4 b0 = tb[0], b1 = tb[1], a0 = tb[256], a1 = tb[257], i0 = tb[512], i1 = tb[513];
            const float bv[8] = {b0.x, b0.y, b0.z, b0.w, b1.x, b1.y, b1.z, b1.w};
            const float pa[8] = {a0.x, a0.y, a0.z, a0.w, a1.x, a1.y, a1.z, a1.w};
            const float pb[8] = {i0.x, i0.y, i0.z, i0.w, i1.x, i1.y, i1.z, i1.w};
#pragma unroll
            for (int ch = 0; ch < 8; ++ch) {
              v[ch] = a[ch] + bv[ch];
              const float sn = __sinf(v[ch] * pa[ch]);
              v[ch] = fmaf(pb[ch] * sn, sn, v[ch]);
            }
          } else {
#pragma unroll
            for (int ch = 0; ch < 8; ++ch) {
              const int src = (8 * q + ch) & 31;
              const float bv = __shfl_sync(0xffffffffu, pc[q >> 2], src);
              const float pa = __shfl_sync(0xffffffffu, pc[2 + (q >> 2)], src);
              const float pb = __shfl_sync(0xffffffffu, pc[4 + (q >> 2)], src);
              v[ch] = a[ch] + bv;
              const float sn = __sinf(v[ch] * pa);
              v[ch] = fmaf(pb * sn, sn, v[ch]);
            }
          }
          uint32_t off = static_cast<uint32_t>(r * 128 + q * 16);
          off ^= ((off >> 7) & 7u) << 4;
          *reinterpret_cast<uint4*>(mid_c + off) = pack8<T>(v);
        }
        if constexpr (!PAIR) {
#pragma unroll
          for (int k = 0; k < 6; ++k) pc[k] = pn[k];
        }
        tc_fence_before();
        fence_proxy_async();     // generic-proxy stores -> visible to the tensor core's async-proxy reads
        __syncwarp();
        if (lane == 0) { arrive_mma(&t1_empty[slot]); arrive_mma(&mid_full[slot]); }
      }
      // ---- E2: this warp's 128 output channels [128 half, +128) of the lane's row.  The residual row segment (256 B)
      // is requested before the accumulator barrier, so its latency hides behind the last chunk's MMAs.
      const long row = static_cast<long>(tile) * 128 + r;
      const bool ok = row < g.rows;
      const float m = ok ? __ldg(g.mask + row) : 0.f;
      if (pre) {
        // The x tile is idle once the last M1 has completed (nothing refills it before the next tile's E0), so the residual
        // rows and the output rows are staged through it for coalesced global traffic (rowio.cuh).  The residual is the
        // hidden tensor E0 stored with the same lane -> address mapping, hence plain (coherent) loads.
        const long row0 = static_cast<long>(tile) * 128;
        const int r0 = lane_grp * 32;
        FW(mbar_wait(x_empty, n & 1), ew[1]);
        warp_rows_g2s<T, false>(x_tile, reinterpret_cast<const T*>(g.resid), row0, g.rows, r0, half, lane);
        __syncwarp();
        FW(mbar_wait(t2_full, n & 1), ew[2]);
        tc_fence_after();
#pragma unroll 1
        for (int cb = 0; cb < 4; ++cb) {
          const int n0 = half * 128 + cb * 32;
          float acc[32];
          tmem_ld_32x32(tlane + 256 + n0, acc);
          uint4* xp[4];
          uint4 cur[4];
#pragma unroll
          for (int q = 0; q < 4; ++q) {
            const int i = cb * 4 + q;
            xp[q] = reinterpret_cast<uint4*>(x_tile + (half * 2 + (i >> 3)) * 16384 + xtile_off(r, i & 7));
            cur[q] = *xp[q];
          }
          tmem_ld_wait();
#pragma unroll
          for (int q = 0; q < 4; ++q) {
            float bv[8], rv[8], v[8];
            ldg8f(g.bias2 + n0 + 8 * q, bv);
            unpack8<T>(cur[q], rv);
#pragma unroll
            for (int ch = 0; ch < 8; ++ch) { v[ch] = acc[8 * q + ch] + bv[ch]; v[ch] += rv[ch]; v[ch] *= m; }
            *xp[q] = pack8<T>(v);
          }
        }
        tc_fence_before();
        __syncwarp();
        if (lane == 0) arrive_mma(t2_empty);     // the accumulators are read: the next tile's MMAs may start
        warp_rows_s2g<T>(x_tile, reinterpret_cast<T*>(g.out), row0, g.rows, r0, half, lane);
        __syncwarp();
        continue;
      }
      const T* rp = reinterpret_cast<const T*>(g.resid) + row * 256 + half * 128;
      T* op = reinterpret_cast<T*>(g.out) + row * 256 + half * 128;
      uint4 ru[4];   // plain loads: with the pre-stage this thread wrote the row itself
#pragma unroll
      for (int q = 0; q < 4; ++q) { ru[q] = make_uint4(0, 0, 0, 0); if (ok) ru[q] = *(reinterpret_cast<const uint4*>(rp) + q); }
      FW(mbar_wait(t2_full, n & 1), ew[2]);
      tc_fence_after();
#pragma unroll 1
      for (int cb = 0; cb < 4; ++cb) {      // rolled: see E0
        const int n0 = half * 128 + cb * 32;
        float acc[32];
        tmem_ld_32x32(tlane + 256 + n0, acc);
        uint4 cur[4];
#pragma unroll
        for (int q = 0; q < 4; ++q) cur[q] = ru[q];
        if (cb + 1 < 4) {
#pragma unroll
          for (int q = 0; q < 4; ++q) { ru[q] = make_uint4(0, 0, 0, 0); if (ok) ru[q] = *(reinterpret_cast<const uint4*>(rp) + (cb + 1) * 4 + q); }
        }
        tmem_ld_wait();
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          float bv[8], rv[8], v[8];
          ldg8f(g.bias2 + n0 + 8 * q, bv);
          unpack8<T>(cur[q], rv);
#pragma unroll
          for (int ch = 0; ch < 8; ++ch) { v[ch] = acc[8 * q + ch] + bv[ch]; v[ch] += rv[ch]; v[ch] *= m; }
          if (ok) *reinterpret_cast<uint4*>(op + cb * 32 + 8 * q) = pack8<T>(v);
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) arrive_mma(t2_empty);
    }
    if (g.dbg_out && warp == 2 && lane == 0) { long long* o = g.dbg_out + blockIdx.x * 16; o[9] = clock64() - et0; for (int i = 0; i < 3; ++i) o[10 + i] = ew[i]; o[13] = nt_cta; }
  }
  tc_fence_before();
  if constexpr (PAIR) cluster_sync_all(); else __syncthreads();
  if (warp == 1) { if constexpr (PAIR) tmem_dealloc_pair<512>(tmem_base); else tmem_dealloc<512>(tmem_base); }
}

template <typename T, bool PAIR>
static cudaError_t launch_ff_fused_inst(const FFLaunch& L, cudaStream_t st) {
  auto kern = ff_fused_kernel<T, PAIR>;
  static PerDeviceOnce configured;
  if (configured.needed()) {
    cudaError_t err = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
    if (err != cudaSuccess) return err;
    configured.mark();
  }
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3(L.grid);
  cfg.blockDim = dim3(320);
  cfg.dynamicSmemBytes = L.smem_bytes;
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = 2; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr;
  cfg.numAttrs = PAIR ? 1 : 0;
  return cudaLaunchKernelEx(&cfg, kern, L.tmX, L.tmW1, L.tmW2, L.tmWo, L.g);
}

template <typename T>
cudaError_t launch_ff_fused(const FFLaunch& L, cudaStream_t st) {
  return L.pair ? launch_ff_fused_inst<T, true>(L, st) : launch_ff_fused_inst<T, false>(L, st);
}
template cudaError_t launch_ff_fused<__nv_bfloat16>(const FFLaunch&, cudaStream_t);
template cudaError_t launch_ff_fused<__half>(const FFLaunch&, cudaStream_t);

}  // namespace mb
