// Persistent tcgen05 conv-GEMM (the production tensor-core kernel).
//
// One CTA per SM loops over output tiles (mb x 128 rows x NTILE channels).  Warp roles:
//   warp 0      TMA producer: per K chunk ONE halo load of the A tile (128*mb + 8*extra rows, 128B/64B swizzle) that
//               all taps of the conv reuse, plus the weight tiles (streamed, or loaded once and kept resident when the
//               whole weight matrix fits beside the A ring -- the 64/32-channel HiFi-GAN stages).
//   warp 1      TMEM owner; issues every tcgen05.mma from a flat host-built schedule (ConvGeom2::sched):
//               taps address the halo tile through row-shifted shared-memory descriptors, operands advance by 32-bit
//               adds on the descriptors' low words.  Accumulators are a ring of NACC TMEM buffers.
//   warps 2-9   epilogue (epilogue3.cuh): the eight warps split every tile's 32-column blocks between them, so the
//               epilogue of tile i overlaps the loads and MMAs of tile i+1.
#include "conv.cuh"
#include "epilogue3.cuh"
#include "launch.h"
#include "ptx.cuh"

namespace mb {

#ifdef MB200_PROBE
#define MB_DBG(bit) (g.dbg & (bit))
// cycle accounting of one role: total loop time and time spent inside barrier waits
#define MB_T0() const long long probe_t0 = clock64(); long long probe_w = 0
#define MB_T0E() const long long probe_t0 = clock64(); long long probe_e[6] = {0, 0, 0, 0, 0, 0}; long long& probe_w = probe_e[0]
#define MB_WAIT(stmt) do { const long long probe_a = clock64(); stmt; probe_w += clock64() - probe_a; } while (0)
#define MB_TEND(slot) do { if (lane == 0 && g.dbg_out) { g.dbg_out[blockIdx.x * 16 + 2 * (slot)] = clock64() - probe_t0; g.dbg_out[blockIdx.x * 16 + 2 * (slot) + 1] = probe_w; } } while (0)
#else
#define MB_DBG(bit) false
#define MB_T0()
#define MB_T0E()
#define MB_WAIT(stmt) stmt
#define MB_TEND(slot)
#endif

constexpr int kMaxRing = 8;
constexpr int kBarBytes = 512;   // mbarriers + TMEM slot

// PAIR: two CTAs of a cluster work as one tcgen05 cta_group::2 unit.  Each computes its own 128-row tile, but the MMAs
// (issued by the even CTA for both) read the weight tile HALF from each CTA's shared memory, so a CTA streams only
// NTILE / 2 weight rows per K step from L2: the N = 256 estimator layers are bound by that stream.
template <int NTILE, int KC, int MB, int EG, typename T, uint32_t F, bool PAIR = false>
__global__ void __launch_bounds__(64 + 128 * EG, 1)
conv_gemm_umma2_kernel(const __grid_constant__ CUtensorMap tmA0, const __grid_constant__ CUtensorMap tmA1,
                       const __grid_constant__ CUtensorMap tmA0x, const __grid_constant__ CUtensorMap tmA1x,
                       const __grid_constant__ CUtensorMap tmW, const __grid_constant__ ConvGeom2 g,
                       const __grid_constant__ EpiParams e) {
#ifdef MB200_PROBE
  const long long probe_k0 = clock64();
#endif
  constexpr int ROWB = KC * 2;
  constexpr int W_TILE = (PAIR ? NTILE / 2 : NTILE) * ROWB;   // weight rows of one K chunk held by THIS CTA
  constexpr int TMEM_COLS = 512;   // one CTA per SM: take all of TMEM
  constexpr int SLOT_COLS = MB * NTILE;
  constexpr int NACC = (512 / SLOT_COLS) < 4 ? (512 / SLOT_COLS) : 4;
  constexpr int TROWS = 128 * MB;
  const int A_BYTES = g.a_bytes;
  const int W_STAGE = g.wt * W_TILE;   // streamed weights: one ring stage holds the tiles of g.wt consecutive taps
  extern __shared__ uint8_t smem_raw[];
  // 1024-byte alignment for the swizzled tiles; offsets stay relative to smem_raw so accesses compile to LDS/STS
  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);

  int kchunks = 0;  // K chunks per (phase, n_tile): sum over groups of chunks * taps
  for (int q = 0; q < g.n_groups; ++q) kchunks += g.g_chunks[q] * g.g_ntaps[q];
  const int w_bytes = g.w_resident ? g.n_phases * g.n_tiles * kchunks * W_TILE : g.w_stages * W_STAGE;
  uint8_t* a_ring = smem;
  uint8_t* w_base = smem + g.a_stages * A_BYTES;
  uint8_t* tail = w_base + w_bytes;   // multiple of 1024 from smem
  uint64_t* bars = reinterpret_cast<uint64_t*>(tail);
  uint64_t* a_full = bars;
  uint64_t* a_empty = a_full + kMaxRing;
  uint64_t* w_full = a_empty + kMaxRing;
  uint64_t* w_empty = w_full + kMaxRing;
  uint64_t* t_full = w_empty + kMaxRing;
  uint64_t* t_empty = t_full + 4;
  uint64_t* wres = t_empty + 4;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(wres + 1);
  uint8_t* stage_base = tail + kBarBytes;                                   // 8 warp-private staging tiles
  float2* lnx = reinterpret_cast<float2*>(stage_base + 4 * EG * kStageBytes);    // LayerNorm partial-sum exchange

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int np = g.n_tiles * g.n_phases;
  const int rank = PAIR ? static_cast<int>(cluster_ctarank()) : 0;   // 0 = the pair's leader (issues the MMAs)

  if (threadIdx.x == 0) {
    tma_prefetch_desc(&tmA0); tma_prefetch_desc(&tmA0x); tma_prefetch_desc(&tmW);
    for (int s = 0; s < kMaxRing; ++s) { mbar_init(&a_full[s], 1); mbar_init(&a_empty[s], 1); mbar_init(&w_full[s], 1); mbar_init(&w_empty[s], 1); }
    for (int s = 0; s < 4; ++s) { mbar_init(&t_full[s], 1); mbar_init(&t_empty[s], (PAIR ? 8 : 4) * EG); }   // pair: both CTAs' epilogue warps
    mbar_init(wres, 1);
    fence_barrier_init();
  }
  if (warp == 1) { if constexpr (PAIR) tmem_alloc_pair<TMEM_COLS>(tmem_slot); else tmem_alloc<TMEM_COLS>(tmem_slot); }
  tc_fence_before();
  if constexpr (PAIR) cluster_sync_all(); else __syncthreads();   // pair: the peer's barriers are initialised before anything arrives on them
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
#ifdef MB200_PROBE
  const long long probe_k1 = clock64();
#endif
  // Programmatic dependent launch: everything above (and the resident weight loads below) overlaps the previous
  // kernel's tail; activations are only touched after griddep_wait().
  if (threadIdx.x == 0) griddep_launch();

  if (warp == 0) {
    // ================================================================= TMA producer (converged warp, elected issue)
    if (g.w_resident) {
      if (elect_one()) {
        if constexpr (PAIR) {   // this CTA's half of every weight tile; the leader's barrier counts both halves
          if (rank == 0) mbar_expect_tx(wres, 2u * static_cast<uint32_t>(w_bytes));
          for (int pn = 0; pn < np; ++pn)
            for (int kk = 0; kk < kchunks; ++kk)
              tma_load_2d_pair(w_base + static_cast<size_t>(pn * kchunks + kk) * W_TILE, &tmW, wres, kk * KC,
                               (pn / g.n_tiles) * g.N + (pn % g.n_tiles) * NTILE + rank * (NTILE / 2));
        } else {
          mbar_expect_tx(wres, static_cast<uint32_t>(w_bytes));
          for (int pn = 0; pn < np; ++pn)   // pn = phase * n_tiles + n_tile
            for (int kk = 0; kk < kchunks; ++kk)
              tma_load_2d(w_base + static_cast<size_t>(pn * kchunks + kk) * W_TILE, &tmW, wres, kk * KC,
                          (pn / g.n_tiles) * g.N + (pn % g.n_tiles) * NTILE);
        }
      }
      __syncwarp();
    }
    griddep_wait();
    int as = 0, ws = 0;
    uint32_t aph = 0, wph = 0;
    MB_T0();
    // pair: both CTAs run the same number of iterations; the odd CTA of the last pair may have no tile of its own
    // ("ghost"): it repeats the last tile's loads so that the leader's byte counts hold, and stores nothing.
    for (int tl = PAIR ? (blockIdx.x >> 1) : blockIdx.x; (PAIR ? 2 * tl : tl) < g.total_tiles; tl += PAIR ? (gridDim.x >> 1) : gridDim.x) {
      const int tile = min(PAIR ? 2 * tl + rank : tl, g.total_tiles - 1);
      const int nn = tile % np, mm = tile / np;
      const int n_tile = nn % g.n_tiles, phase = nn / g.n_tiles;
      const int t0 = (mm % g.m_tiles) * TROWS, b = mm / g.m_tiles;
      for (int q = 0; q < g.n_groups; ++q) {
        const CUtensorMap* m128 = g.g_src[q] ? &tmA1 : &tmA0;
        const CUtensorMap* m8 = g.g_src[q] ? &tmA1x : &tmA0x;
        const int row0 = t0 + g.g_rowmin[phase][q], extra = g.g_extra[phase][q];
        const int ntaps = g.g_ntaps[q];
        for (int c = 0; c < g.g_chunks[q]; ++c) {
          MB_WAIT(mbar_wait(&a_empty[as], aph ^ 1));
          if (elect_one()) {
            uint8_t* dst = a_ring + as * A_BYTES;
            const int ch = g.g_ch[q] + c * KC;
            if (MB_DBG(2)) {
              mbar_arrive(&a_full[as]);
            } else if constexpr (PAIR) {   // the leader's barrier counts both CTAs' bytes
              if (rank == 0) mbar_expect_tx(&a_full[as], 2 * (TROWS + 8 * extra) * ROWB);
#pragma unroll
              for (int i = 0; i < MB; ++i) tma_load_3d_pair(dst + i * 128 * ROWB, m128, &a_full[as], ch, row0 + 128 * i, b);
              for (int x = 0; x < extra; ++x) tma_load_3d_pair(dst + (TROWS + 8 * x) * ROWB, m8, &a_full[as], ch, row0 + TROWS + 8 * x, b);
            } else {
              mbar_expect_tx(&a_full[as], (TROWS + 8 * extra) * ROWB);
#pragma unroll
              for (int i = 0; i < MB; ++i) tma_load_3d(dst + i * 128 * ROWB, m128, &a_full[as], ch, row0 + 128 * i, b);
              for (int x = 0; x < extra; ++x) tma_load_3d(dst + (TROWS + 8 * x) * ROWB, m8, &a_full[as], ch, row0 + TROWS + 8 * x, b);
            }
          }
          __syncwarp();
          if (++as == g.a_stages) { as = 0; aph ^= 1; }
          if (!g.w_resident) {
            for (int tp = 0; tp < ntaps; tp += g.wt) {   // one ring stage = up to wt taps of this K chunk
              const int nt = min(g.wt, ntaps - tp);
              MB_WAIT(mbar_wait(&w_empty[ws], wph ^ 1));
              if (elect_one()) {
                if (MB_DBG(1)) {
                  mbar_arrive(&w_full[ws]);
                } else if constexpr (PAIR) {   // this CTA's half of the weight rows
                  if (rank == 0) mbar_expect_tx(&w_full[ws], 2 * nt * W_TILE);
                  for (int j = 0; j < nt; ++j)
                    tma_load_2d_pair(w_base + ws * W_STAGE + j * W_TILE, &tmW, &w_full[ws], g.g_wk[q][tp + j] + c * KC,
                                     phase * g.N + n_tile * NTILE + rank * (NTILE / 2));
                } else {
                  mbar_expect_tx(&w_full[ws], nt * W_TILE);
                  for (int j = 0; j < nt; ++j)
                    tma_load_2d(w_base + ws * W_STAGE + j * W_TILE, &tmW, &w_full[ws], g.g_wk[q][tp + j] + c * KC,
                                phase * g.N + n_tile * NTILE);
                }
              }
              __syncwarp();
              if (++ws == g.w_stages) { ws = 0; wph ^= 1; }
            }
          }
        }
      }
    }
    MB_TEND(0);
  } else if (warp == 1) {
    // ================================================================= MMA issuer (converged warp, elected issue; pair: leader only)
    if (rank == 0) {
    constexpr uint32_t idesc = umma_idesc_f16(std::is_same<T, __nv_bfloat16>::value ? 1 : 0, PAIR ? 256 : 128, NTILE);
    auto mma = [](uint32_t d, uint32_t a, uint32_t b, uint32_t hi, uint32_t id, uint32_t acc) {
      if constexpr (PAIR) umma_f16_lohi_pair(d, a, b, hi, id, acc); else umma_f16_lohi(d, a, b, hi, id, acc);
    };
    auto commit = [](uint64_t* bar) { if constexpr (PAIR) umma_commit_pair(bar); else umma_commit(bar); };
    constexpr uint32_t W16 = W_TILE >> 4, BLK16 = (128 * ROWB) >> 4;
    const uint32_t desc_hi = static_cast<uint32_t>(umma_smem_desc<ROWB>(0) >> 32);
    const uint32_t a_lo0 = static_cast<uint32_t>(umma_smem_desc<ROWB>(smem_u32(a_ring)));
    const uint32_t w_lo0 = static_cast<uint32_t>(umma_smem_desc<ROWB>(smem_u32(w_base)));
    const uint32_t a_stage16 = static_cast<uint32_t>(A_BYTES) >> 4, w_stage16 = static_cast<uint32_t>(W_STAGE) >> 4;
    const int n_steps = g.n_steps, resident = g.w_resident;
    griddep_wait();
    if (resident) { mbar_wait(wres, 0); tc_fence_after(); }
    int as = 0, ws = 0, slot = 0;
    uint32_t aph = 0, wph = 0, tph = 0;
    MB_T0();
    for (int tl = PAIR ? (blockIdx.x >> 1) : blockIdx.x; (PAIR ? 2 * tl : tl) < g.total_tiles; tl += PAIR ? (gridDim.x >> 1) : gridDim.x) {
      const int tile = PAIR ? 2 * tl : tl;   // pair: both tiles share (n tile, phase) (np == 1 in pair mode)
      const int nn = tile % np;
      const uint32_t* sch = g.sched + (nn / g.n_tiles) * n_steps;
      MB_WAIT(mbar_wait(&t_empty[slot], tph ^ 1));
      const uint32_t dtm = tmem_base + slot * SLOT_COLS;
      const uint32_t wres_lo = w_lo0 + static_cast<uint32_t>(nn * kchunks) * W16;
      uint32_t a_lo = 0;
#pragma unroll 1
      for (int s = 0; s < n_steps; s += 2) {   // two schedule entries (taps) per iteration
        const uint32_t en0 = sch[s], en1 = sch[s + 1];
        const bool two = !(en1 & kSchedNull);
        if (en0 & kSchedFirst) {
          MB_WAIT(mbar_wait(&a_full[as], aph));
          a_lo = a_lo0 + as * a_stage16;
        }
        uint32_t w_lo_0, w_lo_1;
        int ws1 = ws;          // ring stage of the second entry (== ws when both taps share one stage)
        uint32_t wph1 = wph;
        if (resident) {
          w_lo_0 = wres_lo + ((en0 >> 16) & 0x7ffu) * W16;
          w_lo_1 = wres_lo + ((en1 >> 16) & 0x7ffu) * W16;
        } else {
          MB_WAIT(mbar_wait(&w_full[ws], wph));
          w_lo_0 = w_lo0 + ws * w_stage16 + ((en0 >> 16) & 0x7ffu) * W16;
          if (two && (en1 & kSchedWFirst)) {
            if (++ws1 == g.w_stages) { ws1 = 0; wph1 ^= 1; }
            MB_WAIT(mbar_wait(&w_full[ws1], wph1));
          }
          w_lo_1 = w_lo0 + ws1 * w_stage16 + ((en1 >> 16) & 0x7ffu) * W16;
        }
        tc_fence_after();
        if (elect_one()) {
          const uint32_t a0 = a_lo + (en0 & 0xffffu), a1 = a_lo + (en1 & 0xffffu);
          const uint32_t acc0 = s != 0;
          // The weight tile is reused by every 128-row block of the tile; consecutive MMAs go to DIFFERENT accumulators
          // (k outer, block inner): back-to-back accumulation into one TMEM tile stalls on the previous MMA at narrow N.
#pragma unroll
          for (int k = 0; k < KC / 16; ++k) {
#pragma unroll
            for (int i = 0; i < MB; ++i)
              if (!MB_DBG(4)) mma(dtm + i * NTILE, a0 + i * BLK16 + 2 * k, w_lo_0 + 2 * k, desc_hi, idesc, k == 0 ? acc0 : 1u);
          }
          if (!resident && (en0 & kSchedWLast)) commit(&w_empty[ws]);
          if (two) {
#pragma unroll
            for (int k = 0; k < KC / 16; ++k) {
#pragma unroll
              for (int i = 0; i < MB; ++i)
                if (!MB_DBG(4)) mma(dtm + i * NTILE, a1 + i * BLK16 + 2 * k, w_lo_1 + 2 * k, desc_hi, idesc, 1u);
            }
            if (!resident && (en1 & kSchedWLast)) commit(&w_empty[ws1]);
          }
          if ((en0 | en1) & kSchedLast) commit(&a_empty[as]);
          if (s + 2 >= n_steps) commit(&t_full[slot]);
        }
        __syncwarp();
        if (!resident) {
          ws = ws1; wph = wph1;
          if ((two ? en1 : en0) & kSchedWLast) { if (++ws == g.w_stages) { ws = 0; wph ^= 1; } }
        }
        if ((en0 | en1) & kSchedLast) { if (++as == g.a_stages) { as = 0; aph ^= 1; } }
      }
      if (++slot == NACC) { slot = 0; tph ^= 1; }
    }
    MB_TEND(1);
    }
  } else {
    // ================================================================= epilogue: warps 2-5 (group 0), 6-9 (group 1)
    const int grp = (warp - 2) >> 2;
    const int lane_grp = warp & 3;         // TMEM lanes [32*(warp%4), +32)
    uint8_t* stg = stage_base + (warp - 2) * kStageBytes;
    griddep_wait();
    int slot = 0, it = 0;
    uint32_t tph = 0;
    MB_T0E();
    for (int tl = PAIR ? (blockIdx.x >> 1) : blockIdx.x; (PAIR ? 2 * tl : tl) < g.total_tiles; tl += PAIR ? (gridDim.x >> 1) : gridDim.x, ++it) {
      const int tile_raw = PAIR ? 2 * tl + rank : tl;
      const bool ghost = tile_raw >= g.total_tiles;   // pair only: this CTA has no tile in the last round
      const int tile = ghost ? g.total_tiles - 1 : tile_raw;
      const int nn = tile % np, mm = tile / np;
      const int n_tile = nn % g.n_tiles, phase = nn / g.n_tiles;
      const int t0 = (mm % g.m_tiles) * TROWS, b = mm / g.m_tiles;
      const uint32_t taddr = tmem_base + (static_cast<uint32_t>(lane_grp * 32) << 16) + slot * SLOT_COLS;
#ifdef MB200_PROBE
      if (MB_DBG(8) || (PAIR && ghost)) {
        MB_WAIT(mbar_wait(&t_full[slot], tph));
      } else {
        epilogue_tile3<NTILE, MB, EG, T, F>(e, b, t0, g.T_out, phase, n_tile * NTILE, taddr, stg, lnx + (it & 1) * (EG * 128), lane,
                                 lane_grp, grp, &t_full[slot], tph, probe_e);
      }
#else
      if (PAIR && ghost) {
        mbar_wait(&t_full[slot], tph);     // consume the phase; the accumulator holds a copy of the last tile: nothing to store
      } else {
        epilogue_tile3<NTILE, MB, EG, T, F>(e, b, t0, g.T_out, phase, n_tile * NTILE, taddr, stg, lnx + (it & 1) * (EG * 128), lane,
                                 lane_grp, grp, &t_full[slot], tph);
      }
#endif
      tc_fence_before();
      __syncwarp();
      if (lane == 0) { if constexpr (PAIR) mbar_arrive_leader(&t_empty[slot]); else mbar_arrive(&t_empty[slot]); }
      if (++slot == NACC) { slot = 0; tph ^= 1; }
    }
    if constexpr ((F & kF_STATS) != 0) {
      if (e.stats_out && lane == 0) gn_stats_publish();   // this lane stored the warp's partial GroupNorm sums
    }
    if (warp == 2) { MB_TEND(2); }
#ifdef MB200_PROBE
    if (warp == 2 && lane == 0 && g.dbg_out) { for (int k = 1; k < 6; ++k) g.dbg_out[blockIdx.x * 16 + 9 + k] = probe_e[k]; }
#endif
    if (warp == 6) { MB_TEND(3); }
  }
  tc_fence_before();
  if constexpr (PAIR) cluster_sync_all(); else __syncthreads();   // pair: nothing of the peer's may still be in flight towards this CTA
  if constexpr ((F & kF_STATS) != 0) {
    if (e.stats_out && !e.stats_nofinish) gn_stats_finish(e, g.B, gridDim.x, reinterpret_cast<int*>(tmem_slot + 1));   // ordered GroupNorm sums by the last CTA to get here
  }
#ifdef MB200_PROBE
  if (threadIdx.x == 0 && g.dbg_out) { g.dbg_out[blockIdx.x * 16 + 8] = clock64() - probe_k0; g.dbg_out[blockIdx.x * 16 + 9] = probe_k1 - probe_k0; }
#endif
  if (warp == 1) { if constexpr (PAIR) tmem_dealloc_pair<TMEM_COLS>(tmem_base); else tmem_dealloc<TMEM_COLS>(tmem_base); }
}

template <int NTILE, int KC, int MB, int EG, typename T, uint32_t F, bool PAIR = false>
static cudaError_t launch_umma2_inst(const ConvLaunch& L, cudaStream_t st) {
  auto kern = conv_gemm_umma2_kernel<NTILE, KC, MB, EG, T, F, PAIR>;
  static PerDeviceOnce configured;
  if (configured.needed()) {
    cudaError_t err = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
    if (err != cudaSuccess) return err;
    configured.mark();
  }
  const bool two = L.g.n_src > 1;
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3(L.grid);
  cfg.blockDim = dim3(64 + 128 * EG);
  cfg.dynamicSmemBytes = L.smem_bytes;
  cfg.stream = st;
  cudaLaunchAttribute attr[2];
  int na = 0;
  if (L.pdl) {
    attr[na].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[na].val.programmaticStreamSerializationAllowed = 1;
    ++na;
  }
  if (PAIR) {   // CTA pairs: clusters of two (grid is even)
    attr[na].id = cudaLaunchAttributeClusterDimension;
    attr[na].val.clusterDim.x = 2; attr[na].val.clusterDim.y = 1; attr[na].val.clusterDim.z = 1;
    ++na;
  }
  cfg.attrs = attr;
  cfg.numAttrs = na;
  return cudaLaunchKernelEx(&cfg, kern, L.tmA[0], L.tmA[two ? 1 : 0], L.tmA8[0], L.tmA8[two ? 1 : 0], L.tmW, L.g2, L.e);
}

// Picks the leanest instantiated epilogue feature class (epilogue3.cuh) that covers what this launch asks for.
template <int NTILE, int KC, int MB, int EG, typename T>
static cudaError_t launch_umma2_feat(const ConvLaunch& L, cudaStream_t st) {
  const uint32_t need = epi_features(L.e);
  if constexpr ((NTILE == 256 && MB == 1) || (NTILE == 128 && MB == 2 && EG == 2)) {
    if (L.g2.pair) {
#define MB_TRYP(CLS) if ((need & ~(CLS)) == 0) return launch_umma2_inst<NTILE, KC, MB, EG, T, (CLS), true>(L, st)
      if constexpr (NTILE == 256 && EG == 3) { MB_TRYP(kFC_STATS); MB_TRYP(kFC_SNAKE); MB_TRYP(kFC_RESLN); MB_TRYP(kFC_GNLN); }
      if constexpr (EG == 2) { MB_TRYP(kFC_VOC); }
      if constexpr (NTILE == 256 && EG == 2) { MB_TRYP(kFC_RESMASK); }
#undef MB_TRYP
      return launch_umma2_inst<NTILE, KC, MB, EG, T, kF_ALL, true>(L, st);
    }
  }
#define MB_TRY(CLS) if ((need & ~(CLS)) == 0) return launch_umma2_inst<NTILE, KC, MB, EG, T, (CLS)>(L, st)
  if constexpr (NTILE == 256 && MB == 1 && EG == 3) { MB_TRY(kFC_STATS); MB_TRY(kFC_SNAKE); MB_TRY(kFC_RESLN); MB_TRY(kFC_GNLN); }
  if constexpr (EG == 2) { MB_TRY(kFC_VOC); }
  if constexpr (NTILE == 256 && MB == 1 && EG == 2) { MB_TRY(kFC_RESMASK); }
#undef MB_TRY
  return launch_umma2_inst<NTILE, KC, MB, EG, T, kF_ALL>(L, st);
}

template <int NTILE, int KC, int MB, typename T>
static cudaError_t launch_umma2_eg(const ConvLaunch& L, cudaStream_t st) {
  if constexpr (MB == 1 && NTILE >= 128) {   // epilogue-bound estimator layers: twelve epilogue warps
    if (L.g2.eg == 3) return launch_umma2_feat<NTILE, KC, MB, 3, T>(L, st);
  }
  return launch_umma2_feat<NTILE, KC, MB, 2, T>(L, st);
}

template <typename T>
cudaError_t launch_conv_umma2(const ConvLaunch& L, cudaStream_t st) {
  const int nt = L.ntile, kc = L.kc, mb = L.g2.mb;
  if (kc == 64 && nt == 256 && mb == 1) return launch_umma2_eg<256, 64, 1, T>(L, st);
  if (kc == 64 && nt == 128 && mb == 1) return launch_umma2_eg<128, 64, 1, T>(L, st);
  if (kc == 64 && nt == 128 && mb == 2) return launch_umma2_eg<128, 64, 2, T>(L, st);
  if (kc == 64 && nt == 64 && mb == 1) return launch_umma2_eg<64, 64, 1, T>(L, st);
  if (kc == 64 && nt == 64 && mb == 2) return launch_umma2_eg<64, 64, 2, T>(L, st);
  if (kc == 64 && nt == 32 && mb == 1) return launch_umma2_eg<32, 64, 1, T>(L, st);
  if (kc == 64 && nt == 32 && mb == 4) return launch_umma2_eg<32, 64, 4, T>(L, st);
  if (kc == 32 && nt == 32 && mb == 1) return launch_umma2_eg<32, 32, 1, T>(L, st);
  if (kc == 32 && nt == 32 && mb == 4) return launch_umma2_eg<32, 32, 4, T>(L, st);
  return cudaErrorInvalidConfiguration;
}
template cudaError_t launch_conv_umma2<__nv_bfloat16>(const ConvLaunch&, cudaStream_t);
template cudaError_t launch_conv_umma2<__half>(const ConvLaunch&, cudaStream_t);

}  // namespace mb
