// Persistent tcgen05 conv-GEMM (the production tensor-core kernel).
//
// One CTA per SM loops over output tiles (128 rows x NTILE channels).  Warp roles:
//   warp 0      TMA producer: per K chunk ONE halo load of the A tile (128 + 8*extra rows, 128B/64B swizzle) that all
//               taps of the conv reuse, plus the weight tiles (streamed, or loaded once and kept resident when the
//               whole weight matrix fits beside the A ring -- the 64/32-channel HiFi-GAN stages).
//   warp 1      TMEM owner + single-thread tcgen05.mma issuer; taps address the halo tile through row-shifted
//               shared-memory descriptors.  Accumulators are a ring of NACC TMEM buffers.
//   warps 2-5, 6-9   two epilogue groups that drain alternate tiles (tcgen05.ld -> fused epilogue -> global), so the
//               epilogue of tile i overlaps the loads and MMAs of tiles i+1, i+2.
#include "conv.cuh"
#include "epilogue2.cuh"
#include "launch.h"
#include "ptx.cuh"

namespace mb {

constexpr int kARows = 192;  // halo capacity: 128 + 8*8 rows (largest span on the path is 50: k=11, dilation 5)
constexpr int kMaxRing = 8;

template <int NTILE, typename T>
__device__ __forceinline__ void epilogue_tile(const EpiParams& e, int b, int t, bool valid, int phase, int nbase,
                                              uint32_t taddr) {
  float v[32];
  if (e.ln_out == nullptr || NTILE != 256) {
#pragma unroll 1
    for (int c = 0; c < NTILE / 32; ++c) {
      tmem_ld_32x32(taddr + c * 32, v);
      tmem_ld_wait();
      epilogue_chunk<T>(e, b, t, phase, nbase + c * 32, valid, true, v);
    }
  } else {
    float s = 0.f, ss = 0.f;
#pragma unroll 1
    for (int c = 0; c < NTILE / 32; ++c) {
      tmem_ld_32x32(taddr + c * 32, v);
      tmem_ld_wait();
      epilogue_chunk<T>(e, b, t, phase, nbase + c * 32, valid, true, v);
#pragma unroll
      for (int i = 0; i < 32; ++i) { s += v[i]; ss += v[i] * v[i]; }
    }
    const float mean = s * (1.f / 256.f);
    const float rstd = rsqrtf(fmaxf(ss * (1.f / 256.f) - mean * mean, 0.f) + 1e-5f);
    const long row = static_cast<long>(b) * e.T_o + t * e.o_mul + phase * e.o_phase_step;
#pragma unroll 1
    for (int c = 0; c < NTILE / 32; ++c) {
      tmem_ld_32x32(taddr + c * 32, v);
      tmem_ld_wait();
      epilogue_chunk<T>(e, b, t, phase, nbase + c * 32, valid, false, v);
      if (valid) {
        T* dst = reinterpret_cast<T*>(e.ln_out) + row * e.ln_ld + c * 32;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          float o[8];
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            const int n = c * 32 + 8 * j + i;
            o[i] = (v[8 * j + i] - mean) * rstd * __ldg(e.ln_g + n) + __ldg(e.ln_b + n);
          }
          Elem<T>::st8(dst + 8 * j, o);
        }
      }
    }
  }
}

template <int NTILE, int KC, typename T>
__global__ void __launch_bounds__(320, 1)
conv_gemm_umma2_kernel(const __grid_constant__ CUtensorMap tmA0, const __grid_constant__ CUtensorMap tmA1,
                       const __grid_constant__ CUtensorMap tmA0x, const __grid_constant__ CUtensorMap tmA1x,
                       const __grid_constant__ CUtensorMap tmW, const __grid_constant__ ConvGeom2 g,
                       const __grid_constant__ EpiParams e) {
  constexpr int ROWB = KC * 2;
  constexpr int W_TILE = NTILE * ROWB;
  constexpr int TMEM_COLS = 512;   // one CTA per SM: take all of TMEM
  const int A_BYTES = g.a_bytes;
  const int SLOT_COLS = g.mb * NTILE;
  const int NACC = (512 / SLOT_COLS) < 4 ? (512 / SLOT_COLS) : 4;
  const int TROWS = 128 * g.mb;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));

  int kchunks = 0;  // K chunks per (phase, n_tile): sum over groups of chunks * taps
  for (int q = 0; q < g.n_groups; ++q) kchunks += g.g_chunks[q] * g.g_ntaps[q];
  const int w_tiles = g.w_resident ? g.n_phases * g.n_tiles * kchunks : g.w_stages;
  uint8_t* a_ring = smem;
  uint8_t* w_base = smem + g.a_stages * A_BYTES;
  uint64_t* bars = reinterpret_cast<uint64_t*>(w_base + static_cast<size_t>(w_tiles) * W_TILE);
  uint64_t* a_full = bars;
  uint64_t* a_empty = a_full + kMaxRing;
  uint64_t* w_full = a_empty + kMaxRing;
  uint64_t* w_empty = w_full + kMaxRing;
  uint64_t* t_full = w_empty + kMaxRing;
  uint64_t* t_empty = t_full + 4;
  uint64_t* wres = t_empty + 4;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(wres + 1);
  // 8 warp-private epilogue staging tiles, 16-byte aligned
  uint8_t* stage_base = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(wres + 2) + 15) & ~uintptr_t(15));

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int np = g.n_tiles * g.n_phases;

  if (threadIdx.x == 0) {
    tma_prefetch_desc(&tmA0); tma_prefetch_desc(&tmA0x); tma_prefetch_desc(&tmW);
    for (int s = 0; s < kMaxRing; ++s) { mbar_init(&a_full[s], 1); mbar_init(&a_empty[s], 1); mbar_init(&w_full[s], 1); mbar_init(&w_empty[s], 1); }
    for (int s = 0; s < 4; ++s) { mbar_init(&t_full[s], 1); mbar_init(&t_empty[s], 4); }
    mbar_init(wres, 1);
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc<TMEM_COLS>(tmem_slot);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    // ================================================================= TMA producer
    if (lane == 0) {
      if (g.w_resident) {
        mbar_expect_tx(wres, static_cast<uint32_t>(w_tiles) * W_TILE);
        for (int pn = 0; pn < np; ++pn)   // pn = phase * n_tiles + n_tile
          for (int kk = 0; kk < kchunks; ++kk)
            tma_load_2d(w_base + static_cast<size_t>(pn * kchunks + kk) * W_TILE, &tmW, wres, kk * KC,
                        (pn / g.n_tiles) * g.N + (pn % g.n_tiles) * NTILE);
      }
      int as = 0, ws = 0;
      uint32_t aph = 0, wph = 0;
      for (int tile = blockIdx.x; tile < g.total_tiles; tile += gridDim.x) {
        const int nn = tile % np, mm = tile / np;
        const int n_tile = nn % g.n_tiles, phase = nn / g.n_tiles;
        const int t0 = (mm % g.m_tiles) * TROWS, b = mm / g.m_tiles;
        for (int q = 0; q < g.n_groups; ++q) {
          const CUtensorMap* m128 = g.g_src[q] ? &tmA1 : &tmA0;
          const CUtensorMap* m8 = g.g_src[q] ? &tmA1x : &tmA0x;
          const int row0 = t0 + g.g_rowmin[phase][q], extra = g.g_extra[phase][q];
          for (int c = 0; c < g.g_chunks[q]; ++c) {
            mbar_wait(&a_empty[as], aph ^ 1);
            uint8_t* dst = a_ring + as * A_BYTES;
            mbar_expect_tx(&a_full[as], (TROWS + 8 * extra) * ROWB);
            const int ch = g.g_ch[q] + c * KC;
            for (int i = 0; i < g.mb; ++i) tma_load_3d(dst + i * 128 * ROWB, m128, &a_full[as], ch, row0 + 128 * i, b);
            for (int x = 0; x < extra; ++x) tma_load_3d(dst + (TROWS + 8 * x) * ROWB, m8, &a_full[as], ch, row0 + TROWS + 8 * x, b);
            if (++as == g.a_stages) { as = 0; aph ^= 1; }
            if (!g.w_resident) {
              for (int tp = 0; tp < g.g_ntaps[q]; ++tp) {
                mbar_wait(&w_empty[ws], wph ^ 1);
                mbar_expect_tx(&w_full[ws], W_TILE);
                tma_load_2d(w_base + ws * W_TILE, &tmW, &w_full[ws], g.g_wk[q][tp] + c * KC, phase * g.N + n_tile * NTILE);
                if (++ws == g.w_stages) { ws = 0; wph ^= 1; }
              }
            }
          }
        }
      }
    }
  } else if (warp == 1) {
    // ================================================================= MMA issuer
    constexpr uint32_t idesc = umma_idesc_f16(std::is_same<T, __nv_bfloat16>::value ? 1 : 0, 128, NTILE);
    if (g.w_resident) mbar_wait(wres, 0);
    int as = 0, ws = 0, it = 0;
    uint32_t aph = 0, wph = 0;
    for (int tile = blockIdx.x; tile < g.total_tiles; tile += gridDim.x, ++it) {
      const int nn = tile % np;
      const int phase = nn / g.n_tiles;
      const int slot = it % NACC;
      mbar_wait(&t_empty[slot], (((it / NACC) & 1) ^ 1));
      tc_fence_after();
      const uint32_t dtm = tmem_base + slot * SLOT_COLS;
      uint32_t first = 1;
      for (int q = 0; q < g.n_groups; ++q) {
        for (int c = 0; c < g.g_chunks[q]; ++c) {
          mbar_wait(&a_full[as], aph);
          tc_fence_after();
          const uint32_t a_addr = smem_u32(a_ring + as * A_BYTES);
          for (int tp = 0; tp < g.g_ntaps[q]; ++tp) {
            uint32_t w_addr;
            if (g.w_resident) {
              w_addr = smem_u32(w_base + static_cast<size_t>(nn * kchunks + (g.g_wk[q][tp] / KC) + c) * W_TILE);
            } else {
              mbar_wait(&w_full[ws], wph);
              tc_fence_after();
              w_addr = smem_u32(w_base + ws * W_TILE);
            }
            if (lane == 0) {
              const uint64_t bdesc = umma_smem_desc<ROWB>(w_addr);
              for (int i = 0; i < g.mb; ++i) {   // the weight tile is reused by every 128-row block of the tile
                const uint32_t sa = a_addr + (g.g_shift[phase][q][tp] + 128 * i) * ROWB;
                uint64_t adesc = umma_smem_desc<ROWB>(sa);
                if (g.base_off_mode) adesc |= static_cast<uint64_t>((sa >> 7) & 7) << 49;
#pragma unroll
                for (int k = 0; k < KC / 16; ++k)
                  umma_f16(dtm + i * NTILE, adesc + 2 * k, bdesc + 2 * k, idesc, (first && k == 0) ? 0u : 1u);
              }
              if (!g.w_resident) umma_commit(&w_empty[ws]);
            }
            __syncwarp();
            if (!g.w_resident) { if (++ws == g.w_stages) { ws = 0; wph ^= 1; } }
            first = 0;
          }
          if (lane == 0) umma_commit(&a_empty[as]);
          __syncwarp();
          if (++as == g.a_stages) { as = 0; aph ^= 1; }
        }
      }
      if (lane == 0) umma_commit(&t_full[slot]);
      __syncwarp();
    }
  } else {
    // ================================================================= epilogue groups
    const int grp = (warp - 2) >> 2;       // 0: warps 2-5, 1: warps 6-9
    const int lane_grp = warp & 3;         // TMEM lanes [32*(warp%4), +32)
    int it = grp;
    for (int tile = blockIdx.x + grp * gridDim.x; tile < g.total_tiles; tile += 2 * gridDim.x, it += 2) {
      const int nn = tile % np, mm = tile / np;
      const int n_tile = nn % g.n_tiles, phase = nn / g.n_tiles;
      const int t0 = (mm % g.m_tiles) * TROWS, b = mm / g.m_tiles;
      const int slot = it % NACC;
      mbar_wait(&t_full[slot], (it / NACC) & 1);
      tc_fence_after();
      for (int i = 0; i < g.mb; ++i) {
        const int tw = t0 + 128 * i + lane_grp * 32;
        if (tw >= g.T_out) break;   // warp-uniform: nothing of this block exists
        const uint32_t taddr = tmem_base + (static_cast<uint32_t>(lane_grp * 32) << 16) + slot * SLOT_COLS + i * NTILE;
        if (epi2_supported(e))
          epilogue_tile2<NTILE, T>(e, b, tw, g.T_out, phase, n_tile * NTILE, taddr, stage_base + (warp - 2) * kStageBytes, lane);
        else
          epilogue_tile<NTILE, T>(e, b, tw + lane, tw + lane < g.T_out, phase, n_tile * NTILE, taddr);
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&t_empty[slot]);
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc<TMEM_COLS>(tmem_base);
}

template <int NTILE, int KC, typename T>
static cudaError_t launch_umma2_inst(const ConvLaunch& L, cudaStream_t st) {
  auto kern = conv_gemm_umma2_kernel<NTILE, KC, T>;
  static int configured = 0;
  if (configured < L.smem_bytes) {
    cudaError_t err = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
    if (err != cudaSuccess) return err;
    configured = 227 * 1024;
  }
  const bool two = L.g.n_src > 1;
  kern<<<L.grid, 320, L.smem_bytes, st>>>(L.tmA[0], L.tmA[two ? 1 : 0], L.tmA8[0], L.tmA8[two ? 1 : 0], L.tmW, L.g2, L.e);
  return cudaGetLastError();
}

template <typename T>
cudaError_t launch_conv_umma2(const ConvLaunch& L, cudaStream_t st) {
  const int nt = L.ntile, kc = L.kc;
  if (kc == 64 && nt == 256) return launch_umma2_inst<256, 64, T>(L, st);
  if (kc == 64 && nt == 128) return launch_umma2_inst<128, 64, T>(L, st);
  if (kc == 64 && nt == 64) return launch_umma2_inst<64, 64, T>(L, st);
  if (kc == 64 && nt == 32) return launch_umma2_inst<32, 64, T>(L, st);
  if (kc == 32 && nt == 32) return launch_umma2_inst<32, 32, T>(L, st);
  return cudaErrorInvalidConfiguration;
}
template cudaError_t launch_conv_umma2<__nv_bfloat16>(const ConvLaunch&, cudaStream_t);
template cudaError_t launch_conv_umma2<__half>(const ConvLaunch&, cudaStream_t);

}  // namespace mb
