// The second half of a ResnetBlock1D and the first half of the transformer block that follows it, as ONE persistent
// tcgen05 kernel (decoder.py:54-60 res_conv + block2 tail, transformer.py:262-271 norm1 + to_q/to_k/to_v):
//
//     res = x Wres^T + b + mish(GroupNorm(raw2)) * mask          -> HBM  (the block output, residual of the attention)
//     qkv = LayerNorm1(res) [Wq/8 | Wk | Wv]^T                    -> HBM  (attention input)
//
// Run as two conv-GEMM launches, the first is a K <= 512 GEMM whose epilogue (GroupNorm apply, Mish, LayerNorm) is 10x
// longer than its MMAs with one or two tiles per CTA, and the LayerNorm output makes a round trip through HBM for a
// K = 256 GEMM.  Here, per tile of 128 rows of the flat B*L frame index:
//     TMA   x chunks [128][64] (1 or 2 sources = channel concat) + Wres chunks [256][64]  -> rings
//     M0    TMEM[0..255]  = x . Wres^T                                   (N = 256, K = 192 / 256 / 512)
//     E0    + bias + mish(GN(raw2)) * mask -> res (HBM); LayerNorm1 (row sums exchanged between the two warps of a row)
//           -> x tile in the UMMA K-major 128B-swizzled layout (shared memory)
//     M1    TMEM[0..383]  = x_tile . Wqkv^T                              (3 x N = 128, K = 256)
//     E1    -> qkv (HBM)
// Same MMA order and epilogue arithmetic as the two launches it replaces (LayerNorm row sums are added in another order).
//
// Warp roles (320 threads): 0 TMA producer, 1 MMA issuer, 2-9 epilogue (warp % 4 = TMEM lane quarter, (warp-2)/4 = column half).
#include "conv.cuh"
#include "epilogue3.cuh"
#include "launch.h"
#include "ptx.cuh"
#include "resqkv_epilogue.cuh"

namespace mb {

namespace {
constexpr int kRQAStage = 16384, kRQAStages = 3;   // x chunk ring
constexpr int kRQWStage = 32768, kRQWStages = 3;   // weight ring
constexpr int kRQXBytes = 4 * 16384;               // LayerNorm output tile (A operand of M1)
constexpr int kRQXchBytes = 2048;                  // LayerNorm row-sum exchange [2 halves][128 rows] float2
constexpr int kRQBarBytes = 512;
constexpr int kRQParBytes = 5 * 1024;               // bias, GroupNorm gamma/beta, LayerNorm gamma/beta: [5][256] fp32
}  // namespace

int resqkv_fused_smem_bytes() {
  return 1024 + kRQAStages * kRQAStage + kRQWStages * kRQWStage + kRQXBytes + kRQXchBytes + kRQBarBytes + kRQParBytes;
}

// PAIR: two CTAs of a cluster as one tcgen05 cta_group::2 unit (conv_gemm2.cu): each CTA runs its own 128-row tile and holds
// half of every weight box (res_conv: 128 of 256 output channels, QKV: 64 of a block's 128), the even CTA issues the MMAs.
template <typename T, bool PAIR>
__global__ void __launch_bounds__(320, 1)
resqkv_fused_kernel(const __grid_constant__ CUtensorMap tmA0, const __grid_constant__ CUtensorMap tmA1,
                    const __grid_constant__ CUtensorMap tmWres, const __grid_constant__ CUtensorMap tmWqkv,
                    const __grid_constant__ RQGeom g) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  uint8_t* a_ring = smem;
  uint8_t* w_ring = a_ring + kRQAStages * kRQAStage;
  uint8_t* x_tile = w_ring + kRQWStages * kRQWStage;
  float2* xch = reinterpret_cast<float2*>(x_tile + kRQXBytes);
  uint64_t* bars = reinterpret_cast<uint64_t*>(x_tile + kRQXBytes + kRQXchBytes);
  uint64_t* a_full = bars;            // [3]
  uint64_t* a_empty = bars + 3;       // [3]
  uint64_t* w_full = bars + 6;        // [3]
  uint64_t* w_empty = bars + 9;       // [3]
  uint64_t* t0_full = bars + 12;      // M0 complete
  uint64_t* x_full = bars + 13;       // LayerNorm tile written (8 warps)
  uint64_t* t1_full = bars + 14;      // M1 complete
  uint64_t* t_free = bars + 15;       // E1 has read the accumulators (8 warps)
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 16);
  float* par = reinterpret_cast<float*>(x_tile + kRQXBytes + kRQXchBytes + kRQBarBytes);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int rank = PAIR ? static_cast<int>(cluster_ctarank()) : 0;
  const int units = PAIR ? (g.total_tiles + 1) / 2 : g.total_tiles;   // tiles, or pairs of tiles (2 u, 2 u + 1)
  const int u0 = PAIR ? static_cast<int>(blockIdx.x >> 1) : static_cast<int>(blockIdx.x);
  const int ustep = PAIR ? static_cast<int>(gridDim.x >> 1) : static_cast<int>(gridDim.x);
  const int nt_cta = units > u0 ? (units - 1 - u0) / ustep + 1 : 0;
  const int nk = g.nk0 + g.nk1;
  constexpr int WSTAGE = PAIR ? kRQWStage / 2 : kRQWStage;            // this CTA's share of a weight ring stage
  auto expect = [&](uint64_t* bar, uint32_t bytes) { if constexpr (PAIR) { if (rank == 0) mbar_expect_tx(bar, 2 * bytes); } else mbar_expect_tx(bar, bytes); };
  auto load2 = [](void* dst, const CUtensorMap* m, uint64_t* bar, int c0, int c1) {
    if constexpr (PAIR) tma_load_2d_pair(dst, m, bar, c0, c1); else tma_load_2d(dst, m, bar, c0, c1);
  };
  auto load3 = [](void* dst, const CUtensorMap* m, uint64_t* bar, int c0, int c1, int c2) {
    if constexpr (PAIR) tma_load_3d_pair(dst, m, bar, c0, c1, c2); else tma_load_3d(dst, m, bar, c0, c1, c2);
  };

  if (threadIdx.x == 0) {
    tma_prefetch_desc(&tmA0); tma_prefetch_desc(&tmA1); tma_prefetch_desc(&tmWres); tma_prefetch_desc(&tmWqkv);
    for (int s = 0; s < 3; ++s) { mbar_init(&a_full[s], 1); mbar_init(&a_empty[s], 1); mbar_init(&w_full[s], 1); mbar_init(&w_empty[s], 1); }
    mbar_init(t0_full, 1); mbar_init(x_full, PAIR ? 16 : 8); mbar_init(t1_full, 1); mbar_init(t_free, PAIR ? 16 : 8);
    fence_barrier_init();
  }
  // per-channel parameters -> shared memory once per CTA: with 217 KB of the SM's 256 KB carved out as shared memory the
  // L1 is too small to keep them against the row-per-lane global traffic (measured: every parameter load in the epilogue
  // went to L2, 2 100 cycles per 8-channel group)
  for (int i = threadIdx.x; i < 5 * 256; i += 320) {
    const float* src = i < 256 ? g.bias : (i < 512 ? g.gn_gamma : (i < 768 ? g.gn_beta : (i < 1024 ? g.ln_g : g.ln_b)));
    par[i] = __ldg(src + (i & 255));
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
    for (int n = 0; n < nt_cta; ++n) {
      const int tile = PAIR ? 2 * (u0 + n * ustep) + rank : u0 + n * ustep;
      for (int c = 0; c < nk; ++c) {
        const bool second = c >= g.nk0;
        mbar_wait(&a_empty[as], aph ^ 1);
        mbar_wait(&w_empty[ws], wph ^ 1);
        if (elect_one()) {
          expect(&a_full[as], kRQAStage);
          load3(a_ring + as * kRQAStage, second ? &tmA1 : &tmA0, &a_full[as], (second ? c - g.nk0 : c) * 64, tile * 128, 0);
          expect(&w_full[ws], WSTAGE);
          load2(w_ring + ws * WSTAGE, &tmWres, &w_full[ws], c * 64, PAIR ? rank * 128 : 0);
        }
        __syncwarp();
        if (++as == kRQAStages) { as = 0; aph ^= 1; }
        if (++ws == kRQWStages) { ws = 0; wph ^= 1; }
      }
      for (int nb = 0; nb < 3; ++nb)        // Wqkv rows [128 nb, +128): two stages of two [128 n][64 k] boxes
        for (int s = 0; s < 2; ++s) {
          mbar_wait(&w_empty[ws], wph ^ 1);
          if (elect_one()) {
            expect(&w_full[ws], WSTAGE);
            const int n0 = nb * 128 + (PAIR ? rank * 64 : 0);
            load2(w_ring + ws * WSTAGE, &tmWqkv, &w_full[ws], (2 * s) * 64, n0);
            load2(w_ring + ws * WSTAGE + WSTAGE / 2, &tmWqkv, &w_full[ws], (2 * s + 1) * 64, n0);
          }
          __syncwarp();
          if (++ws == kRQWStages) { ws = 0; wph ^= 1; }
        }
    }
  } else if (warp == 1) {
    // ================================================================= MMA issuer (pair: the leader's only)
    if (rank == 0) {
    constexpr int fmt = std::is_same<T, __nv_bfloat16>::value ? 1 : 0;
    constexpr int MM = PAIR ? 256 : 128;
    constexpr uint32_t idesc128 = umma_idesc_f16(fmt, MM, 128), idesc256 = umma_idesc_f16(fmt, MM, 256);
    auto umma_f16_lohi = [](uint32_t d, uint32_t a, uint32_t b, uint32_t hi, uint32_t id, uint32_t acc) {
      if constexpr (PAIR) mb::umma_f16_lohi_pair(d, a, b, hi, id, acc); else mb::umma_f16_lohi(d, a, b, hi, id, acc);
    };
    auto umma_commit = [](uint64_t* bar) { if constexpr (PAIR) mb::umma_commit_pair(bar); else mb::umma_commit(bar); };
    const uint32_t desc_hi = static_cast<uint32_t>(umma_smem_desc<128>(0) >> 32);
    const uint32_t a_lo0 = static_cast<uint32_t>(umma_smem_desc<128>(smem_u32(a_ring)));
    const uint32_t x_lo = static_cast<uint32_t>(umma_smem_desc<128>(smem_u32(x_tile)));
    const uint32_t w_lo0 = static_cast<uint32_t>(umma_smem_desc<128>(smem_u32(w_ring)));
    int as = 0, ws = 0;
    uint32_t aph = 0, wph = 0;
    for (int n = 0; n < nt_cta; ++n) {
      mbar_wait(t_free, (n & 1) ^ 1);       // the previous tile's E1 no longer reads TMEM
      tc_fence_after();
      for (int c = 0; c < nk; ++c) {        // ---- M0
        mbar_wait(&a_full[as], aph);
        mbar_wait(&w_full[ws], wph);
        tc_fence_after();
        if (elect_one()) {
          const uint32_t a = a_lo0 + as * (kRQAStage >> 4), w = w_lo0 + ws * (WSTAGE >> 4);
#pragma unroll
          for (int k = 0; k < 4; ++k) umma_f16_lohi(tmem_base, a + 2 * k, w + 2 * k, desc_hi, idesc256, (c | k) != 0 ? 1u : 0u);
          umma_commit(&a_empty[as]);
          umma_commit(&w_empty[ws]);
          if (c == nk - 1) umma_commit(t0_full);
        }
        __syncwarp();
        if (++as == kRQAStages) { as = 0; aph ^= 1; }
        if (++ws == kRQWStages) { ws = 0; wph ^= 1; }
      }
      mbar_wait(x_full, n & 1);             // ---- M1: three 128-column blocks of q | k | v
      tc_fence_after();
      for (int nb = 0; nb < 3; ++nb)
        for (int s = 0; s < 2; ++s) {
          mbar_wait(&w_full[ws], wph);
          tc_fence_after();
          if (elect_one()) {
            const uint32_t w = w_lo0 + ws * (WSTAGE >> 4);
#pragma unroll
            for (int c = 0; c < 2; ++c) {
              const uint32_t a = x_lo + (2 * s + c) * 1024u, wc = w + c * ((WSTAGE / 2) >> 4);
#pragma unroll
              for (int k = 0; k < 4; ++k) umma_f16_lohi(tmem_base + nb * 128, a + 2 * k, wc + 2 * k, desc_hi, idesc128, (s | c | k) != 0 ? 1u : 0u);
            }
            umma_commit(&w_empty[ws]);
            if (nb == 2 && s == 1) umma_commit(t1_full);
          }
          __syncwarp();
          if (++ws == kRQWStages) { ws = 0; wph ^= 1; }
        }
    }
    }
  } else {
    rq_epilogue<T, PAIR>(g, par, x_tile, xch, t0_full, x_full, t1_full, t_free, tmem_base, warp, lane, nt_cta, u0, ustep, rank);
  }
  tc_fence_before();
  if constexpr (PAIR) cluster_sync_all(); else __syncthreads();
  if (warp == 1) { if constexpr (PAIR) tmem_dealloc_pair<512>(tmem_base); else tmem_dealloc<512>(tmem_base); }
}

template <typename T, bool PAIR>
static cudaError_t launch_resqkv_inst(const RQLaunch& L, cudaStream_t st) {
  auto kern = resqkv_fused_kernel<T, PAIR>;
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
  return cudaLaunchKernelEx(&cfg, kern, L.tmA0, L.tmA1, L.tmWres, L.tmWqkv, L.g);
}

template <typename T>
cudaError_t launch_resqkv_fused(const RQLaunch& L, cudaStream_t st) {
  return L.pair ? launch_resqkv_inst<T, true>(L, st) : launch_resqkv_inst<T, false>(L, st);
}
template cudaError_t launch_resqkv_fused<__nv_bfloat16>(const RQLaunch&, cudaStream_t);
template cudaError_t launch_resqkv_fused<__half>(const RQLaunch&, cudaStream_t);

}  // namespace mb
