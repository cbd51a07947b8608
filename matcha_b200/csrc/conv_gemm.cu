// Conv-as-GEMM kernels for sm_100a.
//   conv_gemm_umma_kernel : TMA (cp.async.bulk.tensor, 128B/64B swizzle) -> shared memory ring -> tcgen05.mma
//                           (bf16/f16 operands, fp32 accumulators in TMEM) -> tcgen05.ld -> fused epilogue.
//                           Warp roles: warp 0 = TMA producer, warp 1 = TMEM owner + MMA issuer, warps 2..5 = epilogue.
//   conv_gemm_simt_kernel : plain FFMA implementation of the same contract (fp32 mode, and a device-side
//                           cross-check for the tensor-core kernel).
#include "conv.cuh"
#include "launch.h"
#include "ptx.cuh"

namespace mb {

// =================================================================================================== tcgen05
template <int NTILE, int KC, int STAGES>
struct UmmaSmem {
  static constexpr int A_BYTES = 128 * KC * 2;
  static constexpr int W_BYTES = NTILE * KC * 2;
  static constexpr int STAGE_BYTES = A_BYTES + W_BYTES;
  static constexpr int BAR_OFFSET = STAGES * STAGE_BYTES;
  static constexpr int TOTAL = BAR_OFFSET + (2 * STAGES + 1) * 8 + 16 + 1024;  // + alignment slack
};

template <int NTILE, int KC, int STAGES, typename T>
__global__ void __launch_bounds__(192)
conv_gemm_umma_kernel(const __grid_constant__ CUtensorMap tmA0, const __grid_constant__ CUtensorMap tmA1,
                      const __grid_constant__ CUtensorMap tmW, const __grid_constant__ ConvGeom g,
                      const __grid_constant__ EpiParams e) {
  using S = UmmaSmem<NTILE, KC, STAGES>;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint64_t* full = reinterpret_cast<uint64_t*>(smem + S::BAR_OFFSET);
  uint64_t* empty = full + STAGES;
  uint64_t* accum = empty + STAGES;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(accum + 1);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int n_tiles = g.N / NTILE;
  const int n_tile = blockIdx.y % n_tiles, phase = blockIdx.y / n_tiles;
  const int b = blockIdx.z;
  const int t0 = blockIdx.x * 128;
  const int chunks0 = g.src_ch[0] / KC;
  const int per_tap = chunks0 + (g.n_src > 1 ? g.src_ch[1] / KC : 0);
  const int num_k = g.n_taps * per_tap;

  if (threadIdx.x == 0) {
    tma_prefetch_desc(&tmA0);
    if (g.n_src > 1) tma_prefetch_desc(&tmA1);
    tma_prefetch_desc(&tmW);
    for (int s = 0; s < STAGES; ++s) { mbar_init(&full[s], 1); mbar_init(&empty[s], 1); }
    mbar_init(accum, 1);
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc<(NTILE < 32 ? 32 : NTILE)>(tmem_slot);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    if (lane == 0) {
      for (int it = 0; it < num_k; ++it) {
        const int s = it % STAGES;
        const uint32_t ph = (it / STAGES) & 1;
        mbar_wait(&empty[s], ph ^ 1);
        const int tap = it / per_tap, r = it - tap * per_tap;
        const bool second = r >= chunks0;
        const int c = second ? r - chunks0 : r;
        uint8_t* sa = smem + s * S::STAGE_BYTES;
        mbar_expect_tx(&full[s], S::STAGE_BYTES);
        tma_load_3d(sa, second ? &tmA1 : &tmA0, &full[s], g.tap_ch[phase][tap] + c * KC, t0 + g.tap_row[phase][tap], b);
        tma_load_2d(sa + S::A_BYTES, &tmW, &full[s], it * KC, phase * g.N + n_tile * NTILE);
      }
    }
  } else if (warp == 1) {
    constexpr uint32_t idesc = umma_idesc_f16(std::is_same<T, __nv_bfloat16>::value ? 1 : 0, 128, NTILE);
    for (int it = 0; it < num_k; ++it) {
      const int s = it % STAGES;
      const uint32_t ph = (it / STAGES) & 1;
      mbar_wait(&full[s], ph);
      tc_fence_after();
      if (lane == 0) {
        const uint32_t sa = smem_u32(smem + s * S::STAGE_BYTES);
        const uint64_t adesc = umma_smem_desc<KC * 2>(sa);
        const uint64_t bdesc = umma_smem_desc<KC * 2>(sa + S::A_BYTES);
#pragma unroll
        for (int k = 0; k < KC / 16; ++k)  // +32 bytes per K=16 step inside the swizzle span (>>4 => +2)
          umma_f16(tmem_base, adesc + 2 * k, bdesc + 2 * k, idesc, (it | k) != 0 ? 1u : 0u);
        umma_commit(&empty[s]);
        if (it == num_k - 1) umma_commit(accum);
      }
      __syncwarp();
    }
  } else {
    mbar_wait(accum, 0);
    tc_fence_after();
    const int lane_grp = warp & 3;  // a warp may only touch TMEM lanes [32*(warp%4), +32)
    const int t = t0 + lane_grp * 32 + lane;
    const bool valid = t < g.T_out;
    const uint32_t taddr = tmem_base + (static_cast<uint32_t>(lane_grp * 32) << 16);
    const int nbase = n_tile * NTILE;
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
  tc_fence_before();
  __syncthreads();
  __shared__ int stats_last;
  if (e.stats_out && (threadIdx.x & 31) == 0) gn_stats_publish();
  if (e.stats_out && !e.stats_nofinish) gn_stats_finish(e, g.B, gridDim.x * gridDim.y * gridDim.z, &stats_last);   // ordered GroupNorm sums by the last CTA
  if (warp == 1) tmem_dealloc<(NTILE < 32 ? 32 : NTILE)>(tmem_base);
}

template <int NTILE, int KC, int STAGES, typename T>
static cudaError_t launch_umma_inst(const ConvLaunch& L, cudaStream_t st) {
  using S = UmmaSmem<NTILE, KC, STAGES>;
  auto kern = conv_gemm_umma_kernel<NTILE, KC, STAGES, T>;
  static PerDeviceOnce configured;
  if (configured.needed()) {
    cudaError_t err = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, S::TOTAL);
    if (err != cudaSuccess) return err;
    configured.mark();
  }
  dim3 grid((L.g.T_out + 127) / 128, (L.g.N / NTILE) * L.g.n_phases, L.g.B);
  kern<<<grid, 192, S::TOTAL, st>>>(L.tmA[0], L.tmA[L.g.n_src > 1 ? 1 : 0], L.tmW, L.g, L.e);
  return cudaGetLastError();
}

template <typename T>
cudaError_t launch_conv_umma(const ConvLaunch& L, cudaStream_t st) {
  const int nt = L.ntile, kc = L.kc;
  if (kc == 64 && nt == 256) return launch_umma_inst<256, 64, 2, T>(L, st);
  if (kc == 64 && nt == 128) return launch_umma_inst<128, 64, 3, T>(L, st);
  if (kc == 64 && nt == 64) return launch_umma_inst<64, 64, 4, T>(L, st);
  if (kc == 64 && nt == 32) return launch_umma_inst<32, 64, 4, T>(L, st);
  if (kc == 32 && nt == 32) return launch_umma_inst<32, 32, 4, T>(L, st);
  return cudaErrorInvalidConfiguration;
}
template cudaError_t launch_conv_umma<__nv_bfloat16>(const ConvLaunch&, cudaStream_t);
template cudaError_t launch_conv_umma<__half>(const ConvLaunch&, cudaStream_t);

// =================================================================================================== SIMT
// Tile 128 rows x 64 channels, 256 threads; thread (row = tid & 127, half = tid >> 7) owns 32 channels of one
// row -- the same row-slice ownership as the TMEM epilogue, so both kernels share epilogue_chunk().
template <typename T>
__global__ void __launch_bounds__(256)
conv_gemm_simt_kernel(const T* __restrict__ a0, const T* __restrict__ a1, const T* __restrict__ w,
                      const __grid_constant__ ConvGeom g, const __grid_constant__ EpiParams e) {
  __shared__ float As[32][129];
  __shared__ __align__(16) float Ws[32][64];
  const int tid = threadIdx.x;
  const int n_tiles = (g.N + 63) / 64;
  const int n_tile = blockIdx.y % n_tiles, phase = blockIdx.y / n_tiles;
  const int b = blockIdx.z;
  const int t0 = blockIdx.x * 128;
  const int row = tid & 127, half = tid >> 7;
  float acc[32];
#pragma unroll
  for (int i = 0; i < 32; ++i) acc[i] = 0.f;

  const int a_row = tid >> 1, a_seg = (tid & 1) * 16;   // A tile loader: 16 channels of one row
  const int w_n = tid >> 2, w_seg = (tid & 3) * 8;      // W tile loader: 8 k-values of one output channel
  long kofs = 0;
  for (int tap = 0; tap < g.n_taps; ++tap) {
    for (int src = 0; src < g.n_src; ++src) {
      const T* a = src ? a1 : a0;
      const int srow = t0 + a_row + g.tap_row[phase][tap];
      const bool in = srow >= 0 && srow < g.T_in;
      const T* ap = a + (static_cast<long>(b) * g.T_in + srow) * g.src_ld[src] + g.tap_ch[phase][tap] + a_seg;
      const int wn = n_tile * 64 + w_n;
      const T* wp = w + (static_cast<long>(phase) * g.N + wn) * g.K_total + kofs + w_seg;
      for (int c = 0; c < g.src_ch[src]; c += 32) {
        float x[16], y[8];
        if (in) { Elem<T>::ld8(ap + c, x); Elem<T>::ld8(ap + c + 8, x + 8); }
        else {
#pragma unroll
          for (int i = 0; i < 16; ++i) x[i] = 0.f;
        }
        if (wn < g.N) Elem<T>::ld8(wp + c, y);
        else {
#pragma unroll
          for (int i = 0; i < 8; ++i) y[i] = 0.f;
        }
        __syncthreads();
#pragma unroll
        for (int i = 0; i < 16; ++i) As[a_seg + i][a_row] = x[i];
#pragma unroll
        for (int i = 0; i < 8; ++i) Ws[w_seg + i][w_n] = y[i];
        __syncthreads();
#pragma unroll 4
        for (int k = 0; k < 32; ++k) {
          const float av = As[k][row];
          const float4* wr = reinterpret_cast<const float4*>(&Ws[k][half * 32]);
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            const float4 wv = wr[j];
            acc[4 * j + 0] = fmaf(av, wv.x, acc[4 * j + 0]);
            acc[4 * j + 1] = fmaf(av, wv.y, acc[4 * j + 1]);
            acc[4 * j + 2] = fmaf(av, wv.z, acc[4 * j + 2]);
            acc[4 * j + 3] = fmaf(av, wv.w, acc[4 * j + 3]);
          }
        }
      }
      kofs += g.src_ch[src];
    }
  }
  const int t = t0 + row;
  const int n0 = n_tile * 64 + half * 32;
  if (n0 < g.N) epilogue_chunk<T>(e, b, t, phase, n0, t < g.T_out, true, acc);  // n0 is warp-uniform
  __shared__ int stats_last;
  if (e.stats_out && (threadIdx.x & 31) == 0) gn_stats_publish();
  if (e.stats_out && !e.stats_nofinish) gn_stats_finish(e, g.B, gridDim.x * gridDim.y * gridDim.z, &stats_last);   // ordered GroupNorm sums by the last CTA
}

template <typename T>
cudaError_t launch_conv_simt(const ConvLaunch& L, cudaStream_t st) {
  dim3 grid((L.g.T_out + 127) / 128, ((L.g.N + 63) / 64) * L.g.n_phases, L.g.B);
  conv_gemm_simt_kernel<T><<<grid, 256, 0, st>>>(reinterpret_cast<const T*>(L.a[0]), reinterpret_cast<const T*>(L.a[1]),
                                                 reinterpret_cast<const T*>(L.w), L.g, L.e);
  return cudaGetLastError();
}
template cudaError_t launch_conv_simt<float>(const ConvLaunch&, cudaStream_t);
template cudaError_t launch_conv_simt<__nv_bfloat16>(const ConvLaunch&, cudaStream_t);
template cudaError_t launch_conv_simt<__half>(const ConvLaunch&, cudaStream_t);

}  // namespace mb
