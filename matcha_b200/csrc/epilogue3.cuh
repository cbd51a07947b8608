// Epilogue of the persistent tcgen05 kernel.
//
// tcgen05.ld hands every thread one accumulator ROW (32 lanes = 32 rows x 32 fp32 columns).  Doing the fused
// element-wise work in that layout makes every global access touch 32 different lines and needs one scalar load per
// channel parameter.  Instead each warp transposes its 32x32 fp32 block once through a private shared-memory tile
// (pitch 144 B, conflict-free both ways) into a COALESCED layout:
//     lane = 4*rq + p   ->  rows rq, rq+8, rq+16, rq+24 ; channels 8p .. 8p+7 of the 32-column block
// so that every residual / GroupNorm input / output access is a 16-byte piece of a 64-byte row segment (four lanes per
// row, fully used sectors), channel parameters are two float4 loads per block, and the first input stream of the next
// block is requested before the current block is processed (and before the accumulator barrier is waited on).
//
// A tile's (128-row block, 32-column block) items are split between the two epilogue groups, so all eight warps drain
// every tile.  LayerNorm row sums are exchanged between the two warps that share a row through shared memory.
#pragma once
#include "conv.cuh"
#include "ptx.cuh"

namespace mb {

// Compile-time feature classes.  The epilogue is one function with every fusion the path needs; a launch only uses a
// few of them, and the tests on the others (parameter loads, uniform branches, live registers) cost ~40 % of the
// instructions of this latency-bound code.  F is a SUPERSET of what the launch's EpiParams may ask for: features outside
// F are compiled out, features inside F are still tested at run time.  kF_ALL is the fully generic epilogue.
constexpr uint32_t kF_GN = 1u, kF_RESID = 2u, kF_RESID2 = 4u, kF_SNAKE = 8u, kF_MASK = 16u, kF_STATS = 32u, kF_XSTATE = 64u,
                   kF_LN = 128u, kF_OUT1 = 256u, kF_SCALE = 512u, kF_F32 = 1024u, kF_SLOPE = 2048u;
constexpr uint32_t kF_NOFAST = 4096u;   // the packed-half path is known not to apply (estimator layers)
constexpr uint32_t kF_ALL = 0xfffu;
// classes the launchers instantiate
constexpr uint32_t kFC_STATS = kF_STATS | kF_NOFAST;                       // Block1D convs: bias + GroupNorm partial sums
constexpr uint32_t kFC_GNLN = kF_GN | kF_MASK | kF_LN | kF_NOFAST;         // res_conv + mish(GN(raw)) + LayerNorm
constexpr uint32_t kFC_RESLN = kF_RESID | kF_LN | kF_NOFAST;               // attention out-projection + residual + LayerNorm
constexpr uint32_t kFC_SNAKE = kF_SNAKE | kF_NOFAST;                       // feed-forward up projection + SnakeBeta
constexpr uint32_t kFC_RESMASK = kF_RESID | kF_MASK | kF_NOFAST;           // feed-forward down projection, masked convs
constexpr uint32_t kFC_VOC = kF_RESID | kF_RESID2 | kF_SCALE | kF_SLOPE | kF_OUT1;   // every HiFi-GAN conv (packed-half path in f16)
__host__ inline uint32_t epi_features(const EpiParams& e) {
  uint32_t f = 0;
  if (e.gn_raw) f |= kF_GN;
  if (e.resid) f |= kF_RESID;
  if (e.resid2) f |= kF_RESID2;
  if (e.snake_a) f |= kF_SNAKE;
  if (e.mask || e.mask_out) f |= kF_MASK;
  if (e.stats_out) f |= kF_STATS;
  if (e.x_state) f |= kF_XSTATE;
  if (e.ln_out) f |= kF_LN;
  if (e.out1) f |= kF_OUT1 | kF_SLOPE;
  if (e.scale != 1.f) f |= kF_SCALE;
  if (e.out0_f32 || e.resid_f32 || e.resid2_f32 || e.all_f32 || e.out0_lo || e.out1_lo) f |= kF_F32;
  if (e.out0_slope != 1.f) f |= kF_SLOPE;
  return f;
}

constexpr int kStagePitch = 144;                 // bytes per staged row: 32 fp32 + 16 B pad
constexpr int kStageBytes = 32 * kStagePitch;    // per epilogue warp
constexpr int kLnXBytes = 2 * 4 * 128 * 8;        // [tile parity][group <= 4][row] partial LayerNorm sums (float2)

// mish(x) = x tanh(softplus(x)) = x w / (w + 2),  w = e^x (e^x + 2).  Branch-free on purpose: written with an early
// `if (x > 20) return x`, every element became its own divergent region (BSSY / BRA / BSYNC) and the compiler could not
// interleave the 13-instruction dependent chains of neighbouring channels - 41 cycles per element in the fused epilogues
// (ncu source view).  Clamping the exponent argument keeps w finite and gives x (1 - 2e-17) = x for x >= 20; ex2 / rcp with
// .ftz skip the denormal fix-up sequences (w + 2 >= 2 is never denormal, e^x underflowing to 0 gives mish = 0 as it should).
__device__ __forceinline__ float mish_fast(float x) {
  float n, rc;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(n) : "f"(fminf(x, 20.f) * 1.4426950408889634f));
  const float w = n * (n + 2.f);
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(rc) : "f"(w + 2.f));
  return x * (w * rc);
}

template <typename T> __device__ __forceinline__ void unpack8(const uint4& u, float* f);
template <> __device__ __forceinline__ void unpack8<__half>(const uint4& u, float* f) {
  const __half2* h = reinterpret_cast<const __half2*>(&u);
#pragma unroll
  for (int i = 0; i < 4; ++i) { const float2 x = __half22float2(h[i]); f[2 * i] = x.x; f[2 * i + 1] = x.y; }
}
template <> __device__ __forceinline__ void unpack8<__nv_bfloat16>(const uint4& u, float* f) {
  const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&u);
#pragma unroll
  for (int i = 0; i < 4; ++i) { const float2 x = __bfloat1622float2(h[i]); f[2 * i] = x.x; f[2 * i + 1] = x.y; }
}
template <typename T> __device__ __forceinline__ uint4 pack8(const float* f);
template <> __device__ __forceinline__ uint4 pack8<__half>(const float* f) {
  uint4 u; __half2* h = reinterpret_cast<__half2*>(&u);
#pragma unroll
  for (int i = 0; i < 4; ++i) h[i] = __floats2half2_rn(f[2 * i], f[2 * i + 1]);
  return u;
}
template <> __device__ __forceinline__ uint4 pack8<__nv_bfloat16>(const float* f) {
  uint4 u; __nv_bfloat162* h = reinterpret_cast<__nv_bfloat162*>(&u);
#pragma unroll
  for (int i = 0; i < 4; ++i) h[i] = __floats2bfloat162_rn(f[2 * i], f[2 * i + 1]);
  return u;
}
__device__ __forceinline__ void ldg8f(const float* p, float* f) {
  const float4 a = __ldg(reinterpret_cast<const float4*>(p)), b = __ldg(reinterpret_cast<const float4*>(p) + 1);
  f[0] = a.x; f[1] = a.y; f[2] = a.z; f[3] = a.w; f[4] = b.x; f[5] = b.y; f[6] = b.z; f[7] = b.w;
}
// 8 consecutive elements of an activation row, stored as T or as fp32
template <typename T>
__device__ __forceinline__ void load8(const void* base, long elem, int f32, float* f) {
  if (f32) {
    const float4* p = reinterpret_cast<const float4*>(reinterpret_cast<const float*>(base) + elem);
    const float4 a = p[0], b = p[1];
    f[0] = a.x; f[1] = a.y; f[2] = a.z; f[3] = a.w; f[4] = b.x; f[5] = b.y; f[6] = b.z; f[7] = b.w;
  } else {
    unpack8<T>(*reinterpret_cast<const uint4*>(reinterpret_cast<const T*>(base) + elem), f);
  }
}
// split-precision store: f = hi + lo with hi, lo in bf16 (16 mantissa bits kept)
__device__ __forceinline__ void store8_split(void* hi_base, void* lo_base, long elem, const float* f) {
  __nv_bfloat162 h[4], l[4];
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    const __nv_bfloat16 h0 = __float2bfloat16_rn(f[2 * k]), h1 = __float2bfloat16_rn(f[2 * k + 1]);
    h[k] = __halves2bfloat162(h0, h1);
    l[k] = __floats2bfloat162_rn(f[2 * k] - __bfloat162float(h0), f[2 * k + 1] - __bfloat162float(h1));
  }
  *reinterpret_cast<uint4*>(reinterpret_cast<__nv_bfloat16*>(hi_base) + elem) = *reinterpret_cast<const uint4*>(h);
  *reinterpret_cast<uint4*>(reinterpret_cast<__nv_bfloat16*>(lo_base) + elem) = *reinterpret_cast<const uint4*>(l);
}
template <typename T>
__device__ __forceinline__ void store8(void* base, long elem, int f32, const float* f) {
  if (f32) {
    float4* p = reinterpret_cast<float4*>(reinterpret_cast<float*>(base) + elem);
    p[0] = make_float4(f[0], f[1], f[2], f[3]);
    p[1] = make_float4(f[4], f[5], f[6], f[7]);
  } else {
    *reinterpret_cast<uint4*>(reinterpret_cast<T*>(base) + elem) = pack8<T>(f);
  }
}
__device__ __forceinline__ void named_bar_sync(int id, int threads) {
  asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(threads) : "memory");
}

// Everything one epilogue warp does for one tile.  `tmem_tile` already holds the warp's lane offset and the accumulator
// slot; `lnx` is this tile's LayerNorm exchange buffer ([group][128 rows]).  The accumulator barrier is waited on here
// so that the first input stream can be requested ahead of it.
//
// The profile of this function is flat and latency-bound (two epilogue warps per scheduler), so the code is organised
// to keep the instruction count per 32x32 item low: options are tested once per item (feature-major, row-minor loops),
// row offsets are 32-bit element indices computed once, and memory operations are predicated rather than branched.
#ifdef MB200_PROBE   // probe_wait[0] = wait for the accumulator; [1..5] = item header, TMEM load, staging, arithmetic, stores
#define EP_MARK(k) do { if (probe_wait) { const long long ep_now = clock64(); probe_wait[k] += ep_now - ep_last; ep_last = ep_now; } } while (0)
#else
#define EP_MARK(k) do {} while (0)
#endif
template <int NTILE, int MB, int EG, typename T, uint32_t F = kF_ALL>
__device__ __forceinline__ void epilogue_tile3(const EpiParams& ep, const int b, const int t0, const int T_out,
                                               const int phase, const int nbase, const uint32_t tmem_tile,
                                               uint8_t* stg, float2* lnx, const int lane, const int lane_grp,
                                               const int grp, uint64_t* t_full, const uint32_t t_parity,
                                               long long* probe_wait = nullptr) {
  constexpr int NP = NTILE / 32;            // 32-column blocks per 128-row block
  constexpr int ITEMS = MB * NP;
  // masked copy of the parameters: everything outside the feature class F is a compile-time constant from here on
  EpiParams e = ep;
  if constexpr (!(F & kF_GN)) e.gn_raw = nullptr;
  if constexpr (!(F & kF_RESID)) e.resid = nullptr;
  if constexpr (!(F & kF_RESID2)) e.resid2 = nullptr;
  if constexpr (!(F & kF_SNAKE)) e.snake_a = nullptr;
  if constexpr (!(F & kF_MASK)) { e.mask = nullptr; e.mask_out = 0; }
  if constexpr (!(F & kF_STATS)) e.stats_out = nullptr;
  if constexpr (!(F & kF_XSTATE)) e.x_state = nullptr;
  if constexpr (!(F & kF_LN)) e.ln_out = nullptr;
  if constexpr (!(F & kF_OUT1)) e.out1 = nullptr;
  if constexpr (!(F & kF_SCALE)) e.scale = 1.f;
  if constexpr (!(F & kF_F32)) { e.out0_f32 = 0; e.resid_f32 = 0; e.resid2_f32 = 0; e.all_f32 = 0; e.out0_lo = nullptr; e.out1_lo = nullptr; }
  if (e.all_f32) { e.out0_f32 = 1; e.resid_f32 = 1; e.resid2_f32 = 1; }
  const int af = e.all_f32;
  if constexpr (!(F & kF_SLOPE)) e.out0_slope = 1.f;
  const int p = lane & 3, rq = lane >> 2;
  const int row_base = b * e.T_o + phase * e.o_phase_step;   // tensors on this path stay below 2^31 elements
  const int tw0 = t0 + lane_grp * 32 + rq;  // this lane's first row of block 0
  // stream A = the first half-precision input stream; it is requested one item ahead
  const T* inA = af ? nullptr : reinterpret_cast<const T*>(e.gn_raw ? e.gn_raw : (e.resid_f32 ? nullptr : e.resid));
  const int ldA = e.gn_raw ? e.gn_ld : e.resid_ld;
  const bool resid_is_A = (e.gn_raw == nullptr) && (inA != nullptr);
  const int n_valid = e.n_valid, o_mul = e.o_mul;

  auto issueA = [&](int item, uint4 (&u)[4]) {
    const int i = item / NP, c = item % NP;
    const int n0 = nbase + c * 32 + 8 * p;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const int t = tw0 + 128 * i + 8 * k;
      u[k] = make_uint4(0, 0, 0, 0);
      if (t < T_out && n0 < n_valid) u[k] = __ldg(reinterpret_cast<const uint4*>(inA + static_cast<size_t>(row_base + t * o_mul) * ldA + n0));
    }
  };

  uint4 nxt[4];
  int item = grp;
  if (inA && item < ITEMS) issueA(item, nxt);
  if (probe_wait) {
    const long long a = clock64();
    mbar_wait(t_full, t_parity);
    *probe_wait += clock64() - a;
  } else {
    mbar_wait(t_full, t_parity);
  }
  tc_fence_after();
#ifdef MB200_PROBE
  long long ep_last = clock64();
#endif

  const bool do_ln = (NTILE == 256) && (e.ln_out != nullptr);
  // fp16 layers whose epilogue is only bias / residuals / scale / LeakyReLU (every HiFi-GAN conv) finish in packed half
  // arithmetic: a third of the instructions of the general fp32 path; the extra roundings are fp16 ulps of values that
  // are stored as fp16 anyway.  bf16 keeps the fp32 path.
  const bool fast = !(F & kF_NOFAST) && std::is_same<T, __half>::value && !e.gn_raw && !e.snake_a && !e.mask && !e.mask_out && !e.stats_out &&
                    !e.x_state && !do_ln && !e.out0_f32 && !e.resid_f32 && !e.resid2_f32;
  float ln_s[4] = {0.f, 0.f, 0.f, 0.f}, ln_ss[4] = {0.f, 0.f, 0.f, 0.f};

#pragma unroll 1
  for (; item < ITEMS; item += EG) {
    const int i = item / NP, c = item % NP;
    uint4 cur[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) cur[k] = nxt[k];
    if (inA && item + EG < ITEMS) issueA(item + EG, nxt);
    if (t0 + 128 * i + lane_grp * 32 >= T_out) continue;   // warp-uniform: no row of this block exists

    const int n0 = nbase + c * 32 + 8 * p;
    const bool cv = n0 < n_valid;
    bool ok[4];     // row exists and this lane's channels exist: gates every load and store
    int grow[4];    // output row index of the lane's four rows
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const int t = tw0 + 128 * i + 8 * k;
      ok[k] = (t < T_out) && cv;
      grow[k] = row_base + t * o_mul;
    }
    // per-channel parameters: requested here so that they are in flight while the accumulators are fetched
    float bv[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    float pa[8], pb[8];   // GroupNorm gamma/beta or SnakeBeta a / 1/b (never both on one layer)
    if (e.bias) ldg8f(e.bias + n0, bv);
    float gsum = 0.f, gsq = 0.f;   // GroupNorm partial sums of this item's 32-channel group
    if (e.gn_raw) {
      ldg8f(e.gn_gamma + n0, pa); ldg8f(e.gn_beta + n0, pb);
      const float* st = e.gn_stats + (b * 8 + (n0 >> 5)) * 2;
      gsum = st[0]; gsq = st[1];
    }
    else if (e.snake_a) { ldg8f(e.snake_a + n0, pa); ldg8f(e.snake_invb + n0, pb); }

    // ---- accumulators: TMEM -> registers (row per lane) -> staging -> coalesced layout
    float v[4][8];
    EP_MARK(1);
    {
      float acc[32];
      tmem_ld_32x32(tmem_tile + i * NTILE + c * 32, acc);
      tmem_ld_wait();
      EP_MARK(2);
      float4* srow = reinterpret_cast<float4*>(stg + lane * kStagePitch);
#pragma unroll
      for (int j = 0; j < 8; ++j) srow[j] = make_float4(acc[4 * j], acc[4 * j + 1], acc[4 * j + 2], acc[4 * j + 3]);
      __syncwarp();
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const float4* q = reinterpret_cast<const float4*>(stg + (rq + 8 * k) * kStagePitch + p * 32);
        const float4 x = q[0], y = q[1];
        v[k][0] = x.x + bv[0]; v[k][1] = x.y + bv[1]; v[k][2] = x.z + bv[2]; v[k][3] = x.w + bv[3];
        v[k][4] = y.x + bv[4]; v[k][5] = y.y + bv[5]; v[k][6] = y.z + bv[6]; v[k][7] = y.w + bv[7];
      }
      __syncwarp();
    }
    EP_MARK(3);
    if (fast) {
      // ---- packed-half path: bias + residual(s) + scale + LeakyReLU copies, two channels per instruction
      if constexpr (std::is_same<T, __half>::value) {
        const bool inv_resid = e.resid && e.resid_inv != 1.f;
        if (inv_resid) {   // the stored stream is lrelu(x): undo it in fp32 (exact inverse of the stored slope) and add
          const __half* rp = reinterpret_cast<const __half*>(e.resid);
          const int ld = e.resid_ld;
          const float inv = e.resid_inv;
#pragma unroll
          for (int k = 0; k < 4; ++k) {
            uint4 ru = cur[k];
            if (!resid_is_A) { ru = make_uint4(0, 0, 0, 0); if (ok[k]) ru = *reinterpret_cast<const uint4*>(rp + static_cast<size_t>(grow[k]) * ld + n0); }
            float r[8];
            unpack8<__half>(ru, r);
#pragma unroll
            for (int ch = 0; ch < 8; ++ch) v[k][ch] += fminf(r[ch], r[ch] * inv);
          }
        }
        uint4 hu[4];
#pragma unroll
        for (int k = 0; k < 4; ++k) hu[k] = pack8<__half>(v[k]);
        if (e.resid && !inv_resid) {
          const __half* rp = reinterpret_cast<const __half*>(e.resid);
          const int ld = e.resid_ld;
#pragma unroll
          for (int k = 0; k < 4; ++k) {
            uint4 ru = cur[k];
            if (!resid_is_A) { ru = make_uint4(0, 0, 0, 0); if (ok[k]) ru = *reinterpret_cast<const uint4*>(rp + static_cast<size_t>(grow[k]) * ld + n0); }
            __half2* h = reinterpret_cast<__half2*>(&hu[k]);
            const __half2* r = reinterpret_cast<const __half2*>(&ru);
#pragma unroll
            for (int j = 0; j < 4; ++j) h[j] = __hadd2(h[j], r[j]);
          }
        }
        if (e.resid2) {
          const __half* rp = reinterpret_cast<const __half*>(e.resid2);
          const int ld = e.resid2_ld;
          uint4 ru[4];
#pragma unroll
          for (int k = 0; k < 4; ++k) { ru[k] = make_uint4(0, 0, 0, 0); if (ok[k]) ru[k] = *reinterpret_cast<const uint4*>(rp + static_cast<size_t>(grow[k]) * ld + n0); }
#pragma unroll
          for (int k = 0; k < 4; ++k) {
            __half2* h = reinterpret_cast<__half2*>(&hu[k]);
            const __half2* r = reinterpret_cast<const __half2*>(&ru[k]);
#pragma unroll
            for (int j = 0; j < 4; ++j) h[j] = __hadd2(h[j], r[j]);
          }
        }
        if (e.scale != 1.f) {
          const __half2 sc = __float2half2_rn(e.scale);
#pragma unroll
          for (int k = 0; k < 4; ++k) {
            __half2* h = reinterpret_cast<__half2*>(&hu[k]);
#pragma unroll
            for (int j = 0; j < 4; ++j) h[j] = __hmul2(h[j], sc);
          }
        }
        if (e.out0) {   // LeakyReLU(x) = max(x, slope * x) for 0 < slope <= 1
          __half* op = reinterpret_cast<__half*>(e.out0);
          const int ld = e.out0_ld;
          const __half2 sl = __float2half2_rn(e.out0_slope);
          const bool plain = e.out0_slope == 1.f;
#pragma unroll
          for (int k = 0; k < 4; ++k) {
            uint4 ou = hu[k];
            if (!plain) {
              __half2* o = reinterpret_cast<__half2*>(&ou);
              const __half2* h = reinterpret_cast<const __half2*>(&hu[k]);
#pragma unroll
              for (int j = 0; j < 4; ++j) o[j] = __hmax2(h[j], __hmul2(h[j], sl));
            }
            if (ok[k]) *reinterpret_cast<uint4*>(op + static_cast<size_t>(grow[k]) * ld + n0) = ou;
          }
        }
        if (e.out1) {
          __half* op = reinterpret_cast<__half*>(e.out1);
          const int ld = e.out1_ld;
          const __half2 sl = __float2half2_rn(e.out1_slope);
#pragma unroll
          for (int k = 0; k < 4; ++k) {
            uint4 ou;
            __half2* o = reinterpret_cast<__half2*>(&ou);
            const __half2* h = reinterpret_cast<const __half2*>(&hu[k]);
#pragma unroll
            for (int j = 0; j < 4; ++j) o[j] = __hmax2(h[j], __hmul2(h[j], sl));
            if (ok[k]) *reinterpret_cast<uint4*>(op + static_cast<size_t>(grow[k]) * ld + n0) = ou;
          }
        }
      }
      continue;
    }
    // ---- general fp32 path
    float m[4] = {1.f, 1.f, 1.f, 1.f};
    if (e.mask) {
#pragma unroll
      for (int k = 0; k < 4; ++k) if (ok[k]) m[k] = __ldg(e.mask + grow[k]);
    }
    if (e.gn_raw) {
      const float mean = gsum * e.gn_inv_count;
      const float rstd = rsqrtf(fmaxf(gsq * e.gn_inv_count - mean * mean, 0.f) + 1e-5f);
#pragma unroll
      for (int ch = 0; ch < 8; ++ch) { pa[ch] *= rstd; pb[ch] -= mean * pa[ch]; }
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        float r[8];
        if (af) {
#pragma unroll
          for (int ch = 0; ch < 8; ++ch) r[ch] = 0.f;
          if (ok[k]) load8<T>(e.gn_raw, static_cast<long>(grow[k]) * e.gn_ld + n0, 1, r);
        } else {
          unpack8<T>(cur[k], r);
        }
#pragma unroll
        for (int ch = 0; ch < 8; ++ch) v[k][ch] += mish_fast(fmaf(r[ch], pa[ch], pb[ch])) * m[k];
      }
    }
    if (e.resid) {
      const int ld = e.resid_ld, f32 = e.resid_f32;
      const float inv = e.resid_inv;
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        float r[8];
        if (resid_is_A) {
          unpack8<T>(cur[k], r);
        } else {
#pragma unroll
          for (int ch = 0; ch < 8; ++ch) r[ch] = 0.f;
          if (ok[k]) load8<T>(e.resid, static_cast<long>(grow[k]) * ld + n0, f32, r);
        }
        if (inv != 1.f) {
#pragma unroll
          for (int ch = 0; ch < 8; ++ch) r[ch] = fminf(r[ch], r[ch] * inv);
        }
#pragma unroll
        for (int ch = 0; ch < 8; ++ch) v[k][ch] += r[ch];
      }
    }
    if (e.resid2) {
      const int ld = e.resid2_ld, f32 = e.resid2_f32;
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        float r[8];
#pragma unroll
        for (int ch = 0; ch < 8; ++ch) r[ch] = 0.f;
        if (ok[k]) load8<T>(e.resid2, static_cast<long>(grow[k]) * ld + n0, f32, r);
#pragma unroll
        for (int ch = 0; ch < 8; ++ch) v[k][ch] += r[ch];
      }
    }
    if (e.scale != 1.f) {
      const float sc = e.scale;
#pragma unroll
      for (int k = 0; k < 4; ++k)
#pragma unroll
        for (int ch = 0; ch < 8; ++ch) v[k][ch] *= sc;
    }
    if (e.snake_a) {
#pragma unroll
      for (int k = 0; k < 4; ++k)
#pragma unroll
        for (int ch = 0; ch < 8; ++ch) {
          const float sn = __sinf(v[k][ch] * pa[ch]);
          v[k][ch] = fmaf(pb[ch] * sn, sn, v[k][ch]);
        }
    }
    if (e.mask_out) {
#pragma unroll
      for (int k = 0; k < 4; ++k)
#pragma unroll
        for (int ch = 0; ch < 8; ++ch) v[k][ch] *= m[k];
    }
    if (e.stats_out) {   // one GroupNorm group per 32-column block; warp-collective
      float sm = 0.f, ss = 0.f;
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const float w = ok[k] ? 1.f : 0.f;
#pragma unroll
        for (int ch = 0; ch < 8; ++ch) { sm = fmaf(w, v[k][ch], sm); ss = fmaf(w * v[k][ch], v[k][ch], ss); }
      }
      sm = warp_sum(sm);
      ss = warp_sum(ss);
      if (lane == 0) gn_stats_store(e, b, (nbase + c * 32) >> 5, (t0 + 128 * i + lane_grp * 32) >> 5, sm, ss);
    }
    if (e.x_state) {     // Euler update of the fp32 state; the stored copy below is x * mask
      const int ld = e.x_ld;
      const float dt = e.dt;
#pragma unroll
      for (int k = 0; k < 4; ++k)
        if (ok[k]) {
          float x[8];
          const long at = static_cast<long>(grow[k]) * ld + n0;
          load8<T>(e.x_state, at, 1, x);
#pragma unroll
          for (int ch = 0; ch < 8; ++ch) x[ch] = fmaf(dt, v[k][ch], x[ch]);
          store8<T>(e.x_state, at, 1, x);
#pragma unroll
          for (int ch = 0; ch < 8; ++ch) v[k][ch] = x[ch] * m[k];
        }
    }
    if (do_ln) {
#pragma unroll
      for (int k = 0; k < 4; ++k)
#pragma unroll
        for (int ch = 0; ch < 8; ++ch) { ln_s[k] += v[k][ch]; ln_ss[k] = fmaf(v[k][ch], v[k][ch], ln_ss[k]); }
    }
    EP_MARK(4);
    if (e.out0) {
      const int ld = e.out0_ld, f32 = e.out0_f32;
      const float sl = e.out0_slope;
      if (e.out0_lo) {
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          float o[8];
#pragma unroll
          for (int ch = 0; ch < 8; ++ch) o[ch] = fmaxf(v[k][ch], v[k][ch] * sl);   // sl == 1: the identity
          if (ok[k]) store8_split(e.out0, e.out0_lo, static_cast<long>(grow[k]) * ld + n0, o);
        }
      } else if (sl == 1.f) {
#pragma unroll
        for (int k = 0; k < 4; ++k)
          if (ok[k]) store8<T>(e.out0, static_cast<long>(grow[k]) * ld + n0, f32, v[k]);
      } else {
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          float o[8];
#pragma unroll
          for (int ch = 0; ch < 8; ++ch) o[ch] = fmaxf(v[k][ch], v[k][ch] * sl);
          if (ok[k]) store8<T>(e.out0, static_cast<long>(grow[k]) * ld + n0, f32, o);
        }
      }
    }
    if (e.out1) {
      const int ld = e.out1_ld;
      const float sl = e.out1_slope;
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        float o[8];
#pragma unroll
        for (int ch = 0; ch < 8; ++ch) o[ch] = fmaxf(v[k][ch], v[k][ch] * sl);
        if (ok[k]) {
          if (e.out1_lo) store8_split(e.out1, e.out1_lo, static_cast<long>(grow[k]) * ld + n0, o);
          else store8<T>(e.out1, static_cast<long>(grow[k]) * ld + n0, af, o);
        }
      }
    }
    EP_MARK(5);
  }

  if (do_ln) {
    // Row sums: four lanes share a row inside the warp, the other half of the columns belongs to the partner warp of
    // the other epilogue group.  The finished row is re-read from out0 (this thread's own stores, operand dtype).
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      ln_s[k] += __shfl_xor_sync(0xffffffffu, ln_s[k], 1);
      ln_ss[k] += __shfl_xor_sync(0xffffffffu, ln_ss[k], 1);
      ln_s[k] += __shfl_xor_sync(0xffffffffu, ln_s[k], 2);
      ln_ss[k] += __shfl_xor_sync(0xffffffffu, ln_ss[k], 2);
      if (p == 0) lnx[grp * 128 + lane_grp * 32 + rq + 8 * k] = make_float2(ln_s[k], ln_ss[k]);
    }
    named_bar_sync(1 + lane_grp, 32 * EG);
    float mean[4], rstd[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      float sm = 0.f, sq = 0.f;   // same summation order in every group: bit-identical statistics
#pragma unroll
      for (int gi = 0; gi < EG; ++gi) {
        const float2 o = lnx[gi * 128 + lane_grp * 32 + rq + 8 * k];
        sm += o.x; sq += o.y;
      }
      mean[k] = sm * (1.f / 256.f);
      rstd[k] = rsqrtf(fmaxf(sq * (1.f / 256.f) - mean[k] * mean[k], 0.f) + 1e-5f);
    }
#pragma unroll 1
    for (int c = grp; c < NP; c += EG) {
      const int n0 = c * 32 + 8 * p;
      float ga[8], be[8];
      ldg8f(e.ln_g + n0, ga);
      ldg8f(e.ln_b + n0, be);
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const int t = tw0 + 8 * k;
        if (t < T_out) {
          const long row = row_base + t * o_mul;
          float x[8];
          load8<T>(e.out0, row * e.out0_ld + n0, af, x);
#pragma unroll
          for (int ch = 0; ch < 8; ++ch) x[ch] = fmaf((x[ch] - mean[k]) * rstd[k], ga[ch], be[ch]);
          store8<T>(e.ln_out, row * e.ln_ld + n0, af, x);
        }
      }
    }
  }
}

}  // namespace mb
