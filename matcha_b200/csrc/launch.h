// Host-visible launch interface of the kernels (internal; the public boundary is include/matcha_b200.h).
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>

#include <atomic>

#include "conv.cuh"

namespace mb {

// cudaFuncSetAttribute(MaxDynamicSharedMemorySize) is a per-DEVICE setting: every kernel instantiation remembers the
// devices it has been configured on (a plan may live on any GPU of the process).  Setting it twice is harmless, so two
// threads racing through the first launch on a device need no lock.
struct PerDeviceOnce {
  std::atomic<unsigned long long> done{0};
  int dev = 0;
  bool needed() {
    if (cudaGetDevice(&dev) != cudaSuccess) dev = 0;
    return !((done.load(std::memory_order_acquire) >> (dev & 63)) & 1ull);
  }
  void mark() { done.fetch_or(1ull << (dev & 63), std::memory_order_release); }
};

struct ConvLaunch {
  ConvGeom g;
  EpiParams e;
  const void* a[2];   // sources (SIMT path)
  const void* w;      // packed weights [n_phases*N][K_total]
  CUtensorMap tmA[2]; // TMA views (tcgen05 path)
  CUtensorMap tmW;
  int ntile, kc;      // tcgen05 tile: N per CTA, K chunk (channels per TMA box)
  // persistent kernel (conv_gemm2.cu)
  ConvGeom2 g2;
  CUtensorMap tmA8[2];  // same views, 8-row boxes (halo rows beyond the first 128)
  int smem_bytes, grid;
  int pdl;              // launch with programmatic stream serialization (prologue overlaps the previous kernel's tail)
};

struct PairLaunch {
  CUtensorMap tmA, tmA8, tmW1, tmW2;
  PairGeom g;
  EpiParams e;        // c2's epilogue (bias2, residual, branch sum, scale, output)
  int C, mb, smem_bytes, grid;
  int cta_pair;       // CTA pairs (cta_group::2): weight tiles split between the two CTAs of a cluster
};

struct FFLaunch {
  CUtensorMap tmX, tmW1, tmW2, tmWo;   // tmX: LayerNorm output [rows][256], or the attention output [rows][128] with the pre-stage
  FFGeom g;
  int smem_bytes, grid;
  int pair;                            // CTA pairs (cta_group::2): weight boxes split between the two CTAs of a cluster
};
struct RQLaunch {
  CUtensorMap tmA0, tmA1, tmWres, tmWqkv;
  RQGeom g;
  int smem_bytes, grid;
  int pair;                            // CTA pairs (cta_group::2)
};
int resqkv_fused_smem_bytes();
template <typename T> cudaError_t launch_resqkv_fused(const RQLaunch& L, cudaStream_t st);
int ff_fused_smem_bytes();
template <typename T> cudaError_t launch_ff_fused(const FFLaunch& L, cudaStream_t st);
template <typename T> cudaError_t launch_resblock_pair(const PairLaunch& L, cudaStream_t st);
template <typename T> cudaError_t launch_conv_umma(const ConvLaunch& L, cudaStream_t st);
template <typename T> cudaError_t launch_conv_simt(const ConvLaunch& L, cudaStream_t st);
template <typename T> cudaError_t launch_conv_umma2(const ConvLaunch& L, cudaStream_t st);

// ---- elementwise / small kernels (elementwise.cu)
template <typename T>
cudaError_t launch_gn_mish(const T* raw, const float* stats, const float2* part, int slots, const float* gamma, const float* beta,
                           const float* tadd, int tadd_bstride, const float* mask, T* out, int B, int Tn, cudaStream_t st);
template <typename T>
cudaError_t launch_layernorm(const T* in, const float* g, const float* b, T* out, long rows, cudaStream_t st);
cudaError_t launch_split_bf16x2(const float* s0, int ld0, int c0, const float* s1, int ld1, int c1, long rows, void* hi, void* lo,
                                cudaStream_t st);   // fp32 -> bf16 hi / lo (split-precision fp32 mode), optional concat
cudaError_t launch_bct_to_btc(const float* in, float* out, int B, int C, int Tn, cudaStream_t st);
cudaError_t launch_btc_to_bct(const float* in, float* out, int B, int C, int Tn, cudaStream_t st);
template <typename T>
cudaError_t launch_pack_xin(const float* x_ntc, const float* mu_ntc, const float* spk, int spk_dim, const float* mask,
                            T* xin, int ld, int B, int Tn, cudaStream_t st);
cudaError_t launch_mask_half(const float* mask, float* mask_h, int B, int Tn, cudaStream_t st);
// sinusoidal embedding + time MLP + the six per-ResNet projections (decoder.py:20-29, 105-117, 49)
cudaError_t launch_sinusoidal(const float* t_vals, int n_t, int c_in, int ld, float* emb, cudaStream_t st);
// small dense layer y[t][n] = act(sum_k x[t][k] w[n][k] + b[n]); act: 0 none, 1 SiLU, 2 Mish
cudaError_t launch_small_linear(const float* x, const float* w, const float* b, float* y, int n_t, int K, int N, int act,
                                cudaStream_t st);
// mel (B,80,Tsrc) fp32, first L frames -> operand [B][L][ld] = value*scale+shift, channels >= 80 zero
template <typename T>
cudaError_t launch_pack_mel(const float* mel_bct, int Tsrc, float scale, float shift, T* out, int ld, int B, int L,
                            cudaStream_t st);
// out (B,C,L) = in (B,C,Tsrc)[:, :, :L] * scale + shift   (crop + denormalize: matcha_tts.py:136,145)
cudaError_t launch_affine_crop(const float* in, int Tsrc, float scale, float shift, float* out, int B, int C, int L,
                               cudaStream_t st);
template <typename T>
cudaError_t launch_conv_post(const T* in, const float* w /*[7][32]*/, float bias, float* wav, int B, long L, cudaStream_t st);

// ---- attention (attention.cu): qkv [B][L][384] (q | k | v, 2 heads x 64), additive key mask, out [B][L][128]
template <typename T>
cudaError_t launch_attention_simt(const T* qkv, const float* mask, T* out, int B, int L, cudaStream_t st);
template <typename T>   // mma.sync tensor-core version (attention_mma.cu), half types only; kept as the A/B cross-check (MB200_ATTN=mma)
cudaError_t launch_attention_mma(const T* qkv, const float* mask, T* out, int B, int L, cudaStream_t st);
// tcgen05 / TMEM / TMA version (attention_umma.cu), half types only: the production kernel
struct AttnLaunch {
  CUtensorMap tmQKV;    // 3-D view (384 channels, L rows, B) of the qkv buffer, box 64 x 128 x 1, 128B swizzle
  const float* mask;    // [B][L] float 0/1, added to the logits
  void* out;            // [B][L][128] operand dtype
  int B, L, grid;
  long long* dbg;       // developer probe (MB200_ATTN_DBG): per-phase cycle counters of CTA 0, else null
};
template <typename T> cudaError_t launch_attention_umma(const AttnLaunch& A, cudaStream_t st);

}  // namespace mb
