// HBM-bound helper kernels of the hot path: GroupNorm-apply + Mish (+ time projection) + mask, LayerNorm (fp32 mode),
// boundary layout changes (B,C,T) <-> [B][T][C], estimator input packing, timestep embedding, vocoder head.
// All are one pass, 16-byte vector accesses along the contiguous channel axis.
#include <math.h>
#include <algorithm>

#include "conv.cuh"
#include "launch.h"

namespace mb {

// ---------------------------------------------------------------- GroupNorm apply + Mish (+temb) + mask
// out = (mish(gn(raw)) + tadd) * mask   (decoder.py:37-43 followed by :58 and the next block's x*mask, :41)
// fp32 mode keeps the accurate expf / division; the half modes round the result to 11 / 8 bits anyway and use the SFU forms
template <typename T> __device__ __forceinline__ float mish_sel(float x) { return mish_f(x); }
template <> __device__ __forceinline__ float mish_sel<__half>(float x) {
  if (x > 20.f) return x;
  const float n = __expf(x), w = n * (n + 2.f);
  return x * __fdividef(w, w + 2.f);
}
template <> __device__ __forceinline__ float mish_sel<__nv_bfloat16>(float x) { return mish_sel<__half>(x); }

// GroupNorm statistics arrive either finished ([B][8] (sum, sum of squares): part == nullptr) or as the producing conv's
// per-32-row partial sums ([B][8][slots] float2): the CTA then adds the slots of the (utterance, group) pairs its rows need
// in index order - a fixed order, so results stay bit-identical run to run - and the conv launch carries no ordered-sum
// tail (gn_stats_finish: +5.7 us per launch).  A CTA handles kGnRows rows (<= kGnRows utterances, <= 8 * kGnRows pairs).
constexpr int kGnRows = 8;    // one row per thread: more, smaller CTAs measured faster than 32 rows with four rows per thread (12.7 vs 16.9 us)
template <typename T>
__global__ void __launch_bounds__(256)
gn_mish_kernel(const T* __restrict__ raw, const float* __restrict__ stats, const float2* __restrict__ part, int slots,
               const float* __restrict__ gamma, const float* __restrict__ beta, const float* __restrict__ tadd, int tadd_bstride,
               const float* __restrict__ mask, T* __restrict__ out, int Tn, long total_rows, float inv_count) {
  __shared__ float2 st_s[8 * kGnRows];
  const long row_first = static_cast<long>(blockIdx.x) * kGnRows;
  const long row_last = min(row_first + kGnRows, total_rows) - 1;
  const int b_first = static_cast<int>(row_first / Tn), b_last = static_cast<int>(row_last / Tn);
  const int pairs = (b_last - b_first + 1) * 8;
  const int c0 = (threadIdx.x & 31) * 8;                    // this thread's 8 channels of a 256-channel row
  const float4 g0 = __ldg(reinterpret_cast<const float4*>(gamma + c0)), g1 = __ldg(reinterpret_cast<const float4*>(gamma + c0 + 4));
  const float4 b0 = __ldg(reinterpret_cast<const float4*>(beta + c0)), b1 = __ldg(reinterpret_cast<const float4*>(beta + c0 + 4));
  const float ga[8] = {g0.x, g0.y, g0.z, g0.w, g1.x, g1.y, g1.z, g1.w};
  const float be[8] = {b0.x, b0.y, b0.z, b0.w, b1.x, b1.y, b1.z, b1.w};
  // all four rows' loads are issued before any arithmetic and before the statistics are summed: one memory latency per CTA
  constexpr int NR = kGnRows / 8;
  long rows_[NR];
  float m_[NR];
  float r_[NR][8];
#pragma unroll
  for (int it = 0; it < NR; ++it) {
    rows_[it] = row_first + it * 8 + (threadIdx.x >> 5);
    const bool ok = rows_[it] < total_rows;
    m_[it] = ok ? __ldg(mask + rows_[it]) : 0.f;
#pragma unroll
    for (int k = 0; k < 8; ++k) r_[it][k] = 0.f;
    if (ok) Elem<T>::ld8(raw + rows_[it] * 256 + c0, r_[it]);
  }
  for (int j = threadIdx.x; j < pairs; j += 256) {
    const int bg = b_first * 8 + j;
    float2 acc = make_float2(0.f, 0.f);
    if (part) {
      const float2* pp = part + static_cast<long>(bg) * slots;
      for (int i = 0; i < slots; ++i) { const float2 v = __ldcg(pp + i); acc.x += v.x; acc.y += v.y; }
    } else {
      acc = __ldg(reinterpret_cast<const float2*>(stats) + bg);
    }
    st_s[j] = acc;
  }
  __syncthreads();
#pragma unroll
  for (int it = 0; it < NR; ++it) {
    const long row = rows_[it];
    if (row >= total_rows) continue;
    const int b = static_cast<int>(row / Tn);
    const float m = m_[it];
    float o[8];
    if (m == 0.f) {  // exact: anything times a zero mask is zero (inputs are finite)
#pragma unroll
      for (int k = 0; k < 8; ++k) o[k] = 0.f;
    } else {
      const float2 st = st_s[(b - b_first) * 8 + (c0 >> 5)];
      float4 t0 = make_float4(0.f, 0.f, 0.f, 0.f), t1 = t0;
      if (tadd) {
        const float* tp = tadd + static_cast<long>(b) * tadd_bstride + c0;
        t0 = __ldg(reinterpret_cast<const float4*>(tp));
        t1 = __ldg(reinterpret_cast<const float4*>(tp + 4));
      }
      const float mean = st.x * inv_count;
      const float rstd = rsqrtf(fmaxf(st.y * inv_count - mean * mean, 0.f) + 1e-5f);
      const float ta[8] = {t0.x, t0.y, t0.z, t0.w, t1.x, t1.y, t1.z, t1.w};
#pragma unroll
      for (int k = 0; k < 8; ++k) {
        const float y = (r_[it][k] - mean) * rstd * ga[k] + be[k];
        o[k] = (mish_sel<T>(y) + ta[k]) * m;
      }
    }
    Elem<T>::st8(out + row * 256 + c0, o);
  }
}

template <typename T>
cudaError_t launch_gn_mish(const T* raw, const float* stats, const float2* part, int slots, const float* gamma, const float* beta,
                           const float* tadd, int tadd_bstride, const float* mask, T* out, int B, int Tn, cudaStream_t st) {
  const long rows = static_cast<long>(B) * Tn;
  gn_mish_kernel<T><<<static_cast<unsigned>((rows + kGnRows - 1) / kGnRows), 256, 0, st>>>(
      raw, stats, part, slots, gamma, beta, tadd, tadd_bstride, mask, out, Tn, rows, 1.f / (32.f * Tn));
  return cudaGetLastError();
}
template cudaError_t launch_gn_mish<float>(const float*, const float*, const float2*, int, const float*, const float*, const float*, int,
                                           const float*, float*, int, int, cudaStream_t);
template cudaError_t launch_gn_mish<__nv_bfloat16>(const __nv_bfloat16*, const float*, const float2*, int, const float*, const float*,
                                                   const float*, int, const float*, __nv_bfloat16*, int, int, cudaStream_t);
template cudaError_t launch_gn_mish<__half>(const __half*, const float*, const float2*, int, const float*, const float*, const float*, int,
                                            const float*, __half*, int, int, cudaStream_t);

// ---------------------------------------------------------------- LayerNorm(256), one warp per row, two-pass variance
template <typename T>
__global__ void __launch_bounds__(256)
layernorm_kernel(const T* __restrict__ in, const float* __restrict__ g, const float* __restrict__ bta,
                 T* __restrict__ out, long rows) {
  const long row = static_cast<long>(blockIdx.x) * 8 + (threadIdx.x >> 5);
  if (row >= rows) return;
  const int lane = threadIdx.x & 31;
  float v[8];
  Elem<T>::ld8(in + row * 256 + lane * 8, v);
  float s = 0.f;
#pragma unroll
  for (int k = 0; k < 8; ++k) s += v[k];
  const float mean = warp_sum(s) * (1.f / 256.f);
  float ss = 0.f;
#pragma unroll
  for (int k = 0; k < 8; ++k) { const float d = v[k] - mean; ss += d * d; }
  const float rstd = rsqrtf(warp_sum(ss) * (1.f / 256.f) + 1e-5f);
  float o[8];
#pragma unroll
  for (int k = 0; k < 8; ++k) o[k] = (v[k] - mean) * rstd * __ldg(g + lane * 8 + k) + __ldg(bta + lane * 8 + k);
  Elem<T>::st8(out + row * 256 + lane * 8, o);
}
template <typename T>
cudaError_t launch_layernorm(const T* in, const float* g, const float* b, T* out, long rows, cudaStream_t st) {
  layernorm_kernel<T><<<static_cast<unsigned>((rows + 7) / 8), 256, 0, st>>>(in, g, b, out, rows);
  return cudaGetLastError();
}
template cudaError_t launch_layernorm<float>(const float*, const float*, const float*, float*, long, cudaStream_t);
template cudaError_t launch_layernorm<__nv_bfloat16>(const __nv_bfloat16*, const float*, const float*, __nv_bfloat16*,
                                                     long, cudaStream_t);
template cudaError_t launch_layernorm<__half>(const __half*, const float*, const float*, __half*, long, cudaStream_t);

// ---------------------------------------------------------------- (B,C,T) <-> [B][T][C] fp32, 32x32 shared-memory tiles
__global__ void __launch_bounds__(256) transpose_kernel(const float* __restrict__ in, float* __restrict__ out, int R, int C) {
  // in: [batch][R][C] -> out: [batch][C][R]
  __shared__ float tile[32][33];
  const int b = blockIdx.z;
  const int c0 = blockIdx.x * 32, r0 = blockIdx.y * 32;
  const float* src = in + static_cast<long>(b) * R * C;
  float* dst = out + static_cast<long>(b) * R * C;
  for (int i = threadIdx.y; i < 32; i += 8) {
    const int r = r0 + i, c = c0 + threadIdx.x;
    tile[i][threadIdx.x] = (r < R && c < C) ? src[static_cast<long>(r) * C + c] : 0.f;
  }
  __syncthreads();
  for (int i = threadIdx.y; i < 32; i += 8) {
    const int c = c0 + i, r = r0 + threadIdx.x;
    if (r < R && c < C) dst[static_cast<long>(c) * R + r] = tile[threadIdx.x][i];
  }
}
cudaError_t launch_bct_to_btc(const float* in, float* out, int B, int C, int Tn, cudaStream_t st) {
  dim3 grid((Tn + 31) / 32, (C + 31) / 32, B);
  transpose_kernel<<<grid, dim3(32, 8), 0, st>>>(in, out, C, Tn);
  return cudaGetLastError();
}
cudaError_t launch_btc_to_bct(const float* in, float* out, int B, int C, int Tn, cudaStream_t st) {
  dim3 grid((C + 31) / 32, (Tn + 31) / 32, B);
  transpose_kernel<<<grid, dim3(32, 8), 0, st>>>(in, out, Tn, C);
  return cudaGetLastError();
}

// ---------------------------------------------------------------- estimator input [x | mu | spk | 0] * mask  (decoder.py:384-388, :41)
template <typename T>
__global__ void __launch_bounds__(256)
pack_xin_kernel(const float* __restrict__ x, const float* __restrict__ mu, const float* __restrict__ spk, int spk_dim,
                const float* __restrict__ mask, T* __restrict__ xin, int ld, int Tn, long total_vec) {
  const long i = static_cast<long>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (i >= total_vec) return;
  const int vpr = ld / 8;
  const long row = i / vpr;
  const int c0 = static_cast<int>(i % vpr) * 8;
  const int b = static_cast<int>(row / Tn);
  const float m = __ldg(mask + row);
  float o[8];
#pragma unroll
  for (int k = 0; k < 8; ++k) {
    const int c = c0 + k;
    float v = 0.f;
    if (c < 80) v = x[row * 80 + c];
    else if (c < 160) v = mu[row * 80 + c - 80];
    else if (c < 160 + spk_dim) v = spk[static_cast<long>(b) * spk_dim + c - 160];
    o[k] = v * m;
  }
  Elem<T>::st8(xin + row * ld + c0, o);
}
template <typename T>
cudaError_t launch_pack_xin(const float* x_ntc, const float* mu_ntc, const float* spk, int spk_dim, const float* mask,
                            T* xin, int ld, int B, int Tn, cudaStream_t st) {
  const long total = static_cast<long>(B) * Tn * (ld / 8);
  pack_xin_kernel<T><<<static_cast<unsigned>((total + 255) / 256), 256, 0, st>>>(x_ntc, mu_ntc, spk, spk_dim, mask, xin,
                                                                                 ld, Tn, total);
  return cudaGetLastError();
}
template cudaError_t launch_pack_xin<float>(const float*, const float*, const float*, int, const float*, float*, int, int,
                                            int, cudaStream_t);
template cudaError_t launch_pack_xin<__nv_bfloat16>(const float*, const float*, const float*, int, const float*,
                                                    __nv_bfloat16*, int, int, int, cudaStream_t);
template cudaError_t launch_pack_xin<__half>(const float*, const float*, const float*, int, const float*, __half*, int,
                                             int, int, cudaStream_t);

// mask pyramid: mask[:, :, ::2] (decoder.py:407)
__global__ void mask_half_kernel(const float* __restrict__ mask, float* __restrict__ mh, int Tn, long total) {
  const long i = static_cast<long>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (i >= total) return;
  const int th = Tn / 2;
  const long b = i / th, t = i % th;
  mh[i] = mask[b * Tn + 2 * t];
}
cudaError_t launch_mask_half(const float* mask, float* mask_h, int B, int Tn, cudaStream_t st) {
  const long total = static_cast<long>(B) * (Tn / 2);
  mask_half_kernel<<<static_cast<unsigned>((total + 255) / 256), 256, 0, st>>>(mask, mask_h, Tn, total);
  return cudaGetLastError();
}

// ---------------------------------------------------------------- timestep embedding
// SinusoidalPosEmb (decoder.py:20-29): emb = cat(sin(e), cos(e)), e = (1000 t) * exp(j * -(ln 1e4 / (half-1))).
// exp/sin/cos are evaluated in double and rounded once, i.e. correctly rounded fp32 functions of fp32 arguments.
__global__ void sinusoidal_kernel(const float* __restrict__ t_vals, int c_in, int ld, float* __restrict__ emb) {
  const int half = c_in / 2;
  const int j = threadIdx.x;
  const float t = t_vals[blockIdx.x];
  if (j < half) {
    const float step = static_cast<float>(-(log(10000.0) / (half - 1)));
    const float w = static_cast<float>(exp(static_cast<double>(static_cast<float>(j) * step)));
    const float e = (1000.f * t) * w;
    emb[static_cast<long>(blockIdx.x) * ld + j] = static_cast<float>(sin(static_cast<double>(e)));
    emb[static_cast<long>(blockIdx.x) * ld + half + j] = static_cast<float>(cos(static_cast<double>(e)));
  }
}
cudaError_t launch_sinusoidal(const float* t_vals, int n_t, int c_in, int ld, float* emb, cudaStream_t st) {
  sinusoidal_kernel<<<n_t, 128, 0, st>>>(t_vals, c_in, ld, emb);
  return cudaGetLastError();
}

// y[t][n] = act(x[t] . w[n] + b[n]): one warp per output, fp32 (TimestepEmbedding decoder.py:105-117; mlp decoder.py:49)
__global__ void __launch_bounds__(256)
small_linear_kernel(const float* __restrict__ x, const float* __restrict__ w, const float* __restrict__ b,
                    float* __restrict__ y, int K, int N, int act) {
  const int n = blockIdx.x * 8 + (threadIdx.x >> 5);
  const int t = blockIdx.y;
  if (n >= N) return;
  const int lane = threadIdx.x & 31;
  const float* xr = x + static_cast<long>(t) * K;
  const float* wr = w + static_cast<long>(n) * K;
  float s = 0.f;
  for (int k = lane; k < K; k += 32) s = fmaf(xr[k], wr[k], s);
  s = warp_sum(s);
  if (lane == 0) {
    s += b[n];
    if (act == 1) s = s / (1.f + expf(-s));
    else if (act == 2) s = mish_f(s);
    y[static_cast<long>(t) * N + n] = s;
  }
}
cudaError_t launch_small_linear(const float* x, const float* w, const float* b, float* y, int n_t, int K, int N, int act,
                                cudaStream_t st) {
  dim3 grid((N + 7) / 8, n_t);
  small_linear_kernel<<<grid, 256, 0, st>>>(x, w, b, y, K, N, act);
  return cudaGetLastError();
}

// ---------------------------------------------------------------- vocoder boundary
template <typename T>
__global__ void __launch_bounds__(256)
pack_mel_kernel(const float* __restrict__ mel, int Tsrc, float scale, float shift, T* __restrict__ out, int ld, int L,
                long total) {
  const long i = static_cast<long>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (i >= total) return;
  const int vpr = ld / 8;
  const int t = static_cast<int>(i % L);           // t fastest: coalesced reads of the (B,C,T) source
  const int cg = static_cast<int>((i / L) % vpr);
  const long b = i / (static_cast<long>(L) * vpr);
  float o[8];
#pragma unroll
  for (int k = 0; k < 8; ++k) {
    const int c = cg * 8 + k;
    o[k] = c < 80 ? mel[(b * 80 + c) * Tsrc + t] * scale + shift : 0.f;
  }
  Elem<T>::st8(out + (b * L + t) * ld + cg * 8, o);
}
template <typename T>
cudaError_t launch_pack_mel(const float* mel_bct, int Tsrc, float scale, float shift, T* out, int ld, int B, int L,
                            cudaStream_t st) {
  const long total = static_cast<long>(B) * L * (ld / 8);
  pack_mel_kernel<T><<<static_cast<unsigned>((total + 255) / 256), 256, 0, st>>>(mel_bct, Tsrc, scale, shift, out, ld, L,
                                                                                 total);
  return cudaGetLastError();
}
template cudaError_t launch_pack_mel<float>(const float*, int, float, float, float*, int, int, int, cudaStream_t);
template cudaError_t launch_pack_mel<__nv_bfloat16>(const float*, int, float, float, __nv_bfloat16*, int, int, int,
                                                    cudaStream_t);
template cudaError_t launch_pack_mel<__half>(const float*, int, float, float, __half*, int, int, int, cudaStream_t);

__global__ void affine_crop_kernel(const float* __restrict__ in, int Tsrc, float scale, float shift,
                                   float* __restrict__ out, int L, long total) {
  const long i = static_cast<long>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (i >= total) return;
  const long bc = i / L;
  const int t = static_cast<int>(i % L);
  out[i] = in[bc * Tsrc + t] * scale + shift;
}
cudaError_t launch_affine_crop(const float* in, int Tsrc, float scale, float shift, float* out, int B, int C, int L,
                               cudaStream_t st) {
  const long total = static_cast<long>(B) * C * L;
  affine_crop_kernel<<<static_cast<unsigned>((total + 255) / 256), 256, 0, st>>>(in, Tsrc, scale, shift, out, L, total);
  return cudaGetLastError();
}

// conv_post Conv1d(32,1,7,pad 3) + tanh (hifigan/models.py:177,194-195); the input already carries leaky_relu(0.01).
// conv_post (32 -> 1, k = 7, pad 3) + tanh (hifigan/models.py:177,193-195; cli.py:118's clamp is a no-op after tanh).
// One block = 256 consecutive samples of one utterance: the 262 input rows are staged once in shared memory with
// coalesced 16-byte loads (row pitch padded to 20 words: conflict-free 16-byte reads), then every thread does its 224 FMAs.
template <typename T>
__global__ void __launch_bounds__(256)
conv_post_kernel(const T* __restrict__ in, const float* __restrict__ w, float bias, float* __restrict__ wav, long L) {
  constexpr int kRows = 256 + 6;
  constexpr int kPitch = 32 * sizeof(T) + 16;          // bytes
  __shared__ float ws[7 * 32];
  __shared__ __align__(16) unsigned char tile[kRows * kPitch];
  for (int i = threadIdx.x; i < 7 * 32; i += blockDim.x) ws[i] = w[i];
  const long t0 = static_cast<long>(blockIdx.x) * 256;
  const long b = blockIdx.y;
  constexpr int kVec = 32 * sizeof(T) / 16;            // 16-byte pieces per row
  for (int i = threadIdx.x; i < kRows * kVec; i += blockDim.x) {
    const int r = i / kVec, q = i % kVec;
    const long srow = t0 + r - 3;
    uint4 v = make_uint4(0, 0, 0, 0);                  // zero padding outside [0, L)
    if (srow >= 0 && srow < L) v = __ldg(reinterpret_cast<const uint4*>(in + (b * L + srow) * 32) + q);
    *reinterpret_cast<uint4*>(tile + r * kPitch + q * 16) = v;
  }
  __syncthreads();
  const long t = t0 + threadIdx.x;
  if (t >= L) return;
  float acc = bias;
#pragma unroll
  for (int j = 0; j < 7; ++j) {
    const T* p = reinterpret_cast<const T*>(tile + (threadIdx.x + j) * kPitch);
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      float v[8];
      Elem<T>::ld8(p + 8 * q, v);
#pragma unroll
      for (int k = 0; k < 8; ++k) acc = fmaf(v[k], ws[j * 32 + 8 * q + k], acc);
    }
  }
  wav[b * L + t] = tanhf(acc);
}
template <typename T>
cudaError_t launch_conv_post(const T* in, const float* w, float bias, float* wav, int B, long L, cudaStream_t st) {
  dim3 grid(static_cast<unsigned>((L + 255) / 256), B);
  conv_post_kernel<T><<<grid, 256, 0, st>>>(in, w, bias, wav, L);
  return cudaGetLastError();
}
template cudaError_t launch_conv_post<float>(const float*, const float*, float, float*, int, long, cudaStream_t);
template cudaError_t launch_conv_post<__nv_bfloat16>(const __nv_bfloat16*, const float*, float, float*, int, long,
                                                     cudaStream_t);
template cudaError_t launch_conv_post<__half>(const __half*, const float*, float, float*, int, long, cudaStream_t);

// ---------------------------------------------------------------------------------------------------------------------
// Split-precision fp32 mode: x (fp32) -> hi = bf16(x), lo = bf16(x - hi), so that x = hi + lo to 16 mantissa bits and
// a . w ~= a_hi w_hi + a_hi w_lo + a_lo w_hi runs as three bf16 tcgen05 products with fp32 accumulation (conv_op).
// Up to two sources ([rows][ld0], [rows][ld1], the first c0 / c1 columns of each) are written side by side into rows of
// c0 + c1 columns: the concat of a skip connection costs nothing extra.
__global__ void split_bf16x2_kernel(const float* __restrict__ s0, int ld0, int c0, const float* __restrict__ s1, int ld1,
                                    int c1, long rows, __nv_bfloat16* __restrict__ hi, __nv_bfloat16* __restrict__ lo) {
  const int per_row = (c0 + c1) >> 3;   // 8-column units per output row
  const long total = rows * per_row;
  for (long i = blockIdx.x * static_cast<long>(blockDim.x) + threadIdx.x; i < total; i += static_cast<long>(gridDim.x) * blockDim.x) {
    const long r = i / per_row;
    const int c = static_cast<int>(i - r * per_row) << 3;
    const float* src = c < c0 ? s0 + r * ld0 + c : s1 + r * ld1 + (c - c0);
    const float4 a = __ldg(reinterpret_cast<const float4*>(src)), b = __ldg(reinterpret_cast<const float4*>(src) + 1);
    const float x[8] = {a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w};
    __nv_bfloat162 h[4], l[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const __nv_bfloat16 h0 = __float2bfloat16_rn(x[2 * k]), h1 = __float2bfloat16_rn(x[2 * k + 1]);
      h[k] = __halves2bfloat162(h0, h1);
      l[k] = __floats2bfloat162_rn(x[2 * k] - __bfloat162float(h0), x[2 * k + 1] - __bfloat162float(h1));
    }
    const long o = r * (c0 + c1) + c;
    *reinterpret_cast<uint4*>(hi + o) = *reinterpret_cast<const uint4*>(h);
    *reinterpret_cast<uint4*>(lo + o) = *reinterpret_cast<const uint4*>(l);
  }
}

cudaError_t launch_split_bf16x2(const float* s0, int ld0, int c0, const float* s1, int ld1, int c1, long rows,
                                void* hi, void* lo, cudaStream_t st) {
  if ((c0 | c1 | ld0 | ld1) & 7) return cudaErrorInvalidValue;
  const long total = rows * ((c0 + c1) >> 3);
  const int grid = static_cast<int>(std::min<long>((total + 255) / 256, 148L * 16));
  split_bf16x2_kernel<<<std::max(grid, 1), 256, 0, st>>>(s0, ld0, c0, s1, ld1, c1, rows, static_cast<__nv_bfloat16*>(hi),
                                                         static_cast<__nv_bfloat16*>(lo));
  return cudaGetLastError();
}

}  // namespace mb
