// HiFi-GAN bias denoiser (matcha/hifigan/denoiser.py:10-64; run right after the vocoder by cli.py:117-120) as one kernel:
//
//     S = STFT(audio; n_fft 1024, hop 256, periodic Hann, center/reflect)        denoiser.py:26-44
//     S' = S * max(|S| - strength * bias, 0) / |S|                                denoiser.py:61-63 (magnitude/phase form)
//     out = iSTFT(S')  (window, overlap-add, / sum of squared windows, trim 512)  denoiser.py:45-51
//
// The reference goes through torch.stft -> sqrt/atan2 -> cos/sin -> torch.istft with (B, 513, frames) complex tensors in
// HBM (5 full passes).  Here a CTA owns 13 output hops of one utterance: it transforms the 16 frames that touch them in
// shared memory (radix-2 complex FFT, 1024 points, fp32), scales each bin by the real gain (cos(atan2(y,x)) = x/|S|, so
// no transcendental is needed), transforms back, overlap-adds in shared memory and writes each output sample once.
// HBM traffic = read audio (x 16/13) + write out.
#include <cmath>

#include "plan.h"

namespace mb {

namespace {
constexpr int kN = 1024, kHop = 256, kBins = 513;
constexpr int kHopsPerCta = 13, kFramesPerCta = kHopsPerCta + 3;

__device__ __forceinline__ int reflect_index(int i, int L) {   // torch.stft center=True, pad_mode="reflect"
  if (i < 0) i = -i;
  if (i >= L) i = 2 * (L - 1) - i;
  return i;
}

// In-place radix-2 decimation-in-time FFT of 1024 complex points in shared memory; input in bit-reversed order.
// sign = -1 forward, +1 inverse (unnormalised).  256 threads, two butterflies per thread per stage.
__device__ __forceinline__ void fft1024(float2* x, const float2* tw, float sign) {
#pragma unroll 1
  for (int s = 1; s <= 10; ++s) {
    const int half = 1 << (s - 1);
    __syncthreads();
#pragma unroll
    for (int q = 0; q < 2; ++q) {
      const int i = threadIdx.x + 256 * q;
      const int pos = i & (half - 1), a = ((i >> (s - 1)) << s) + pos, b = a + half;
      const float2 w = tw[pos << (10 - s)];          // (cos, sin)(2 pi pos / 2^s)
      const float2 xb = x[b], xa = x[a];
      const float cr = w.x, ci = sign * w.y;
      const float mr = xb.x * cr - xb.y * ci, mi = xb.x * ci + xb.y * cr;
      x[a] = make_float2(xa.x + mr, xa.y + mi);
      x[b] = make_float2(xa.x - mr, xa.y - mi);
    }
  }
  __syncthreads();
}

__device__ __forceinline__ int bitrev10(int i) { return static_cast<int>(__brev(static_cast<unsigned>(i)) >> 22); }

__device__ __forceinline__ void make_tables(float2* tw, float* win) {
  for (int k = threadIdx.x; k < 512; k += 256) {
    float s, c;
    sincospif(2.f * k / 1024.f, &s, &c);      // exact argument reduction: (cos, sin)(2 pi k / 1024)
    tw[k] = make_float2(c, s);
  }
  for (int n = threadIdx.x; n < kN; n += 256) win[n] = 0.5f - 0.5f * cospif(2.f * n / 1024.f);   // torch.hann_window (periodic)
}

// grid (chunks, B).  audio (B, L), out (B, Lout), Lout = 256 * (L / 256); bias [513]
__global__ void __launch_bounds__(256) denoise_kernel(const float* __restrict__ audio, int L, int Lout, int n_frames,
                                                      const float* __restrict__ bias, float strength,
                                                      float* __restrict__ out) {
  __shared__ float2 x[kN];
  __shared__ float2 tw[512];
  __shared__ float win[kN];
  __shared__ float ola[kFramesPerCta * kHop + 3 * kHop];
  const int b = blockIdx.y;
  const int hop0 = blockIdx.x * kHopsPerCta;                 // first output hop of this CTA
  const int n_hops = min(kHopsPerCta, Lout / kHop - hop0);
  if (n_hops <= 0) return;
  const float* a = audio + static_cast<long>(b) * L;
  make_tables(tw, win);
  for (int i = threadIdx.x; i < kFramesPerCta * kHop + 3 * kHop; i += 256) ola[i] = 0.f;
  // padded coordinate p = n + 512; the CTA's outputs are p in [P0, P0 + 256 n_hops); frames t = P0/256 - 3 .. P0/256 + n_hops - 1
  const int t_first = hop0 + 2 - 3;                          // P0 / 256 = hop0 + 2
  const int t_last = min(hop0 + 2 + n_hops - 1, n_frames - 1);
  __syncthreads();
  for (int t = max(t_first, 0); t <= t_last; ++t) {
    // frame t covers padded positions [256 t, 256 t + 1024) = audio indices [256 t - 512, ...)
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const int n = threadIdx.x + 256 * q;
      const float v = a[reflect_index(kHop * t - 512 + n, L)] * win[n];
      x[bitrev10(n)] = make_float2(v, 0.f);
    }
    fft1024(x, tw, -1.f);
    // spectral subtraction on the magnitude, phase kept; bins above Nyquist mirror the one-sided gain
    float2 y[4];
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const int k = threadIdx.x + 256 * q;
      const int kb = k <= 512 ? k : kN - k;
      float2 s = x[k];
      if (k == 0 || k == 512) s.y = 0.f;                     // irfft ignores the imaginary part of DC / Nyquist
      const float mag = sqrtf(s.x * s.x + s.y * s.y);
      const float g = mag > 0.f ? fmaxf(mag - bias[kb] * strength, 0.f) / mag : 0.f;
      y[q] = make_float2(s.x * g, s.y * g);
    }
    __syncthreads();
#pragma unroll
    for (int q = 0; q < 4; ++q) x[bitrev10(threadIdx.x + 256 * q)] = y[q];
    fft1024(x, tw, 1.f);
    const int base = (t - t_first) * kHop;                   // offset of this frame in the CTA's overlap-add buffer
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const int n = threadIdx.x + 256 * q;
      ola[base + n] += x[n].x * (1.f / kN) * win[n];         // four frames touch a sample, one at a time: no race
    }
    __syncthreads();
  }
  // outputs: padded positions P0 + i, i in [0, 256 n_hops): ola index = i + 3 * 256
  for (int i = threadIdx.x; i < n_hops * kHop; i += 256) {
    const int p = (hop0 + 2) * kHop + i;
    float env = 0.f;
#pragma unroll
    for (int d = 0; d < 4; ++d) {
      const int t = p / kHop - d;
      if (t >= 0 && t < n_frames) { const float w = win[p - t * kHop]; env += w * w; }
    }
    out[static_cast<long>(b) * Lout + hop0 * kHop + i] = ola[i + 3 * kHop] / env;
  }
}

// |STFT(audio)|[:, frame 0] (denoiser.py:55-57): one CTA
__global__ void __launch_bounds__(256) stft_frame0_mag_kernel(const float* __restrict__ audio, int L, float* __restrict__ mag) {
  __shared__ float2 x[kN];
  __shared__ float2 tw[512];
  __shared__ float win[kN];
  make_tables(tw, win);
  __syncthreads();
#pragma unroll
  for (int q = 0; q < 4; ++q) {
    const int n = threadIdx.x + 256 * q;
    x[bitrev10(n)] = make_float2(audio[reflect_index(n - 512, L)] * win[n], 0.f);
  }
  fft1024(x, tw, -1.f);
  for (int k = threadIdx.x; k < kBins; k += 256) mag[k] = sqrtf(x[k].x * x[k].x + x[k].y * x[k].y);
}
}  // namespace

void Plan::denoiser_bias(const float* audio, int L, float* bias, cudaStream_t st) {
  if (L <= 512) throw Error("denoiser: audio must be longer than n_fft / 2 = 512 samples (reflect padding)");
  stft_frame0_mag_kernel<<<1, 256, 0, st>>>(audio, L, bias);
  MB_CUDA(cudaGetLastError());
  launches += 1;
}

void Plan::denoise(const float* audio, int B, int L, const float* bias, float strength, float* out, cudaStream_t st) {
  if (L <= 512) throw Error("denoiser: audio must be longer than n_fft / 2 = 512 samples (reflect padding)");
  if (B < 1 || B > 65535) throw Error("denoiser: batch out of range");
  const int n_frames = L / kHop + 1, Lout = kHop * (n_frames - 1);
  const int chunks = (n_frames - 1 + kHopsPerCta - 1) / kHopsPerCta;
  denoise_kernel<<<dim3(chunks, B), 256, 0, st>>>(audio, L, Lout, n_frames, bias, strength, out);
  MB_CUDA(cudaGetLastError());
  launches += 1;
}

}  // namespace mb
