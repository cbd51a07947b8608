// Length regulation: per-token frame counts -> monotonic alignment and the frame-rate encoder output, as three small kernels
// (SURVEY 8(f) row 2).  Replaces, on the synthesis path (matcha/models/matcha_tts.py:118-131):
//
//     attn = generate_path(w_ceil, x_mask (x) y_mask)          matcha/utils/model.py:29-41
//     mu_y = (attn^T @ mu_x^T)^T                               a (B, Lx, T) x (B, Lx, C) dense matmul over a 0/1 matrix
//
// attn[b, i, j] = [cum[b, i-1] <= j < cum[b, i]] * x_mask[b, i] * [j < y_len[b]]  with cum = cumsum(w_ceil) in fp32, so
// every frame has at most one owning token and mu_y[b, :, j] is a copy of that token's column (or zero): a gather, no
// FLOPs.  The prefix sums are taken SEQUENTIALLY with a double accumulator, each output rounded to fp32 - what torch.cumsum
// does on the CPU (ATen accumulates float in double there; verified against torch 2.11), which generated the golden
// vectors.  Durations scaled by a non-dyadic length_scale are not exact in fp32, and the comparisons `j < cum` must see
// the same rounded sums to reproduce the reference alignment bit for bit.
#include <cuda_runtime.h>
#include <stdint.h>

#include "plan.h"

namespace mb {

namespace {

// one warp per utterance: coalesced loads of 32 durations, then every lane performs the SAME sequential additions (the
// value of lane k arrives by shuffle) and lane k keeps the k-th partial sum - the order of a scalar loop, without its
// chain of dependent global loads
__global__ void __launch_bounds__(32) lr_cumsum_kernel(const float* __restrict__ w, int Lx, float* __restrict__ cum) {
  const int b = blockIdx.x, lane = threadIdx.x;
  const float* wr = w + static_cast<long>(b) * Lx;
  float* cr = cum + static_cast<long>(b) * Lx;
  double s = 0.0;
  for (int i0 = 0; i0 < Lx; i0 += 32) {
    const float x = (i0 + lane < Lx) ? wr[i0 + lane] : 0.f;
    float mine = 0.f;
#pragma unroll
    for (int k = 0; k < 32; ++k) {
      s += static_cast<double>(__shfl_sync(0xffffffffu, x, k));
      if (lane == k) mine = static_cast<float>(s);
    }
    if (i0 + lane < Lx) cr[i0 + lane] = mine;
  }
}

// grid (frame blocks of 128, channel groups of 16, B); 128 threads, one frame each.  The owner of frame j is the first token with cum > j
// (binary search over the utterance's prefix sums in shared memory); tokens that own no frame (w = 0) are skipped by the
// search exactly as they get an all-zero row from the reference's difference of prefix masks.
__global__ void __launch_bounds__(128)
lr_fill_kernel(const float* __restrict__ mu_x, const float* __restrict__ cum, const float* __restrict__ x_mask,
               const long long* __restrict__ y_len, int C, int Lx, int T, float* __restrict__ mu_y,
               float* __restrict__ y_mask) {
  extern __shared__ float s_cum[];   // [Lx] prefix sums, then [Lx] token mask
  float* s_xm = s_cum + Lx;
  const int b = blockIdx.z, j0 = blockIdx.x * 128, j = j0 + threadIdx.x;
  const int c0 = blockIdx.y * 16, c1 = min(C, c0 + 16);
  for (int i = threadIdx.x; i < Lx; i += 128) {
    s_cum[i] = cum[static_cast<long>(b) * Lx + i];
    s_xm[i] = x_mask[static_cast<long>(b) * Lx + i];
  }
  __syncthreads();
  const long long yl = y_len[b];
  const float fj = static_cast<float>(j);
  int tok = -1;
  float own = 0.f;
  if (j < T) {
    int lo = 0, hi = Lx;              // first i with cum[i] > j
    while (lo < hi) {
      const int mid = (lo + hi) >> 1;
      if (s_cum[mid] > fj) hi = mid; else lo = mid + 1;
    }
    if (lo < Lx && j < yl) { tok = lo; own = s_xm[lo]; }
    if (y_mask && blockIdx.y == 0) y_mask[static_cast<long>(b) * T + j] = j < yl ? 1.f : 0.f;
    // frame-rate encoder output: channel rows are contiguous along T -> coalesced stores, gathered loads (L2-resident)
    const float* mx = mu_x + static_cast<long>(b) * C * Lx;
    float* my = mu_y + static_cast<long>(b) * C * T + j;
    if (tok >= 0 && own != 0.f) {
      for (int c = c0; c < c1; ++c) my[static_cast<long>(c) * T] = __ldg(mx + static_cast<long>(c) * Lx + tok) * own;
    } else {
      for (int c = c0; c < c1; ++c) my[static_cast<long>(c) * T] = 0.f;
    }
  }
}

// The (Lx, T) alignment, written densely (it is part of synthesise()'s result; 35 MB at cfg2, 385 MB at cfg4 - the bulk
// of this operator's HBM traffic).  grid (frame blocks of 512, token blocks of 32, B); a thread owns 4 consecutive frames
// (T is a multiple of 4: fix_len_compatibility) and stores one float4 per token row.  The value is the reference's
// difference of prefix masks, [j < cum[i]] - [j < cum[i-1]], times the token and frame masks.
__global__ void __launch_bounds__(128)
lr_attn_kernel(const float* __restrict__ cum, const float* __restrict__ x_mask, const long long* __restrict__ y_len, int Lx,
               int T, int vec, float* __restrict__ attn) {
  __shared__ float s_cum[33], s_xm[32];
  const int b = blockIdx.z, i0 = blockIdx.y * 32, j = (blockIdx.x * 128 + threadIdx.x) * 4;
  if (threadIdx.x < 33) {
    const int i = i0 - 1 + threadIdx.x;
    s_cum[threadIdx.x] = (i >= 0 && i < Lx) ? cum[static_cast<long>(b) * Lx + i] : 0.f;   // cum[-1] = 0
  } else if (threadIdx.x >= 64 && threadIdx.x < 96) {
    const int i = i0 + threadIdx.x - 64;
    s_xm[threadIdx.x - 64] = i < Lx ? x_mask[static_cast<long>(b) * Lx + i] : 0.f;
  }
  __syncthreads();
  if (j >= T) return;
  const long long yl = y_len[b];
  float fj[4], ym[4];
#pragma unroll
  for (int q = 0; q < 4; ++q) { fj[q] = static_cast<float>(j + q); ym[q] = (j + q) < yl ? 1.f : 0.f; }
  const int ni = min(32, Lx - i0);
  float* ar = attn + (static_cast<long>(b) * Lx + i0) * T + j;
  for (int i = 0; i < ni; ++i) {
    const float lo = s_cum[i], hi = s_cum[i + 1], xm = s_xm[i];
    float v[4];
#pragma unroll
    for (int q = 0; q < 4; ++q) v[q] = ((fj[q] < hi ? 1.f : 0.f) - (fj[q] < lo ? 1.f : 0.f)) * (xm * ym[q]);
    if (vec && j + 3 < T) {
      *reinterpret_cast<float4*>(ar + static_cast<long>(i) * T) = make_float4(v[0], v[1], v[2], v[3]);
    } else {
      for (int q = 0; q < 4 && j + q < T; ++q) ar[static_cast<long>(i) * T + q] = v[q];
    }
  }
}

}  // namespace

void length_regulate(const float* mu_x, const float* w, const float* x_mask, const long long* y_len, int B, int C, int Lx,
                     int T, float* cum, float* mu_y, float* attn, float* y_mask, cudaStream_t st) {
  if (B <= 0 || C <= 0 || Lx <= 0 || T <= 0) throw Error("length_regulate: empty shape");
  if (2 * Lx * sizeof(float) > 48 * 1024) throw Error("length_regulate: Lx > 6144 tokens is not supported");
  lr_cumsum_kernel<<<B, 32, 0, st>>>(w, Lx, cum);
  MB_CUDA(cudaGetLastError());
  lr_fill_kernel<<<dim3((T + 127) / 128, (C + 15) / 16, B), 128, 2 * Lx * sizeof(float), st>>>(mu_x, cum, x_mask, y_len, C, Lx, T, mu_y, y_mask);
  MB_CUDA(cudaGetLastError());
  if (attn) {
    const int vec = (T & 3) == 0 && (reinterpret_cast<uintptr_t>(attn) & 15) == 0;   // rows stay 16-byte aligned
    lr_attn_kernel<<<dim3((T + 511) / 512, (Lx + 31) / 32, B), 128, 0, st>>>(cum, x_mask, y_len, Lx, T, vec, attn);
    MB_CUDA(cudaGetLastError());
  }
}

}  // namespace mb
