// Self-attention of the estimator's BasicTransformerBlock (transformer.py:266-271 -> diffusers 0.25.0
// Attention/AttnProcessor2_0): 2 heads x 64, non-causal, softmax(q k^T / 8 + mask[key]) v where the float 0/1
// mask is ADDED to the logits -- padded keys are attended.  The 1/8 scale is folded
// into W_q at pack time (exact: power of two).
//
// attention_simt_kernel: one thread per query row, K/V tiles broadcast from shared memory, online softmax in
// fp32.  Used by the fp32 mode; the half modes use the tensor-core kernel in attention_mma.cu when present.
#include "conv.cuh"
#include "launch.h"

namespace mb {

template <typename T>
__global__ void __launch_bounds__(128)
attention_simt_kernel(const T* __restrict__ qkv, const float* __restrict__ mask, T* __restrict__ out, int L) {
  __shared__ __align__(16) float Ks[32][64];
  __shared__ __align__(16) float Vs[32][64];
  __shared__ float Ms[32];
  const int b = blockIdx.z, h = blockIdx.y;
  const int tid = threadIdx.x;
  const int qi = blockIdx.x * 128 + tid;
  const bool qvalid = qi < L;
  const T* base = qkv + static_cast<long>(b) * L * 384;

  float q[64], o[64];
#pragma unroll
  for (int d = 0; d < 64; ++d) o[d] = 0.f;
  if (qvalid) {
#pragma unroll
    for (int j = 0; j < 8; ++j) Elem<T>::ld8(base + static_cast<long>(qi) * 384 + h * 64 + 8 * j, q + 8 * j);
  } else {
#pragma unroll
    for (int d = 0; d < 64; ++d) q[d] = 0.f;
  }
  float mrun = -INFINITY, lrun = 0.f;
  const int lk = tid >> 2, lseg = (tid & 3) * 16;
  for (int k0 = 0; k0 < L; k0 += 32) {
    __syncthreads();
    {
      const int kj = k0 + lk;
      float kv[16], vv[16];
      if (kj < L) {
        const T* kp = base + static_cast<long>(kj) * 384 + 128 + h * 64 + lseg;
        Elem<T>::ld8(kp, kv); Elem<T>::ld8(kp + 8, kv + 8);
        Elem<T>::ld8(kp + 128, vv); Elem<T>::ld8(kp + 136, vv + 8);
      } else {
#pragma unroll
        for (int i = 0; i < 16; ++i) { kv[i] = 0.f; vv[i] = 0.f; }
      }
#pragma unroll
      for (int i = 0; i < 16; ++i) { Ks[lk][lseg + i] = kv[i]; Vs[lk][lseg + i] = vv[i]; }
      if (tid < 32) Ms[tid] = (k0 + tid < L) ? mask[static_cast<long>(b) * L + k0 + tid] : 0.f;
    }
    __syncthreads();
    const int nk = min(32, L - k0);
    for (int j0 = 0; j0 < nk; j0 += 8) {
      float s[8];
      float mx = -INFINITY;
#pragma unroll
      for (int jj = 0; jj < 8; ++jj) {
        const int j = j0 + jj;
        float acc = 0.f;
        if (j < nk) {
          const float4* kr = reinterpret_cast<const float4*>(Ks[j]);
#pragma unroll
          for (int d4 = 0; d4 < 16; ++d4) {
            const float4 kk = kr[d4];
            acc = fmaf(q[4 * d4], kk.x, acc); acc = fmaf(q[4 * d4 + 1], kk.y, acc);
            acc = fmaf(q[4 * d4 + 2], kk.z, acc); acc = fmaf(q[4 * d4 + 3], kk.w, acc);
          }
          acc += Ms[j];
        } else {
          acc = -INFINITY;
        }
        s[jj] = acc;
        mx = fmaxf(mx, acc);
      }
      if (mx > mrun) {
        const float sc = expf(mrun - mx);  // exp(-inf) = 0 on the first chunk
        lrun *= sc;
#pragma unroll
        for (int d = 0; d < 64; ++d) o[d] *= sc;
        mrun = mx;
      }
#pragma unroll
      for (int jj = 0; jj < 8; ++jj) {
        const int j = j0 + jj;
        if (j < nk) {
          const float p = expf(s[jj] - mrun);
          lrun += p;
          const float4* vr = reinterpret_cast<const float4*>(Vs[j]);
#pragma unroll
          for (int d4 = 0; d4 < 16; ++d4) {
            const float4 vv = vr[d4];
            o[4 * d4] = fmaf(p, vv.x, o[4 * d4]); o[4 * d4 + 1] = fmaf(p, vv.y, o[4 * d4 + 1]);
            o[4 * d4 + 2] = fmaf(p, vv.z, o[4 * d4 + 2]); o[4 * d4 + 3] = fmaf(p, vv.w, o[4 * d4 + 3]);
          }
        }
      }
    }
  }
  if (qvalid) {
    const float inv = 1.f / lrun;
    T* op = out + (static_cast<long>(b) * L + qi) * 128 + h * 64;
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      float r[8];
#pragma unroll
      for (int i = 0; i < 8; ++i) r[i] = o[8 * j + i] * inv;
      Elem<T>::st8(op + 8 * j, r);
    }
  }
}

template <typename T>
cudaError_t launch_attention_simt(const T* qkv, const float* mask, T* out, int B, int L, cudaStream_t st) {
  dim3 grid((L + 127) / 128, 2, B);
  attention_simt_kernel<T><<<grid, 128, 0, st>>>(qkv, mask, out, L);
  return cudaGetLastError();
}
template cudaError_t launch_attention_simt<float>(const float*, const float*, float*, int, int, cudaStream_t);
template cudaError_t launch_attention_simt<__nv_bfloat16>(const __nv_bfloat16*, const float*, __nv_bfloat16*, int, int,
                                                          cudaStream_t);
template cudaError_t launch_attention_simt<__half>(const __half*, const float*, __half*, int, int, cudaStream_t);

}  // namespace mb
