// Tensor-core self-attention for the half-precision modes (same contract as attention.cu):
//   out[b, q, h*64:(h+1)*64] = softmax_k(q . k + mask[b, k]) v      2 heads x 64, all L keys, additive 0/1 mask.
// Flash-style: one CTA = 64 query rows of one (batch, head); 4 warps x 16 rows; K/V streamed in 64-key tiles through
// double-buffered, XOR-swizzled shared memory (cp.async: the next tile lands while this one is multiplied);
// S = Q K^T and O += P V on mma.sync.m16n8k16 (fp32 accumulate), online softmax in registers; 4 CTAs per SM.  Attention is ~1 % of the path's FLOPs; the tcgen05 budget went to the conv-GEMMs.
#include "conv.cuh"
#include "launch.h"

namespace mb {

namespace {
constexpr float kLog2e = 1.4426950408889634f;
__device__ __forceinline__ float ex2_ftz(float x) {
  float r;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
  return r;
}
}  // namespace


namespace {

template <typename T> struct MmaT;
template <> struct MmaT<__half> {
  static __device__ __forceinline__ void mma(float (&c)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
    asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
                 : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
  }
  static __device__ __forceinline__ uint32_t pack(float lo, float hi) {
    __half2 h = __floats2half2_rn(lo, hi);
    return *reinterpret_cast<uint32_t*>(&h);
  }
};
template <> struct MmaT<__nv_bfloat16> {
  static __device__ __forceinline__ void mma(float (&c)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
    asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
                 : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
  }
  static __device__ __forceinline__ uint32_t pack(float lo, float hi) {
    __nv_bfloat162 h = __floats2bfloat162_rn(lo, hi);
    return *reinterpret_cast<uint32_t*>(&h);
  }
};

__device__ __forceinline__ void ldsm_x4(uint32_t (&r)[4], uint32_t addr) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(addr));
}
__device__ __forceinline__ void ldsm_x4_trans(uint32_t (&r)[4], uint32_t addr) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0,%1,%2,%3}, [%4];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(addr));
}
// Tile [64 rows][64 halves] = 8 chunks of 16 bytes per row; chunk index XOR (row & 7) kills ldmatrix bank conflicts.
__device__ __forceinline__ uint32_t swz(int row, int chunk) { return static_cast<uint32_t>(row * 128 + ((chunk ^ (row & 7)) << 4)); }

__device__ __forceinline__ void cp_async16(uint32_t dst, const void* src, bool valid) {
  const int bytes = valid ? 16 : 0;   // src-size 0: the 16 destination bytes are zero-filled
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(dst), "l"(src), "r"(bytes) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N> __device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

}  // namespace

template <typename T>
__global__ void __launch_bounds__(128, 4)
attention_mma_kernel(const T* __restrict__ qkv, const float* __restrict__ mask, T* __restrict__ out, int L) {
  __shared__ __align__(128) uint8_t Qs[64 * 128];
  __shared__ __align__(128) uint8_t KVs[2][2][64 * 128];   // [buffer][K / V]
  __shared__ float Msb[2][64];
  const int b = blockIdx.z, h = blockIdx.y, q0 = blockIdx.x * 64;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int g = lane >> 2, tq = lane & 3;
  const T* base = qkv + static_cast<long>(b) * L * 384;
  const uint32_t qs = static_cast<uint32_t>(__cvta_generic_to_shared(Qs));
  const uint32_t kv0 = static_cast<uint32_t>(__cvta_generic_to_shared(&KVs[0][0][0]));

  // tile loader: 64 rows x 8 chunks = 512 x 16 B, 4 per thread
  auto load_tile = [&](uint8_t* dst, int row0, int col0) {
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int idx = tid + i * 128, r = idx >> 3, c = idx & 7;
      uint4 v = make_uint4(0, 0, 0, 0);
      if (row0 + r < L) v = *reinterpret_cast<const uint4*>(base + static_cast<long>(row0 + r) * 384 + col0 + c * 8);
      *reinterpret_cast<uint4*>(dst + swz(r, c)) = v;
    }
  };
  // asynchronous K/V tile: 2 x 512 pieces of 16 B, 8 per thread; rows beyond L are zero-filled
  auto issue_kv = [&](int buf, int k0) {
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int idx = tid + i * 128, r = idx >> 3, c = idx & 7;
      const bool ok = k0 + r < L;
      const T* src = base + static_cast<long>(ok ? k0 + r : 0) * 384 + h * 64 + c * 8;
      cp_async16(kv0 + (buf * 2 + 0) * 8192 + swz(r, c), src + 128, ok);
      cp_async16(kv0 + (buf * 2 + 1) * 8192 + swz(r, c), src + 256, ok);
    }
    if (tid < 64) Msb[buf][tid] = (k0 + tid < L) ? mask[static_cast<long>(b) * L + k0 + tid] : -INFINITY;
    cp_async_commit();
  };
  issue_kv(0, 0);
  load_tile(Qs, q0, h * 64);
  __syncthreads();
  uint32_t qa[4][4];  // A fragments of this warp's 16 query rows for the 4 k-steps over d
  {
    const int mi = lane >> 3, r = warp * 16 + (lane & 7) + (mi & 1) * 8;
#pragma unroll
    for (int kk = 0; kk < 4; ++kk) ldsm_x4(qa[kk], qs + swz(r, kk * 2 + (mi >> 1)));
  }
  float o[8][4];
#pragma unroll
  for (int j = 0; j < 8; ++j) { o[j][0] = o[j][1] = o[j][2] = o[j][3] = 0.f; }
  float m0 = -INFINITY, m1 = -INFINITY, l0 = 0.f, l1 = 0.f;  // rows g and g+8

  int buf = 0;
  for (int k0 = 0; k0 < L; k0 += 64, buf ^= 1) {
    if (k0 + 64 < L) {          // the other buffer was released by the barrier that ended the previous iteration
      issue_kv(buf ^ 1, k0 + 64);
      cp_async_wait<1>();
    } else {
      cp_async_wait<0>();
    }
    __syncthreads();            // this tile (and its mask row) is visible to every warp
    const uint32_t ks = kv0 + (buf * 2 + 0) * 8192, vs = kv0 + (buf * 2 + 1) * 8192;
    const float* Ms = Msb[buf];
    // S = Q K^T : 16 x 64 per warp
    float s[8][4];
#pragma unroll
    for (int j = 0; j < 8; ++j) { s[j][0] = s[j][1] = s[j][2] = s[j][3] = 0.f; }
#pragma unroll
    for (int kk = 0; kk < 4; ++kk) {
#pragma unroll
      for (int jp = 0; jp < 4; ++jp) {  // two key n-tiles per ldmatrix.x4
        uint32_t kb[4];
        const int mi = lane >> 3;
        const int r = jp * 16 + (lane & 7) + (mi >> 1) * 8;   // key row
        ldsm_x4(kb, ks + swz(r, kk * 2 + (mi & 1)));
        MmaT<T>::mma(s[2 * jp], qa[kk], kb[0], kb[1]);
        MmaT<T>::mma(s[2 * jp + 1], qa[kk], kb[2], kb[3]);
      }
    }
    // additive mask (and -inf beyond L), running max
    float mx0 = m0, mx1 = m1;
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const float b0 = Ms[8 * j + tq * 2], b1 = Ms[8 * j + tq * 2 + 1];
      s[j][0] += b0; s[j][1] += b1; s[j][2] += b0; s[j][3] += b1;
      mx0 = fmaxf(mx0, fmaxf(s[j][0], s[j][1]));
      mx1 = fmaxf(mx1, fmaxf(s[j][2], s[j][3]));
    }
    mx0 = fmaxf(mx0, __shfl_xor_sync(0xffffffffu, mx0, 1)); mx0 = fmaxf(mx0, __shfl_xor_sync(0xffffffffu, mx0, 2));
    mx1 = fmaxf(mx1, __shfl_xor_sync(0xffffffffu, mx1, 1)); mx1 = fmaxf(mx1, __shfl_xor_sync(0xffffffffu, mx1, 2));
    // exp(s - m) as ex2(s log2e - m log2e): one FFMA + one MUFU per score (the checked __expf expansion is five
    // instructions); m is finite from the first tile on (every tile 0 holds a valid key), -inf scores give ex2(-inf) = 0
    const float nm0 = -mx0 * kLog2e, nm1 = -mx1 * kLog2e;
    const float sc0 = ex2_ftz(fmaf(m0, kLog2e, nm0)), sc1 = ex2_ftz(fmaf(m1, kLog2e, nm1));  // 0 on the first tile (m = -inf)
    m0 = mx0; m1 = mx1;
    float ps0 = 0.f, ps1 = 0.f;
    uint32_t pa[4][4];  // P as A fragments for the 4 k-steps over keys
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const float p0 = ex2_ftz(fmaf(s[j][0], kLog2e, nm0)), p1 = ex2_ftz(fmaf(s[j][1], kLog2e, nm0));
      const float p2 = ex2_ftz(fmaf(s[j][2], kLog2e, nm1)), p3 = ex2_ftz(fmaf(s[j][3], kLog2e, nm1));
      ps0 += p0 + p1; ps1 += p2 + p3;
      pa[j >> 1][(j & 1) * 2] = MmaT<T>::pack(p0, p1);
      pa[j >> 1][(j & 1) * 2 + 1] = MmaT<T>::pack(p2, p3);
    }
    l0 = l0 * sc0 + ps0; l1 = l1 * sc1 + ps1;   // per-thread partial row sums (quad-reduced at the end)
#pragma unroll
    for (int j = 0; j < 8; ++j) { o[j][0] *= sc0; o[j][1] *= sc0; o[j][2] *= sc1; o[j][3] *= sc1; }
    // O += P V
#pragma unroll
    for (int kk = 0; kk < 4; ++kk) {
#pragma unroll
      for (int jp = 0; jp < 4; ++jp) {  // two d n-tiles per ldmatrix.x4.trans
        uint32_t vb[4];
        const int mi = lane >> 3;
        const int r = kk * 16 + (lane & 7) + (mi & 1) * 8;    // key row
        ldsm_x4_trans(vb, vs + swz(r, jp * 2 + (mi >> 1)));
        MmaT<T>::mma(o[2 * jp], pa[kk], vb[0], vb[1]);
        MmaT<T>::mma(o[2 * jp + 1], pa[kk], vb[2], vb[3]);
      }
    }
    __syncthreads();            // every warp is done with this buffer before it is refilled
  }
  l0 += __shfl_xor_sync(0xffffffffu, l0, 1); l0 += __shfl_xor_sync(0xffffffffu, l0, 2);
  l1 += __shfl_xor_sync(0xffffffffu, l1, 1); l1 += __shfl_xor_sync(0xffffffffu, l1, 2);
  const float i0 = 1.f / l0, i1 = 1.f / l1;
  const int r0 = q0 + warp * 16 + g, r1 = r0 + 8;
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    const int col = h * 64 + 8 * j + tq * 2;
    if (r0 < L) *reinterpret_cast<uint32_t*>(out + (static_cast<long>(b) * L + r0) * 128 + col) = MmaT<T>::pack(o[j][0] * i0, o[j][1] * i0);
    if (r1 < L) *reinterpret_cast<uint32_t*>(out + (static_cast<long>(b) * L + r1) * 128 + col) = MmaT<T>::pack(o[j][2] * i1, o[j][3] * i1);
  }
}

template <typename T>
cudaError_t launch_attention_mma(const T* qkv, const float* mask, T* out, int B, int L, cudaStream_t st) {
  dim3 grid((L + 63) / 64, 2, B);
  attention_mma_kernel<T><<<grid, 128, 0, st>>>(qkv, mask, out, L);
  return cudaGetLastError();
}
template cudaError_t launch_attention_mma<__half>(const __half*, const float*, __half*, int, int, cudaStream_t);
template cudaError_t launch_attention_mma<__nv_bfloat16>(const __nv_bfloat16*, const float*, __nv_bfloat16*, int, int,
                                                         cudaStream_t);

}  // namespace mb
