// fp32 FFMA kernels of the training path: BASECFM.compute_loss forward + backward of the U-Net estimator
// (matcha/models/components/flow_matching.py:87-118 over decoder.py:363-443, transformer.py:243-316).
//
// Everything here works on [B][rows][C] fp32 activations (channels contiguous) and on the parameters IN THE REFERENCE
// LAYOUT, read and written in place through strides: a Conv1d / ConvTranspose1d / Linear, its data gradient and its
// weight gradient are all "one strided GEMM per kernel tap" (sgemm_kernel), so the optimiser can keep updating the
// module's own tensors without any repacking between steps.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace mb {
namespace trn {

// C[z](m, n) (op)= alpha * sum_k A[z](m, k) * B[z](k, n), every operand addressed through element strides.
// blockIdx.z = z1 * nz2 + z2 selects the batch (two batch dimensions: utterance and attention head).
struct Gemm {
  const float* A; const float* B; float* C;
  int M, N, K;
  long sAm, sAk, sBk, sBn, sCm, sCn;
  int nz2;
  long bA1, bA2, bB1, bB2, bC1, bC2;
  float alpha;
  int mode;   // 0: C = v, 1: C += v, 2: atomicAdd(C, v) (batch-reduced weight gradients: bC1 = bC2 = 0)
};

constexpr int BM = 64, BN = 64, BK = 16;

__global__ void __launch_bounds__(256) sgemm_kernel(const Gemm p) {
  __shared__ float As[BK][BM + 4];
  __shared__ float Bs[BK][BN + 4];
  const int z1 = blockIdx.z / p.nz2, z2 = blockIdx.z % p.nz2;
  const float* __restrict__ A = p.A + z1 * p.bA1 + z2 * p.bA2;
  const float* __restrict__ B = p.B + z1 * p.bB1 + z2 * p.bB2;
  float* C = p.C + z1 * p.bC1 + z2 * p.bC2;
  const int m0 = blockIdx.x * BM, n0 = blockIdx.y * BN;
  const int tx = threadIdx.x & 15, ty = threadIdx.x >> 4;
  const bool a_kfast = p.sAk == 1, b_kfast = p.sBk == 1;
  float acc[4][4];
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;
  for (int k0 = 0; k0 < p.K; k0 += BK) {
#pragma unroll
    for (int i = threadIdx.x; i < BM * BK; i += 256) {
      int m, k;
      if (a_kfast) { k = i % BK; m = i / BK; } else { m = i % BM; k = i / BM; }
      float v = 0.f;
      if (m0 + m < p.M && k0 + k < p.K) v = A[(m0 + m) * p.sAm + (k0 + k) * p.sAk];
      As[k][m] = v;
    }
#pragma unroll
    for (int i = threadIdx.x; i < BN * BK; i += 256) {
      int n, k;
      if (b_kfast) { k = i % BK; n = i / BK; } else { n = i % BN; k = i / BN; }
      float v = 0.f;
      if (n0 + n < p.N && k0 + k < p.K) v = B[(k0 + k) * p.sBk + (n0 + n) * p.sBn];
      Bs[k][n] = v;
    }
    __syncthreads();
#pragma unroll
    for (int k = 0; k < BK; ++k) {
      const float4 a4 = *reinterpret_cast<const float4*>(&As[k][ty * 4]);
      const float4 b4 = *reinterpret_cast<const float4*>(&Bs[k][tx * 4]);
      const float a[4] = {a4.x, a4.y, a4.z, a4.w}, b[4] = {b4.x, b4.y, b4.z, b4.w};
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(a[i], b[j], acc[i][j]);
    }
    __syncthreads();
  }
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int m = m0 + ty * 4 + i;
    if (m >= p.M) continue;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int n = n0 + tx * 4 + j;
      if (n >= p.N) continue;
      float* c = C + m * p.sCm + n * p.sCn;
      const float v = p.alpha * acc[i][j];
      if (p.mode == 0) *c = v;
      else if (p.mode == 1) *c += v;
      else atomicAdd(c, v);
    }
  }
}

// The same contract on the tensor cores: 128 x 128 x 16 tiles, 8 warps of 64 x 32, mma.sync.m16n8k8 TF32 with fp32
// accumulation, operands through a three-stage cp.async ring (one __syncthreads per k step).  SPLIT = 3 is the error-compensated "3xTF32" product
//     a b ~= a_hi b_hi + a_lo b_hi + a_hi b_lo,   x_hi = tf32(x), x_lo = tf32(x - x_hi)
// which keeps fp32-level accuracy (the dropped a_lo b_lo term is 2^-22 relative) at three MMAs per product - still far
// above the FFMA rate; SPLIT = 1 is plain TF32 (10-bit mantissa operands).
constexpr int TN = 128, TK = 16, TPAD = 8, TSTAGES = 3;
constexpr int tgemm_smem_bytes(int tm) { return TSTAGES * TK * (tm + TPAD + TN + TPAD) * 4; }

// 4-byte asynchronous copy global -> shared (any element stride); `ok == false` writes a zero without reading
__device__ __forceinline__ void cp_async4(float* dst, const float* src, bool ok) {
  const uint32_t d = static_cast<uint32_t>(__cvta_generic_to_shared(dst));
  const int n = ok ? 4 : 0;
  asm volatile("cp.async.ca.shared.global [%0], [%1], 4, %2;" ::"r"(d), "l"(src), "r"(n) : "memory");
}

__device__ __forceinline__ uint32_t to_tf32(float x) {
  uint32_t r;
  asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(r) : "f"(x));
  return r;
}
__device__ __forceinline__ void mma_tf32(float (&d)[4], const uint32_t (&a)[4], const uint32_t (&b)[2]) {
  asm volatile("mma.sync.aligned.m16n8k8.row.col.f32.tf32.tf32.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
               : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
               : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b[0]), "r"(b[1]));
}

// TM = 128 (warps of 64 x 32) or 64 (warps of 32 x 32, for M <= 64)
template <int SPLIT, int TM>
__global__ void __launch_bounds__(256) tgemm_kernel(const Gemm p) {
  constexpr int MI = TM / 32;              // 16-row MMA tiles per warp
  constexpr int RA = TM * TK / 256;        // A elements per thread per k tile
  extern __shared__ float tg_smem[];
  float (*As)[TK][TM + TPAD] = reinterpret_cast<float (*)[TK][TM + TPAD]>(tg_smem);
  float (*Bs)[TK][TN + TPAD] = reinterpret_cast<float (*)[TK][TN + TPAD]>(tg_smem + TSTAGES * TK * (TM + TPAD));
  const int z1 = blockIdx.z / p.nz2, z2 = blockIdx.z % p.nz2;
  const float* __restrict__ A = p.A + z1 * p.bA1 + z2 * p.bA2;
  const float* __restrict__ B = p.B + z1 * p.bB1 + z2 * p.bB2;
  float* C = p.C + z1 * p.bC1 + z2 * p.bC2;
  const int m0 = blockIdx.x * TM, n0 = blockIdx.y * TN;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, g = lane >> 2, t = lane & 3;
  const int wm = (warp >> 2) * (TM / 2), wn = (warp & 3) * 32;
  const bool a_kfast = p.sAk == 1, b_kfast = p.sBk == 1;
  float acc[4][4][4];
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j)
#pragma unroll
      for (int r = 0; r < 4; ++r) acc[i][j][r] = 0.f;
  // Operand addressing hoisted out of the k loop: element r of this thread differs from element 0 by a constant (m, k)
  // step, so one base pointer per operand, one 64-bit step and a pointer bump per k tile replace sixteen
  // multiply-add address computations per tile (measured: with them the kernel spent more time forming addresses than
  // in the MMAs, 21 TFLOP/s).
  int am, ak, a_dm, a_dk, bn, bk, b_dn, b_dk;
  if (a_kfast) { ak = threadIdx.x % TK; am = threadIdx.x / TK; a_dm = 256 / TK; a_dk = 0; }
  else { am = threadIdx.x % TM; ak = threadIdx.x / TM; a_dm = 0; a_dk = 256 / TM; }
  if (b_kfast) { bk = threadIdx.x % TK; bn = threadIdx.x / TK; b_dn = 256 / TK; b_dk = 0; }
  else { bn = threadIdx.x % TN; bk = threadIdx.x / TN; b_dn = 0; b_dk = 256 / TN; }
  const float* pa = A + static_cast<long>(m0 + am) * p.sAm + static_cast<long>(ak) * p.sAk;
  const float* pb = B + static_cast<long>(n0 + bn) * p.sBn + static_cast<long>(bk) * p.sBk;
  const long a_step = static_cast<long>(a_dm) * p.sAm + static_cast<long>(a_dk) * p.sAk;
  const long b_step = static_cast<long>(b_dn) * p.sBn + static_cast<long>(b_dk) * p.sBk;
  const long a_tile = static_cast<long>(TK) * p.sAk, b_tile = static_cast<long>(TK) * p.sBk;
  // Tiles travel global -> shared with 4-byte cp.async (any stride, zero fill outside the matrix) through a ring of
  // TSTAGES buffers: two tiles are in flight while one is multiplied.  With register staging and one tile of look-ahead
  // every k step waited out a memory round trip (measured 3.2 us per k step against 1.2 us of MMA issue).
  auto issue = [&](int k0, int buf) {
    if (k0 < p.K) {
#pragma unroll
      for (int r = 0; r < RA; ++r) {
        const bool oka = m0 + am + r * a_dm < p.M && k0 + ak + r * a_dk < p.K;
        cp_async4(&As[buf][ak + r * a_dk][am + r * a_dm], oka ? pa + r * a_step : A, oka);
      }
#pragma unroll
      for (int r = 0; r < 8; ++r) {
        const bool okb = n0 + bn + r * b_dn < p.N && k0 + bk + r * b_dk < p.K;
        cp_async4(&Bs[buf][bk + r * b_dk][bn + r * b_dn], okb ? pb + r * b_step : B, okb);
      }
      pa += a_tile;
      pb += b_tile;
    }
    asm volatile("cp.async.commit_group;" ::: "memory");
  };
#pragma unroll
  for (int s = 0; s < TSTAGES - 1; ++s) issue(s * TK, s);
  int buf = 0;
  for (int k0 = 0; k0 < p.K; k0 += TK) {
    asm volatile("cp.async.wait_group %0;" ::"n"(TSTAGES - 2) : "memory");
    __syncthreads();      // this tile has landed for every thread; the buffer refilled next is no longer being read
    issue(k0 + (TSTAGES - 1) * TK, (buf + TSTAGES - 1) % TSTAGES);
#pragma unroll
    for (int ks = 0; ks < TK; ks += 8) {
      uint32_t bh[4][2], bl[4][2];
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const float b0 = Bs[buf][ks + t][wn + j * 8 + g], b1 = Bs[buf][ks + t + 4][wn + j * 8 + g];
        bh[j][0] = to_tf32(b0); bh[j][1] = to_tf32(b1);
        if (SPLIT == 3) { bl[j][0] = to_tf32(b0 - __uint_as_float(bh[j][0])); bl[j][1] = to_tf32(b1 - __uint_as_float(bh[j][1])); }
      }
#pragma unroll
      for (int i = 0; i < MI; ++i) {
        const float a[4] = {As[buf][ks + t][wm + i * 16 + g], As[buf][ks + t][wm + i * 16 + g + 8],
                            As[buf][ks + t + 4][wm + i * 16 + g], As[buf][ks + t + 4][wm + i * 16 + g + 8]};
        uint32_t ah[4], al[4];
#pragma unroll
        for (int r = 0; r < 4; ++r) {
          ah[r] = to_tf32(a[r]);
          if (SPLIT == 3) al[r] = to_tf32(a[r] - __uint_as_float(ah[r]));
        }
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          if (SPLIT == 3) {          // small terms first
            mma_tf32(acc[i][j], al, bh[j]);
            mma_tf32(acc[i][j], ah, bl[j]);
          }
          mma_tf32(acc[i][j], ah, bh[j]);
        }
      }
    }
    buf = (buf + 1) % TSTAGES;
  }
  // Epilogue.  `C += v` (conv taps after the first, residual gradients) is done 16 values at a time: all the old
  // values are requested first, then added and stored.  One read-modify-write at a time, each waiting out a full memory
  // round trip, was half of this kernel's run time (64 dependent round trips per thread).
#pragma unroll
  for (int i = 0; i < MI; ++i) {
    float* cp[16];
    float old[16];
#pragma unroll
    for (int j = 0; j < 4; ++j)
#pragma unroll
      for (int r = 0; r < 4; ++r) {
        const int m = m0 + wm + i * 16 + g + (r >> 1) * 8, n = n0 + wn + j * 8 + 2 * t + (r & 1);
        cp[j * 4 + r] = (m < p.M && n < p.N) ? C + m * p.sCm + n * p.sCn : nullptr;
      }
    if (p.mode == 1) {
#pragma unroll
      for (int q = 0; q < 16; ++q) old[q] = cp[q] ? *cp[q] : 0.f;
    }
#pragma unroll
    for (int q = 0; q < 16; ++q) {
      if (!cp[q]) continue;
      const float v = p.alpha * acc[i][q >> 2][q & 3];
      if (p.mode == 0) *cp[q] = v;
      else if (p.mode == 1) *cp[q] = old[q] + v;
      else atomicAdd(cp[q], v);
    }
  }
}

__device__ __forceinline__ float warp_sum_f(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
// sum over a 256-thread block; sh needs 8 floats; result broadcast to every thread
__device__ __forceinline__ float block_sum256(float v, float* sh) {
  v = warp_sum_f(v);
  __syncthreads();
  if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = v;
  __syncthreads();
  float t = 0.f;
#pragma unroll
  for (int i = 0; i < 8; ++i) t += sh[i];
  return t;
}
__device__ __forceinline__ float block_max256(float v, float* sh) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  __syncthreads();
  if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = v;
  __syncthreads();
  float t = sh[0];
#pragma unroll
  for (int i = 1; i < 8; ++i) t = fmaxf(t, sh[i]);
  return t;
}

// Mish (torch: x * tanh(softplus(x)), softplus threshold 20) and its derivative.
__device__ __forceinline__ float mish_t(float x) {
  if (x > 20.f) return x * tanhf(x);
  const float n = expf(x), w = n * (n + 2.f);
  return x * (w / (w + 2.f));
}
__device__ __forceinline__ float mish_grad_t(float x) {
  if (x > 20.f) { const float th = tanhf(x); return th + x * (1.f - th * th); }
  const float n = expf(x), w = n * (n + 2.f);
  const float th = w / (w + 2.f);          // tanh(softplus(x))
  const float sg = n / (1.f + n);          // sigmoid(x) = d softplus / dx
  return th + x * (1.f - th * th) * sg;
}

// Y[r][n] = bias ? bias[n] : 0 over a [rows][ld] view (n < N)
__global__ void fill_rows_kernel(float* __restrict__ Y, const float* __restrict__ bias, long rows, int N, int ld) {
  const long i = blockIdx.x * 256L + threadIdx.x;
  if (i >= rows * N) return;
  const long r = i / N;
  const int n = static_cast<int>(i % N);
  Y[r * ld + n] = bias ? bias[n] : 0.f;
}
// Dropout (decoder.py:206 -> transformer.py:124, :200; rate configs/model/decoder/default.yaml:2).  The keep decision of
// element i at site s is a pure function of (seed, s, i): splitmix64(seed + K (s + 1) + i) >> 32 >= rate * 2^32, so the
// backward pass regenerates the mask instead of storing it and a test can inject the same masks into the oracle.
__host__ __device__ inline unsigned int dropout_bits(unsigned long long seed, unsigned int site, unsigned long long i) {
  unsigned long long z = seed + 0x632BE59BD9B4E019ull * (site + 1ull) + i + 0x9E3779B97F4A7C15ull;
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
  z ^= z >> 31;
  return static_cast<unsigned int>(z >> 32);
}
// out[i] = in[i] * keep(i) / (1 - rate)   (in == out allowed)
__global__ void dropout_kernel(const float* __restrict__ in, float* __restrict__ out, long n, unsigned long long seed,
                               unsigned int site, unsigned int thresh, float inv_keep) {
  const long i = blockIdx.x * 256L + threadIdx.x;
  if (i >= n) return;
  out[i] = dropout_bits(seed, site, static_cast<unsigned long long>(i)) >= thresh ? in[i] * inv_keep : 0.f;
}
// X[r][:] *= mask[r]
__global__ void mul_mask_kernel(float* __restrict__ X, const float* __restrict__ mask, long rows, int N) {
  const long i = blockIdx.x * 256L + threadIdx.x;
  if (i >= rows * N) return;
  X[i] *= mask[i / N];
}
// out[n] += sum_r X[r][n]  (bias gradients); one block per 128 rows
__global__ void colsum_kernel(const float* __restrict__ X, long rows, int N, int ld, float* __restrict__ out) {
  const long r0 = blockIdx.x * 128L, r1 = min(rows, r0 + 128L);
  for (int n = threadIdx.x; n < N; n += 256) {
    float s = 0.f;
    for (long r = r0; r < r1; ++r) s += X[r * ld + n];
    atomicAdd(out + n, s);
  }
}
// out[0] += sum x[i]
__global__ void sum_kernel(const float* __restrict__ x, long n, float* __restrict__ out) {
  __shared__ float sh[8];
  float s = 0.f;
  for (long i = blockIdx.x * 256L + threadIdx.x; i < n; i += 256L * gridDim.x) s += x[i];
  s = block_sum256(s, sh);
  if (threadIdx.x == 0) atomicAdd(out, s);
}

// GroupNorm(8 groups of 32 channels, eps 1e-5, biased variance over 32 x L INCLUDING padded frames: decoder.py:37).
// grid (8, B); stats[b][g] = {mean, rstd}
__global__ void __launch_bounds__(256) gn_stats_kernel(const float* __restrict__ raw, int L, float* __restrict__ stats) {
  __shared__ float sh[8];
  const int g = blockIdx.x, b = blockIdx.y, c = threadIdx.x & 31;
  const float* base = raw + (static_cast<long>(b) * L) * 256 + g * 32 + c;
  float s = 0.f;
  for (int t = threadIdx.x >> 5; t < L; t += 8) s += base[static_cast<long>(t) * 256];
  const float n = 32.f * L;
  const float mean = block_sum256(s, sh) / n;
  float q = 0.f;
  for (int t = threadIdx.x >> 5; t < L; t += 8) { const float d = base[static_cast<long>(t) * 256] - mean; q += d * d; }
  const float var = block_sum256(q, sh) / n;
  if (threadIdx.x == 0) { stats[(b * 8 + g) * 2] = mean; stats[(b * 8 + g) * 2 + 1] = rsqrtf(var + 1e-5f); }
}
// out = (mish(GN(raw)) + tadd[b][c]) * mask    (Block1D tail decoder.py:37-43 + the time projection decoder.py:58)
__global__ void gn_mish_fwd_kernel(const float* __restrict__ raw, const float* __restrict__ stats,
                                   const float* __restrict__ gamma, const float* __restrict__ beta,
                                   const float* __restrict__ tadd, int tadd_bstride, const float* __restrict__ mask,
                                   float* __restrict__ out, int L, long total) {
  const long i = blockIdx.x * 256L + threadIdx.x;
  if (i >= total) return;
  const int c = static_cast<int>(i & 255);
  const long row = i >> 8;
  const int b = static_cast<int>(row / L);
  const float mean = stats[(b * 8 + (c >> 5)) * 2], rstd = stats[(b * 8 + (c >> 5)) * 2 + 1];
  const float y = (raw[i] - mean) * rstd * gamma[c] + beta[c];
  float a = mish_t(y);
  if (tadd) a += tadd[static_cast<long>(b) * tadd_bstride + c];
  out[i] = a * mask[row];
}
// Backward, pass 1: grid (8, B).  sums[b][g] = {sum dxhat, sum dxhat*xhat}; dgamma/dbeta accumulated atomically;
// dtadd[b][c] = sum_t dOut*mask (written, not accumulated).
__global__ void __launch_bounds__(256) gn_mish_bwd_reduce_kernel(
    const float* __restrict__ dout, const float* __restrict__ raw, const float* __restrict__ stats,
    const float* __restrict__ gamma, const float* __restrict__ beta, const float* __restrict__ mask, int L,
    float* __restrict__ sums, float* __restrict__ dgamma, float* __restrict__ dbeta, float* __restrict__ dtadd,
    int dtadd_bstride) {
  __shared__ float sh[8];
  __shared__ float shc[3][8][32];
  const int g = blockIdx.x, b = blockIdx.y, c = g * 32 + (threadIdx.x & 31);
  const float mean = stats[(b * 8 + g) * 2], rstd = stats[(b * 8 + g) * 2 + 1];
  const float gm = gamma[c], bt = beta[c];
  float s1 = 0.f, s2 = 0.f, dg = 0.f, db = 0.f, dt = 0.f;
  for (int t = threadIdx.x >> 5; t < L; t += 8) {
    const long row = static_cast<long>(b) * L + t;
    const float xh = (raw[row * 256 + c] - mean) * rstd;
    const float da = dout[row * 256 + c] * mask[row];
    const float dy = da * mish_grad_t(xh * gm + bt);
    dt += da; dg += dy * xh; db += dy;
    const float dxh = dy * gm;
    s1 += dxh; s2 += dxh * xh;
  }
  const float S1 = block_sum256(s1, sh);
  const float S2 = block_sum256(s2, sh);
  if (threadIdx.x == 0) { sums[(b * 8 + g) * 2] = S1; sums[(b * 8 + g) * 2 + 1] = S2; }
  shc[0][threadIdx.x >> 5][threadIdx.x & 31] = dg;
  shc[1][threadIdx.x >> 5][threadIdx.x & 31] = db;
  shc[2][threadIdx.x >> 5][threadIdx.x & 31] = dt;
  __syncthreads();
  if (threadIdx.x < 32) {
    float a0 = 0.f, a1 = 0.f, a2 = 0.f;
#pragma unroll
    for (int w = 0; w < 8; ++w) { a0 += shc[0][w][threadIdx.x]; a1 += shc[1][w][threadIdx.x]; a2 += shc[2][w][threadIdx.x]; }
    if (dgamma) atomicAdd(dgamma + c, a0);
    if (dbeta) atomicAdd(dbeta + c, a1);
    if (dtadd) dtadd[static_cast<long>(b) * dtadd_bstride + c] = a2;
  }
}
// Backward, pass 2: draw = rstd * (dxhat - S1/n - xhat * S2/n), n = 32 L
__global__ void gn_mish_bwd_apply_kernel(const float* __restrict__ dout, const float* __restrict__ raw,
                                         const float* __restrict__ stats, const float* __restrict__ sums,
                                         const float* __restrict__ gamma, const float* __restrict__ beta,
                                         const float* __restrict__ mask, float* __restrict__ draw, int L, long total) {
  const long i = blockIdx.x * 256L + threadIdx.x;
  if (i >= total) return;
  const int c = static_cast<int>(i & 255);
  const long row = i >> 8;
  const int b = static_cast<int>(row / L), sg = b * 8 + (c >> 5);
  const float mean = stats[sg * 2], rstd = stats[sg * 2 + 1];
  const float xh = (raw[i] - mean) * rstd;
  const float dy = dout[i] * mask[row] * mish_grad_t(xh * gamma[c] + beta[c]);
  const float inv_n = 1.f / (32.f * L);
  draw[i] = rstd * (dy * gamma[c] - sums[sg * 2] * inv_n - xh * sums[sg * 2 + 1] * inv_n);
}

// LayerNorm over 256 channels (eps 1e-5; transformer.py:262,296): one warp per row, 8 channels per lane.
__global__ void __launch_bounds__(256) ln_fwd_kernel(const float* __restrict__ in, const float* __restrict__ g,
                                                     const float* __restrict__ bta, float* __restrict__ out, long rows) {
  const long row = blockIdx.x * 8L + (threadIdx.x >> 5);
  if (row >= rows) return;
  const int lane = threadIdx.x & 31;
  float x[8];
  float s = 0.f;
#pragma unroll
  for (int i = 0; i < 8; ++i) { x[i] = in[row * 256 + lane + 32 * i]; s += x[i]; }
  const float mean = warp_sum_f(s) * (1.f / 256.f);
  float q = 0.f;
#pragma unroll
  for (int i = 0; i < 8; ++i) { const float d = x[i] - mean; q += d * d; }
  const float rstd = rsqrtf(warp_sum_f(q) * (1.f / 256.f) + 1e-5f);
#pragma unroll
  for (int i = 0; i < 8; ++i) { const int c = lane + 32 * i; out[row * 256 + c] = (x[i] - mean) * rstd * g[c] + bta[c]; }
}
// din += LN'(dout) ; dgamma += dout * xhat ; dbeta += dout.  Each warp strides over rows, partial sums in registers.
__global__ void __launch_bounds__(256) ln_bwd_kernel(const float* __restrict__ dout, const float* __restrict__ in,
                                                     const float* __restrict__ g, float* __restrict__ din,
                                                     float* __restrict__ dgamma, float* __restrict__ dbeta, long rows) {
  __shared__ float shg[8][256], shb[8][256];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  float dg[8], db[8], gm[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) { dg[i] = 0.f; db[i] = 0.f; gm[i] = g[lane + 32 * i]; }
  for (long row = blockIdx.x * 8L + warp; row < rows; row += 8L * gridDim.x) {
    float x[8], dy[8];
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < 8; ++i) { x[i] = in[row * 256 + lane + 32 * i]; dy[i] = dout[row * 256 + lane + 32 * i]; s += x[i]; }
    const float mean = warp_sum_f(s) * (1.f / 256.f);
    float q = 0.f;
#pragma unroll
    for (int i = 0; i < 8; ++i) { x[i] -= mean; q += x[i] * x[i]; }
    const float rstd = rsqrtf(warp_sum_f(q) * (1.f / 256.f) + 1e-5f);
    float m1 = 0.f, m2 = 0.f;
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      x[i] *= rstd;                      // xhat
      dg[i] += dy[i] * x[i]; db[i] += dy[i];
      dy[i] *= gm[i];                    // dxhat
      m1 += dy[i]; m2 += dy[i] * x[i];
    }
    m1 = warp_sum_f(m1) * (1.f / 256.f);
    m2 = warp_sum_f(m2) * (1.f / 256.f);
#pragma unroll
    for (int i = 0; i < 8; ++i) din[row * 256 + lane + 32 * i] += rstd * (dy[i] - m1 - x[i] * m2);
  }
#pragma unroll
  for (int i = 0; i < 8; ++i) { shg[warp][lane + 32 * i] = dg[i]; shb[warp][lane + 32 * i] = db[i]; }
  __syncthreads();
  const int c = threadIdx.x;
  float a = 0.f, bsum = 0.f;
#pragma unroll
  for (int w = 0; w < 8; ++w) { a += shg[w][c]; bsum += shb[w][c]; }
  if (dgamma) atomicAdd(dgamma + c, a);
  if (dbeta) atomicAdd(dbeta + c, bsum);
}

// SnakeBeta (transformer.py:64-80): f = y + sin^2(y e^alpha) / (e^beta + 1e-9), 1024 channels
__global__ void snake_fwd_kernel(const float* __restrict__ pre, const float* __restrict__ alpha,
                                 const float* __restrict__ beta, float* __restrict__ out, long total) {
  const long i = blockIdx.x * 256L + threadIdx.x;
  if (i >= total) return;
  const int c = static_cast<int>(i & 1023);
  const float a = expf(alpha[c]), ib = 1.f / (expf(beta[c]) + 1e-9f);
  const float y = pre[i], s = sinf(y * a);
  out[i] = y + ib * s * s;
}
// d (in place: d := dpre), dalpha, dbeta.  One block per 64 rows, 4 channels per thread.
__global__ void __launch_bounds__(256) snake_bwd_kernel(float* __restrict__ d, const float* __restrict__ pre,
                                                        const float* __restrict__ alpha, const float* __restrict__ beta,
                                                        float* __restrict__ dalpha, float* __restrict__ dbeta, long rows) {
  const long r0 = blockIdx.x * 64L, r1 = min(rows, r0 + 64L);
#pragma unroll
  for (int q = 0; q < 4; ++q) {
    const int c = threadIdx.x + 256 * q;
    const float a = expf(alpha[c]), eb = expf(beta[c]), ib = 1.f / (eb + 1e-9f);
    float da = 0.f, db = 0.f;
    for (long r = r0; r < r1; ++r) {
      const float y = pre[r * 1024 + c], g = d[r * 1024 + c];
      float s, co;
      sincosf(y * a, &s, &co);
      const float s2x = 2.f * s * co;                 // sin(2 a y)
      d[r * 1024 + c] = g * (1.f + ib * a * s2x);
      da += g * ib * s2x * y * a;                     // d/dalpha: a' = a
      db -= g * s * s * ib * ib * eb;                 // d/dbeta of 1/(e^beta + eps)
    }
    if (dalpha) atomicAdd(dalpha + c, da);
    if (dbeta) atomicAdd(dbeta + c, db);
  }
}

// softmax over the last axis of S[(b, h, i)][L] with the ADDITIVE float key mask, in place
__global__ void __launch_bounds__(256) softmax_fwd_kernel(float* __restrict__ S, const float* __restrict__ mask, int L, int rows_per_b) {
  __shared__ float sh[8];
  float* row = S + static_cast<long>(blockIdx.x) * L;
  const float* mk = mask + static_cast<long>(blockIdx.x / rows_per_b) * L;
  float mx = -3.0e38f;
  for (int j = threadIdx.x; j < L; j += 256) mx = fmaxf(mx, row[j] + mk[j]);
  mx = block_max256(mx, sh);
  float s = 0.f;
  for (int j = threadIdx.x; j < L; j += 256) { const float e = expf(row[j] + mk[j] - mx); row[j] = e; s += e; }
  const float inv = 1.f / block_sum256(s, sh);
  for (int j = threadIdx.x; j < L; j += 256) row[j] *= inv;
}
// dS = P * (dP - sum_j P dP), in place over dP
__global__ void __launch_bounds__(256) softmax_bwd_kernel(const float* __restrict__ P, float* __restrict__ dP, int L) {
  __shared__ float sh[8];
  const float* p = P + static_cast<long>(blockIdx.x) * L;
  float* d = dP + static_cast<long>(blockIdx.x) * L;
  float s = 0.f;
  for (int j = threadIdx.x; j < L; j += 256) s += p[j] * d[j];
  const float dot = block_sum256(s, sh);
  for (int j = threadIdx.x; j < L; j += 256) d[j] = p[j] * (d[j] - dot);
}

// mode 0 SiLU, 1 Mish (time MLP: decoder.py:105-117, 49)
__global__ void unary_fwd_kernel(const float* __restrict__ in, float* __restrict__ out, long n, int mode) {
  const long i = blockIdx.x * 256L + threadIdx.x;
  if (i >= n) return;
  const float x = in[i];
  out[i] = mode == 0 ? x / (1.f + expf(-x)) : mish_t(x);
}
__global__ void unary_bwd_kernel(float* __restrict__ d, const float* __restrict__ in, long n, int mode) {
  const long i = blockIdx.x * 256L + threadIdx.x;
  if (i >= n) return;
  const float x = in[i];
  float g;
  if (mode == 0) { const float sg = 1.f / (1.f + expf(-x)); g = sg * (1.f + x * (1.f - sg)); }
  else g = mish_grad_t(x);
  d[i] *= g;
}

// x0[b][t][c] = cat(y, mu, spks broadcast)[b][c][t] * mask[b][t]   (decoder.py:384-388 + the block's x*mask)
__global__ void pack_x0_kernel(const float* __restrict__ y, const float* __restrict__ mu, const float* __restrict__ spk,
                               const float* __restrict__ mask, float* __restrict__ x0, int cin, int T, long total) {
  const long i = blockIdx.x * 256L + threadIdx.x;
  if (i >= total) return;
  const int c = static_cast<int>(i % cin);
  const long row = i / cin;
  const int t = static_cast<int>(row % T);
  const long b = row / T;
  float v;
  if (c < 80) v = y[(b * 80 + c) * T + t];
  else if (c < 160) v = mu[(b * 80 + c - 80) * T + t];
  else v = spk[b * (cin - 160) + c - 160];
  x0[i] = v * mask[row];
}
// dmu[b][c][t] = dx0[b][t][80+c] * mask ; dspk[b][k] = sum_t dx0[b][t][160+k] * mask
__global__ void unpack_dx0_kernel(const float* __restrict__ dx0, const float* __restrict__ mask, float* __restrict__ dmu,
                                  float* __restrict__ dspk, int cin, int T) {
  const int b = blockIdx.x;
  if (dmu) {
    for (int i = threadIdx.x; i < 80 * T; i += 256) {
      const int c = i / T, t = i % T;
      dmu[(static_cast<long>(b) * 80 + c) * T + t] = dx0[(static_cast<long>(b) * T + t) * cin + 80 + c] * mask[static_cast<long>(b) * T + t];
    }
  }
  if (dspk && threadIdx.x < cin - 160) {
    float s = 0.f;
    for (int t = 0; t < T; ++t) s += dx0[(static_cast<long>(b) * T + t) * cin + 160 + threadIdx.x] * mask[static_cast<long>(b) * T + t];
    dspk[static_cast<long>(b) * (cin - 160) + threadIdx.x] = s;
  }
}
// y = (1 - (1 - sigma) t_b) z + t_b x1   (flow_matching.py:107), (B,80,T) layout
__global__ void make_y_kernel(const float* __restrict__ x1, const float* __restrict__ z, const float* __restrict__ t,
                              float sigma, float* __restrict__ y, long per_b, long total) {
  const long i = blockIdx.x * 256L + threadIdx.x;
  if (i >= total) return;
  const float tb = t[i / per_b];
  y[i] = (1.f - (1.f - sigma) * tb) * z[i] + tb * x1[i];
}
// loss += sum (v - u)^2 / den ; dv = 2 (v - u) / den * mask, u = x1 - (1 - sigma) z   (flow_matching.py:108,115-117)
// v, dv are [B][T][80]; x1, z are (B,80,T); den_ptr[0] = sum(mask) (multiplied by 80 here)
__global__ void __launch_bounds__(256) loss_kernel(const float* __restrict__ v, const float* __restrict__ x1,
                                                   const float* __restrict__ z, const float* __restrict__ mask,
                                                   const float* __restrict__ den_ptr, float sigma, int T, long total,
                                                   float* __restrict__ dv, float* __restrict__ loss) {
  __shared__ float sh[8];
  const float inv_den = 1.f / (den_ptr[0] * 80.f);
  float s = 0.f;
  for (long i = blockIdx.x * 256L + threadIdx.x; i < total; i += 256L * gridDim.x) {
    const int c = static_cast<int>(i % 80);
    const long row = i / 80;
    const int t = static_cast<int>(row % T);
    const long b = row / T;
    const long j = (b * 80 + c) * T + t;
    const float u = x1[j] - (1.f - sigma) * z[j];
    const float d = v[i] - u;
    s += d * d;
    dv[i] = 2.f * d * inv_den * mask[row];
  }
  s = block_sum256(s, sh);
  if (threadIdx.x == 0) atomicAdd(loss, s * inv_den);
}

}  // namespace trn
}  // namespace mb
