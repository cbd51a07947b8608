// Micro-benchmark: issue rate of tcgen05.mma (kind::f16, M=128, cta_group::1, both operands in shared memory) for the
// tile shapes and row-shifted A descriptors the conv kernel uses.  No TMA, no epilogue: shared memory holds zeros.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -I matcha_b200/csrc tools/umma_rate.cu -o gpurun_out/umma_rate
#include <cstdio>
#include <cstdlib>
#include "ptx.cuh"
using namespace mb;

template <int N, int ROWB>
__global__ void __launch_bounds__(128, 1) rate_kernel(int iters, int shift_rows, int b_shift_rows, int taps, long long* cycles) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  __shared__ uint64_t bar;
  __shared__ uint32_t tmem_slot;
  for (int i = threadIdx.x; i < (200 * 1024) / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(smem)[i] = 0;
  if (threadIdx.x == 0) { mbar_init(&bar, 1); fence_barrier_init(); }
  if (threadIdx.x < 32) tmem_alloc<512>(&tmem_slot);
  fence_proxy_async();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tm = tmem_slot;
  if (threadIdx.x == 0) {
    constexpr uint32_t idesc = umma_idesc_f16(0, 128, N);
    const uint32_t hi = static_cast<uint32_t>(umma_smem_desc<ROWB>(0) >> 32);
    const uint32_t a0 = static_cast<uint32_t>(umma_smem_desc<ROWB>(smem_u32(smem)));
    const uint32_t b0 = static_cast<uint32_t>(umma_smem_desc<ROWB>(smem_u32(smem + 96 * 1024)));
    const long long t0 = clock64();
    for (int it = 0; it < iters; ++it) {
      for (int tp = 0; tp < taps; ++tp) {
        const uint32_t a = a0 + ((tp * shift_rows * ROWB) >> 4);
        const uint32_t b = b0 + ((tp * b_shift_rows * ROWB) >> 4) + (b_shift_rows ? 0 : (tp % 3) * ((N * ROWB) >> 4));
#pragma unroll
        for (int k = 0; k < ROWB / 32; ++k) umma_f16_lohi(tm, a + 2 * k, b + 2 * k, hi, idesc, 1u);
      }
    }
    umma_commit(&bar);
    mbar_wait(&bar, 0);
    cycles[blockIdx.x] = clock64() - t0;
  }
  tc_fence_before();
  __syncthreads();
  if (threadIdx.x < 32) tmem_dealloc<512>(tm);
}

template <int N, int ROWB>
void run(const char* name, int shift, int bshift, int taps) {
  const int iters = 400;
  long long* d;
  cudaMalloc(&d, 148 * sizeof(long long));
  auto k = rate_kernel<N, ROWB>;
  cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, 210 * 1024);
  for (int rep = 0; rep < 2; ++rep) k<<<148, 128, 210 * 1024>>>(iters, shift, bshift, taps, d);
  cudaError_t err = cudaDeviceSynchronize();
  long long h[148];
  cudaMemcpy(h, d, sizeof(h), cudaMemcpyDeviceToHost);
  long long mx = 0;
  for (int i = 0; i < 148; ++i) mx = h[i] > mx ? h[i] : mx;
  const double n_mma = double(iters) * taps * (ROWB / 32);
  printf("%-28s N=%3d rowB=%3d a_shift=%d b_shift=%d taps=%2d : %.1f cyc/MMA (floor %d)  %s\n", name, N, ROWB, shift, bshift, taps,
         mx / n_mma, 128 * N / 256, err == cudaSuccess ? "" : cudaGetErrorString(err));
  cudaFree(d);
}

int main() {
  run<256, 128>("N256 aligned", 0, 0, 11);
  run<256, 128>("N256 A shift 1", 1, 0, 11);
  run<256, 128>("N256 A shift 5", 5, 0, 11);
  run<256, 128>("N256 A shift 8", 8, 0, 11);
  run<128, 128>("N128 aligned", 0, 0, 11);
  run<128, 128>("N128 A shift 1", 1, 0, 11);
  run<128, 128>("N128 A shift 3", 3, 0, 11);
  run<128, 128>("N128 A shift 5", 5, 0, 11);
  run<128, 128>("N128 A shift 8", 8, 0, 11);
  run<64, 128>("N64 aligned", 0, 0, 11);
  run<64, 128>("N64 A shift 5", 5, 0, 11);
  run<32, 64>("N32 rowB64 aligned", 0, 0, 11);
  run<32, 64>("N32 rowB64 A shift 5", 5, 0, 11);
  run<32, 128>("N32 rowB128 aligned", 0, 0, 11);
  // swapped roles: weights are the M=128 operand, activations (row-shifted) the N operand
  run<256, 128>("swap N256 B shift 1", 0, 1, 11);
  run<256, 128>("swap N256 B shift 5", 0, 5, 11);
  run<128, 128>("swap N128 B shift 5", 0, 5, 11);
  return 0;
}
