// Micro-benchmark: instruction-issue cost per tcgen05.mma for different ways of electing the issuing thread, and the
// cost of one producer/consumer ring step when the whole warp takes part in the waits.  N = 32 so that the tensor-pipe
// floor (16 cycles) does not hide the issue cost.
#include <cstdio>
#include "ptx.cuh"
using namespace mb;

__device__ __forceinline__ bool elect_sync_pred() {
  uint32_t pred = 0;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "elect.sync _|p, 0xffffffff;\n\t"
      "@p mov.s32 %0, 1;\n\t}\n"
      : "+r"(pred));
  return pred != 0;
}

template <int MODE>
__global__ void __launch_bounds__(64, 1) issue_kernel(int iters, int taps, int n_wait, long long* cycles) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  __shared__ uint64_t bar, full[8], empty[8];
  __shared__ uint32_t tmem_slot;
  for (int i = threadIdx.x; i < (64 * 1024) / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(smem)[i] = 0;
  if (threadIdx.x == 0) {
    mbar_init(&bar, 1);
    for (int s = 0; s < 8; ++s) { mbar_init(&full[s], 1); mbar_init(&empty[s], 1); }
    fence_barrier_init();
  }
  if (threadIdx.x < 32) tmem_alloc<512>(&tmem_slot);
  fence_proxy_async();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tm = tmem_slot;
  constexpr uint32_t idesc = umma_idesc_f16(0, 128, 32);
  const uint32_t hi = static_cast<uint32_t>(umma_smem_desc<128>(0) >> 32);
  const uint32_t a0 = static_cast<uint32_t>(umma_smem_desc<128>(smem_u32(smem)));
  const uint32_t b0 = static_cast<uint32_t>(umma_smem_desc<128>(smem_u32(smem + 32 * 1024)));
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (warp == 0) {
    // producer for the ring variant: whole warp waits, one elected lane arrives
    if (n_wait) {
      int s = 0; uint32_t ph = 0;
      for (int i = 0; i < iters * taps; ++i) {
        mbar_wait(&empty[s], ph ^ 1);
        if (elect_sync_pred()) mbar_arrive(&full[s]);
        __syncwarp();
        if (++s == 4) { s = 0; ph ^= 1; }
      }
    }
  } else {
    long long t0 = 0;
    if (MODE == 0) {                 // one thread runs the whole loop
      if (lane == 0) {
        t0 = clock64();
        int s = 0; uint32_t ph = 0;
        for (int it = 0; it < iters; ++it)
          for (int tp = 0; tp < taps; ++tp) {
            if (n_wait) mbar_wait(&full[s], ph);
            const uint32_t a = a0 + tp * 40, b = b0 + tp * 256;
#pragma unroll
            for (int k = 0; k < 4; ++k) umma_f16_lohi(tm, a + 2 * k, b + 2 * k, hi, idesc, 1u);
            if (n_wait) { umma_commit(&empty[s]); if (++s == 4) { s = 0; ph ^= 1; } }
          }
        umma_commit(&bar);
      }
    } else if (MODE == 1) {          // converged warp, elect.sync around every MMA group
      t0 = clock64();
      int s = 0; uint32_t ph = 0;
      for (int it = 0; it < iters; ++it)
        for (int tp = 0; tp < taps; ++tp) {
          if (n_wait) mbar_wait(&full[s], ph);
          const uint32_t a = a0 + tp * 40, b = b0 + tp * 256;
          if (elect_sync_pred()) {
#pragma unroll
            for (int k = 0; k < 4; ++k) umma_f16_lohi(tm, a + 2 * k, b + 2 * k, hi, idesc, 1u);
            if (n_wait) umma_commit(&empty[s]);
          }
          __syncwarp();
          if (n_wait) { if (++s == 4) { s = 0; ph ^= 1; } }
        }
      if (elect_sync_pred()) umma_commit(&bar);
      __syncwarp();
    } else {                         // converged warp, leader elected once
      const bool leader = elect_sync_pred();
      t0 = clock64();
      int s = 0; uint32_t ph = 0;
      for (int it = 0; it < iters; ++it)
        for (int tp = 0; tp < taps; ++tp) {
          if (n_wait) mbar_wait(&full[s], ph);
          const uint32_t a = a0 + tp * 40, b = b0 + tp * 256;
          if (leader) {
#pragma unroll
            for (int k = 0; k < 4; ++k) umma_f16_lohi(tm, a + 2 * k, b + 2 * k, hi, idesc, 1u);
            if (n_wait) umma_commit(&empty[s]);
          }
          __syncwarp();
          if (n_wait) { if (++s == 4) { s = 0; ph ^= 1; } }
        }
      if (leader) umma_commit(&bar);
      __syncwarp();
    }
    mbar_wait(&bar, 0);
    if (lane == 0) cycles[blockIdx.x] = clock64() - t0;
  }
  tc_fence_before();
  __syncthreads();
  if (threadIdx.x < 32) tmem_dealloc<512>(tm);
}

template <int MODE>
void run(const char* name, int n_wait) {
  const int iters = 200, taps = 11;
  long long* d;
  cudaMalloc(&d, 148 * sizeof(long long));
  auto k = issue_kernel<MODE>;
  cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, 80 * 1024);
  for (int rep = 0; rep < 2; ++rep) k<<<148, 64, 80 * 1024>>>(iters, taps, n_wait, d);
  cudaError_t err = cudaDeviceSynchronize();
  long long h[148];
  cudaMemcpy(h, d, sizeof(h), cudaMemcpyDeviceToHost);
  long long mx = 0;
  for (int i = 0; i < 148; ++i) mx = h[i] > mx ? h[i] : mx;
  printf("%-44s ring=%d : %.1f cyc per 4-MMA step (tensor floor 64) %s\n", name, n_wait, double(mx) / (iters * taps),
         err == cudaSuccess ? "" : cudaGetErrorString(err));
  cudaFree(d);
}

int main() {
  for (int w = 0; w < 2; ++w) {
    run<0>("single thread (lane==0 branch)", w);
    run<1>("converged warp, elect.sync per step", w);
    run<2>("converged warp, leader elected once", w);
  }
  return 0;
}
