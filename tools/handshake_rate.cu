// Micro-benchmark: cost per step of the producer <-> MMA-issuer ring handshake used by the conv kernel.
//   producer (warp 0, one thread): wait empty[s]; arrive full[s]        (optionally expect_tx + a real TMA-less arrive)
//   consumer (warp 1, one thread): wait full[s]; [n MMAs]; release empty[s] by tcgen05.commit or mbarrier.arrive
#include <cstdio>
#include "ptx.cuh"
using namespace mb;

__global__ void __launch_bounds__(64, 1) hs_kernel(int steps, int stages, int mode, int n_mma, int fence, long long* cycles) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  __shared__ uint64_t full[8], empty[8], done;
  __shared__ uint32_t tmem_slot;
  for (int i = threadIdx.x; i < (64 * 1024) / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(smem)[i] = 0;
  if (threadIdx.x == 0) {
    for (int s = 0; s < 8; ++s) { mbar_init(&full[s], 1); mbar_init(&empty[s], 1); }
    mbar_init(&done, 1);
    fence_barrier_init();
  }
  if (threadIdx.x < 32) tmem_alloc<512>(&tmem_slot);
  fence_proxy_async();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tm = tmem_slot;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (warp == 0 && lane == 0) {
    int s = 0; uint32_t ph = 0;
    for (int i = 0; i < steps; ++i) {
      mbar_wait(&empty[s], ph ^ 1);
      mbar_arrive(&full[s]);
      if (++s == stages) { s = 0; ph ^= 1; }
    }
  } else if (warp == 1 && lane == 0) {
    constexpr uint32_t idesc = umma_idesc_f16(0, 128, 128);
    const uint32_t hi = static_cast<uint32_t>(umma_smem_desc<128>(0) >> 32);
    const uint32_t a0 = static_cast<uint32_t>(umma_smem_desc<128>(smem_u32(smem)));
    const uint32_t b0 = static_cast<uint32_t>(umma_smem_desc<128>(smem_u32(smem + 32 * 1024)));
    int s = 0; uint32_t ph = 0;
    const long long t0 = clock64();
    for (int i = 0; i < steps; ++i) {
      mbar_wait(&full[s], ph);
      if (fence) tc_fence_after();
      for (int k = 0; k < n_mma; ++k) umma_f16_lohi(tm, a0 + 2 * (k & 3), b0 + 2 * (k & 3), hi, idesc, 1u);
      if (mode == 0) umma_commit(&empty[s]); else mbar_arrive(&empty[s]);
      if (++s == stages) { s = 0; ph ^= 1; }
    }
    umma_commit(&done);
    mbar_wait(&done, 0);
    cycles[blockIdx.x] = clock64() - t0;
  }
  tc_fence_before();
  __syncthreads();
  if (threadIdx.x < 32) tmem_dealloc<512>(tm);
}

void run(int stages, int mode, int n_mma, int fence) {
  const int steps = 2000;
  long long* d;
  cudaMalloc(&d, 148 * sizeof(long long));
  cudaFuncSetAttribute(hs_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 80 * 1024);
  for (int rep = 0; rep < 2; ++rep) hs_kernel<<<148, 64, 80 * 1024>>>(steps, stages, mode, n_mma, fence, d);
  cudaError_t err = cudaDeviceSynchronize();
  long long h[148];
  cudaMemcpy(h, d, sizeof(h), cudaMemcpyDeviceToHost);
  long long mx = 0;
  for (int i = 0; i < 148; ++i) mx = h[i] > mx ? h[i] : mx;
  printf("fence=%d stages=%d release=%s mma/step=%d : %.1f cyc/step (MMA floor %d) %s\n", fence, stages, mode == 0 ? "tcgen05.commit" : "mbarrier.arrive",
         n_mma, double(mx) / steps, 64 * n_mma, err == cudaSuccess ? "" : cudaGetErrorString(err));
  cudaFree(d);
}

int main() {
  for (int fence = 0; fence < 2; ++fence) {
    for (int mode = 0; mode < 2; ++mode)
      for (int stages : {2, 6}) run(stages, mode, 0, fence);
    for (int stages : {2, 6}) run(stages, 0, 4, fence);
    for (int stages : {2, 6}) run(stages, 0, 8, fence);
    for (int stages : {2, 6}) run(stages, 0, 2, fence);
  }
  return 0;
}
