// Micro-benchmark: what one kernel boundary costs for a kernel shaped like the persistent conv-GEMM (148 CTAs x 448 threads,
// 220 KB of dynamic shared memory, ~4 KB of __grid_constant__ parameters, barrier init + 512-column TMEM allocation), with
// NO work inside.  Chains of dependent launches: plain stream launches, a captured CUDA graph, and both with programmatic
// dependent launch (griddepcontrol.launch_dependents at the top, griddepcontrol.wait before the first "dependent" access).
//
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -I matcha_b200/csrc tools/launch_overhead.cu -o /tmp/launch_overhead
#include <cstdio>
#include <cuda_runtime.h>
#include "ptx.cuh"
using namespace mb;

struct Params { int words[1000]; };   // 4000 bytes, like five tensor maps + geometry + epilogue parameters

template <int MODE>   // 0: empty kernel; 1: conv-kernel prologue / teardown (barriers, TMEM alloc / dealloc, two block syncs)
__global__ void __launch_bounds__(448, 1) k(const __grid_constant__ Params p, int pdl, int* sink) {
  extern __shared__ uint8_t smem[];
  __shared__ uint64_t bars[40];
  __shared__ uint32_t tmem_slot;
  if (pdl && threadIdx.x == 0) griddep_launch();
  if (MODE == 1) {
    if (threadIdx.x == 0) {
      for (int s = 0; s < 40; ++s) mbar_init(&bars[s], 1);
      fence_barrier_init();
    }
    if (threadIdx.x < 32) tmem_alloc<512>(&tmem_slot);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
  }
  if (pdl) griddep_wait();
  if (p.words[threadIdx.x & 511] == 0x7fffffff) sink[0] = smem[threadIdx.x];   // never true: keeps the parameters alive
  if (MODE == 1) {
    tc_fence_before();
    __syncthreads();
    if (threadIdx.x < 32) tmem_dealloc<512>(tmem_slot);
  }
}

template <int MODE>
static cudaError_t launch(cudaStream_t st, const Params& p, int pdl, int* sink) {
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3(148);
  cfg.blockDim = dim3(448);
  cfg.dynamicSmemBytes = 220 * 1024;
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = pdl ? 1 : 0;
  return cudaLaunchKernelEx(&cfg, k<MODE>, p, pdl, sink);
}

template <int MODE>
static void run(const char* name) {
  cudaFuncSetAttribute(k<MODE>, cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024);
  Params p{};
  int* sink;
  cudaMalloc(&sink, 4);
  cudaStream_t st;
  cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking);
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  const int n = 300;
  for (int pdl = 0; pdl < 2; ++pdl) {
    for (int i = 0; i < 20; ++i) launch<MODE>(st, p, pdl, sink);
    cudaStreamSynchronize(st);
    cudaEventRecord(e0, st);
    for (int i = 0; i < n; ++i) launch<MODE>(st, p, pdl, sink);
    cudaEventRecord(e1, st);
    cudaStreamSynchronize(st);
    float ms = 0.f;
    cudaEventElapsedTime(&ms, e0, e1);
    printf("%-28s stream launches, pdl=%d: %6.2f us per kernel\n", name, pdl, 1e3f * ms / n);
    cudaGraph_t graph;
    cudaGraphExec_t exec;
    cudaStreamBeginCapture(st, cudaStreamCaptureModeThreadLocal);
    for (int i = 0; i < n; ++i) launch<MODE>(st, p, pdl, sink);
    cudaStreamEndCapture(st, &graph);
    cudaError_t err = cudaGraphInstantiate(&exec, graph, 0);
    if (err != cudaSuccess) { printf("graph instantiate failed: %s\n", cudaGetErrorString(err)); continue; }
    for (int i = 0; i < 3; ++i) cudaGraphLaunch(exec, st);
    cudaStreamSynchronize(st);
    cudaEventRecord(e0, st);
    for (int i = 0; i < 10; ++i) cudaGraphLaunch(exec, st);
    cudaEventRecord(e1, st);
    cudaStreamSynchronize(st);
    cudaEventElapsedTime(&ms, e0, e1);
    printf("%-28s CUDA graph,      pdl=%d: %6.2f us per kernel\n", name, pdl, 1e3f * ms / (10 * n));
    cudaGraphExecDestroy(exec);
    cudaGraphDestroy(graph);
  }
  printf("last error: %s\n", cudaGetErrorString(cudaGetLastError()));
}

int main() {
  run<0>("empty kernel");
  run<1>("prologue + teardown only");
  return 0;
}
