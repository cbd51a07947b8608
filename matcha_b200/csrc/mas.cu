// Monotonic alignment search on the GPU: the most likely monotonic path through the (tokens x frames)
// log-prior matrix, one CTA per utterance.  Replaces the reference's only native component, the Cython/OpenMP
// `maximum_path_c` (matcha/utils/monotonic_align/core.pyx:11-47) and its device -> host -> device round trip
// (matcha/utils/monotonic_align/__init__.py:7-22, called from matcha_tts.py:192 every training step).
//
// Reference recurrence, per utterance with t_x tokens and t_y frames (value = log_prior * mask, updated in place):
//     for y in 0..t_y-1, x in [max(0, t_x + y - t_y), min(t_x, y + 1)):
//         V[x, y] = max(x == y ? -1e9 : V[x, y-1],  x == 0 ? (y == 0 ? 0 : -1e9) : V[x-1, y-1]) + value[x, y]
//     index = t_x - 1; for y = t_y-1 .. 0:  path[index, y] = 1;
//         if index != 0 and (index == y or V[index, y-1] < V[index-1, y-1]): index -= 1
// Column y depends on column y-1 only, so the tokens of a column are computed in parallel (one __syncthreads per frame)
// from the previous column kept in shared memory; the input arrives in tiles of TC frames x all tokens with coalesced
// row-segment loads; the back-tracking comparison `V[x, y-1] < V[x-1, y-1]` is evaluated while column y is computed (both
// operands are the previous column) and kept as one bit per (frame, token) - 1/32 of the DP matrix - which a single thread
// then walks backwards.  Same fp32 operations in the same order per cell as the reference: the path is bit-identical.
// Defined, as the reference is, for 1 <= t_x <= t_y (with more tokens than frames the reference reads out of bounds).
#include <cuda_runtime.h>
#include <stdint.h>

#include <cstdlib>

#include "plan.h"

namespace mb {

namespace {

constexpr float kMasNeg = -1e9f;

__global__ void __launch_bounds__(1024)
mas_kernel(const float* __restrict__ value, const float* __restrict__ mask, int Tx, int Ty, int TC, float* __restrict__ path,
           uint32_t* __restrict__ dec) {
  extern __shared__ float sm[];
  float* colA = sm;                    // [Tx] previous column of V
  float* colB = colA + Tx;             // [Tx] current column
  float* tile = colB + Tx;             // [Tx][TC + 1] masked input values of TC frames
  __shared__ int s_cnt[2];
  const int b = blockIdx.x, tid = threadIdx.x, lane = tid & 31, nt = blockDim.x;
  const long base = static_cast<long>(b) * Tx * Ty;
  const int W = (Tx + 31) >> 5;

  // t_x = mask.sum(1)[:, 0], t_y = mask.sum(2)[:, 0]   (monotonic_align/__init__.py:19-20)
  if (tid < 2) s_cnt[tid] = 0;
  __syncthreads();
  int cx = 0, cy = 0;
  for (int x = tid; x < Tx; x += nt) cx += mask[base + static_cast<long>(x) * Ty] != 0.f;
  for (int y = tid; y < Ty; y += nt) cy += mask[base + y] != 0.f;
  atomicAdd(&s_cnt[0], cx);
  atomicAdd(&s_cnt[1], cy);
  // path = 0 everywhere; the back-tracking thread sets the ones afterwards
  for (long i = tid; i < static_cast<long>(Tx) * Ty; i += nt) path[base + i] = 0.f;
  for (int x = tid; x < Tx; x += nt) colA[x] = 0.f;
  __syncthreads();
  const int t_x = s_cnt[0], t_y = s_cnt[1];
  if (t_x <= 0 || t_y <= 0) return;

  float* prev = colA;
  float* cur = colB;
  const int x_round = ((t_x + 31) >> 5) << 5;
  for (int y0 = 0; y0 < t_y; y0 += TC) {
    for (int idx = tid; idx < t_x * TC; idx += nt) {
      const int x = idx / TC, c = idx - x * TC, y = y0 + c;
      float v = 0.f;
      if (y < t_y) v = value[base + static_cast<long>(x) * Ty + y] * mask[base + static_cast<long>(x) * Ty + y];
      tile[x * (TC + 1) + c] = v;
    }
    __syncthreads();
    const int nc = min(TC, t_y - y0);
    for (int c = 0; c < nc; ++c) {
      const int y = y0 + c;
      const int lo = max(0, t_x + y - t_y), hi = min(t_x, y + 1);
      for (int x = tid; x < x_round; x += nt) {
        bool step = false;
        if (x < t_x) {
          const float in = tile[x * (TC + 1) + c];
          float nv = in;                               // cells outside the band keep their input value
          if (x >= lo && x < hi) {
            const float v_cur = (x == y) ? kMasNeg : prev[x];
            const float v_prev = (x == 0) ? (y == 0 ? 0.f : kMasNeg) : prev[x - 1];
            nv = fmaxf(v_cur, v_prev) + in;
          }
          cur[x] = nv;
          if (x != 0 && y != 0) step = (x == y) || (prev[x] < prev[x - 1]);
        }
        const uint32_t bits = __ballot_sync(0xffffffffu, step);
        if (lane == 0) dec[(static_cast<long>(b) * Ty + y) * W + (x >> 5)] = bits;
      }
      __syncthreads();
      float* t = prev; prev = cur; cur = t;
    }
  }
  if (tid == 0) {
    int index = t_x - 1;
    for (int y = t_y - 1; y >= 0; --y) {
      path[base + static_cast<long>(index) * Ty + y] = 1.f;
      if (index != 0) {
        const uint32_t w = dec[(static_cast<long>(b) * Ty + y) * W + (index >> 5)];
        if ((w >> (index & 31)) & 1u) --index;
      }
    }
  }
}

}  // namespace

size_t maximum_path_workspace_bytes(int B, int Tx, int Ty) {
  return static_cast<size_t>(B) * Ty * ((Tx + 31) / 32) * sizeof(uint32_t);
}

void maximum_path(const float* value, const float* mask, int B, int Tx, int Ty, float* path, void* workspace, cudaStream_t st) {
  if (B <= 0 || Tx <= 0 || Ty <= 0) throw Error("maximum_path: empty shape");
  int TC = 32;
  auto bytes = [&](int tc) { return (2 * static_cast<size_t>(Tx) + static_cast<size_t>(Tx) * (tc + 1)) * sizeof(float); };
  while (TC > 1 && bytes(TC) > 200 * 1024) TC >>= 1;
  if (bytes(TC) > 200 * 1024) throw Error("maximum_path: too many tokens for one CTA's shared memory");
  static PerDeviceOnce configured;
  if (configured.needed()) {
    MB_CUDA(cudaFuncSetAttribute(mas_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
    configured.mark();
  }
  // 1024 threads: one CTA per utterance means few resident warps, and both the tile loads and the per-column pass are
  // latency-bound (measured 256 / 512 / 1024 threads: 0.34 / 0.29 / 0.26 ms at 170 x 512, 5.2 / 3.4 / 2.4 ms at 667 x 2001)
  int threads = 1024;
  if (const char* e = std::getenv("MB200_MAS_THREADS")) threads = std::atoi(e);   // developer switch (256 / 512 / 1024)
  if (threads != 256 && threads != 512 && threads != 1024) throw Error("MB200_MAS_THREADS must be 256, 512 or 1024");
  mas_kernel<<<B, threads, bytes(TC), st>>>(value, mask, Tx, Ty, TC, path, static_cast<uint32_t*>(workspace));
  MB_CUDA(cudaGetLastError());
}

// ---------------------------------------------------------------------------------------------------------------------
// Allocation-free variant (mb200_maximum_path_static).  No counterpart in the reference: added
// for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses
// that forbid it).  The kernel's decision-bit scratch lives in a file-scope __device__ array instead of a caller
// allocation; every GPU that loads the library holds these bytes until the process exits.
//
// In host code the array's NAME is a host-side stand-in, not a device address: its device address is taken once per
// device, with that device current, by cudaGetSymbolAddress below, and that pointer is what the kernel receives.
// The array is shared by every call on a device and keeps the last call's contents.  Nothing has to be cleared: each
// decision word the back-tracking walk reads (frame y, token word index >> 5, index < t_x) is written earlier in the
// same launch by the forward pass.  Calls that overlap on one GPU (different streams, or threads) race on it.
constexpr size_t kMasStaticBytes = MB200_MAS_STATIC_SCRATCH_BYTES;
static __device__ __align__(256) unsigned char g_mas_scratch[kMasStaticBytes];

void maximum_path_static(const float* value, const float* mask, int B, int Tx, int Ty, float* path, cudaStream_t st) {
  if (B <= 0 || Tx <= 0 || Ty <= 0) throw Error("maximum_path_static: empty shape");
  const size_t need = maximum_path_workspace_bytes(B, Tx, Ty);
  if (need > kMasStaticBytes)
    throw Error("maximum_path_static: B * Ty * ceil(Tx / 32) * 4 = " + std::to_string(need) + " bytes of scratch exceed the static limit of " +
                std::to_string(kMasStaticBytes) + " bytes (MB200_MAS_STATIC_SCRATCH_BYTES); use mb200_maximum_path with a caller workspace");
  constexpr int kMaxDev = 64;
  static void* scratch_addr[kMaxDev] = {};
  static std::mutex mu;
  int dev = 0;
  MB_CUDA(cudaGetDevice(&dev));
  if (dev < 0 || dev >= kMaxDev) throw Error("maximum_path_static: device ordinal out of range");
  void* ptr = nullptr;
  {
    std::lock_guard<std::mutex> lk(mu);
    if (!scratch_addr[dev]) {   // set-up, once per device, with that device current
      void* p = nullptr;
      MB_CUDA(cudaGetSymbolAddress(&p, g_mas_scratch));
      if (!p) throw Error("maximum_path_static: cudaGetSymbolAddress returned a null address");
      scratch_addr[dev] = p;
    }
    ptr = scratch_addr[dev];
  }
  maximum_path(value, mask, B, Tx, Ty, path, ptr, st);   // same kernel, same launch; no cudaMalloc / cudaFree anywhere
}

}  // namespace mb
