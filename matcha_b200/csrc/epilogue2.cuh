// Coalesced epilogue of the persistent tcgen05 kernel (half-precision modes).
//
// tcgen05.ld hands every thread one output ROW (32 lanes = 32 rows).  Writing rows straight to global memory makes each
// store instruction touch 32 different 128-byte lines; the first kernel was bound by exactly that (L1 wavefronts), not
// by the tensor pipe or HBM.  Here every warp owns a small shared-memory staging tile (32 rows x 128 B, pitch 144 B):
//   inputs  (GroupNorm raw, residuals): coalesced global -> staging (a quarter-warp reads one full 128 B row), then each
//           thread picks up its own row conflict-free;
//   outputs (main, LeakyReLU copy, LayerNorm): thread rows -> staging -> coalesced global stores of full lines.
// LayerNorm keeps the finished fp32 row in TMEM (tcgen05.st) instead of recomputing the epilogue in a second pass.
// Transcendentals use the fast SFU intrinsics: the results are rounded to f16/bf16 anyway.
#pragma once
#include "conv.cuh"
#include "ptx.cuh"

namespace mb {

constexpr int kStagePitch = 144;                 // bytes; 36 words: rows of a quarter-warp land in disjoint banks
constexpr int kStageBytes = 32 * kStagePitch;    // per epilogue warp

__device__ __forceinline__ void tmem_st_32x32(uint32_t taddr, const float* v) {
  const uint32_t* r = reinterpret_cast<const uint32_t*>(v);
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x32.b32 [%0], "
      "{%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16, "
      "%17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31, %32};" ::"r"(taddr),
      "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]), "r"(r[8]), "r"(r[9]),
      "r"(r[10]), "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15]), "r"(r[16]), "r"(r[17]), "r"(r[18]),
      "r"(r[19]), "r"(r[20]), "r"(r[21]), "r"(r[22]), "r"(r[23]), "r"(r[24]), "r"(r[25]), "r"(r[26]), "r"(r[27]),
      "r"(r[28]), "r"(r[29]), "r"(r[30]), "r"(r[31])
      : "memory");
}
__device__ __forceinline__ void tmem_st_wait() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }

__device__ __forceinline__ float mish_fast(float x) {
  if (x > 20.f) return x;
  const float n = __expf(x);
  const float w = n * (n + 2.f);
  return x * __fdividef(w, w + 2.f);
}

// Row geometry of one epilogue warp within its tile.
struct RowMap {
  long row_base;   // b * T_o + phase * o_phase_step
  int t0w;         // first GEMM row (time index) of this warp
  int o_mul;
  int T_out;
  __device__ __forceinline__ long grow(int r) const { return row_base + static_cast<long>(t0w + r) * o_mul; }
  __device__ __forceinline__ bool ok(int r) const { return t0w + r < T_out; }
};

// CP = columns per pass (64: rows of 128 B, 8 lanes per row; 32: rows of 64 B, 4 lanes per row).
template <int CP, typename T>
__device__ __forceinline__ void stage_load(uint8_t* stg, const T* src, int ld, int n0, int n_valid, const RowMap& rm,
                                           int lane) {
  constexpr int PPR = CP / 8;          // 16-byte pieces per row
  constexpr int RPI = 32 / PPR;        // rows per instruction
  const int piece = lane % PPR, rsub = lane / PPR;
#pragma unroll
  for (int i = 0; i < PPR; ++i) {
    const int r = rsub + RPI * i;
    uint4 u = make_uint4(0, 0, 0, 0);
    if (rm.ok(r) && n0 + piece * 8 < n_valid) u = __ldg(reinterpret_cast<const uint4*>(src + rm.grow(r) * ld + n0 + piece * 8));
    *reinterpret_cast<uint4*>(stg + r * kStagePitch + piece * 16) = u;
  }
}
template <int CP, typename T>
__device__ __forceinline__ void stage_store(const uint8_t* stg, T* dst, int ld, int n0, int n_valid, const RowMap& rm,
                                            int lane) {
  constexpr int PPR = CP / 8;
  constexpr int RPI = 32 / PPR;
  const int piece = lane % PPR, rsub = lane / PPR;
#pragma unroll
  for (int i = 0; i < PPR; ++i) {
    const int r = rsub + RPI * i;
    if (rm.ok(r) && n0 + piece * 8 < n_valid)
      *reinterpret_cast<uint4*>(dst + rm.grow(r) * ld + n0 + piece * 8) =
          *reinterpret_cast<const uint4*>(stg + r * kStagePitch + piece * 16);
  }
}
// own-row accessors (lane == row)
template <typename T>
__device__ __forceinline__ void row_get8(const uint8_t* stg, int lane, int j, float* f) {
  Elem<T>::ld8(reinterpret_cast<const T*>(stg + lane * kStagePitch + j * 16), f);
}
template <typename T>
__device__ __forceinline__ void row_put8(uint8_t* stg, int lane, int j, const float* f) {
  Elem<T>::st8(reinterpret_cast<T*>(stg + lane * kStagePitch + j * 16), f);
}

// True when the coalesced path covers this epilogue (everything in the operand dtype).
__device__ __forceinline__ bool epi2_supported(const EpiParams& e) {
  return !e.out0_f32 && !e.resid_f32 && !e.resid2_f32 && e.x_state == nullptr;
}

template <int NTILE, typename T>
__device__ __forceinline__ void epilogue_tile2(const EpiParams& e, int b, int t0w, int T_out, int phase, int nbase,
                                               uint32_t taddr, uint8_t* stg, int lane) {
  constexpr int CP = NTILE >= 64 ? 64 : 32;
  constexpr int NJ = CP / 8;
  RowMap rm{static_cast<long>(b) * e.T_o + phase * e.o_phase_step, t0w, e.o_mul, T_out};
  const bool valid = rm.ok(lane);
  const long myrow = rm.grow(lane);
  const float m = (e.mask && valid) ? __ldg(e.mask + myrow) : 1.f;
  float ln_s = 0.f, ln_ss = 0.f;
  const bool do_ln = (e.ln_out != nullptr) && NTILE == 256;

#pragma unroll 1
  for (int c0 = 0; c0 < NTILE; c0 += CP) {
    const int n0 = nbase + c0;
    float v[CP];
    // the first input stream goes to staging while the accumulator load is in flight
    const void* first_in = e.gn_raw ? e.gn_raw : (e.resid ? e.resid : e.resid2);
    if (first_in) stage_load<CP, T>(stg, reinterpret_cast<const T*>(first_in), e.gn_raw ? e.gn_ld : (e.resid ? e.resid_ld : e.resid2_ld), n0, e.n_valid, rm, lane);
    tmem_ld_32x32(taddr + c0, v);
    if constexpr (CP == 64) tmem_ld_32x32(taddr + c0 + 32, v + (CP == 64 ? 32 : 0));
    tmem_ld_wait();
    if (e.bias) {
#pragma unroll
      for (int i = 0; i < CP; ++i) v[i] += __ldg(e.bias + n0 + i);
    }
    if (e.gn_raw) {
      __syncwarp();
#pragma unroll
      for (int h = 0; h < CP / 32; ++h) {   // one GroupNorm group per 32 channels
        const float* st = e.gn_stats + (static_cast<long>(b) * 8 + ((n0 + 32 * h) >> 5)) * 2;
        const float mean = st[0] * e.gn_inv_count;
        const float rstd = rsqrtf(fmaxf(st[1] * e.gn_inv_count - mean * mean, 0.f) + 1e-5f);
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          float r[8];
          row_get8<T>(stg, lane, 4 * h + j, r);
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            const int n = n0 + 32 * h + 8 * j + i;
            const float y = (r[i] - mean) * rstd * __ldg(e.gn_gamma + n) + __ldg(e.gn_beta + n);
            v[32 * h + 8 * j + i] += mish_fast(y) * m;
          }
        }
      }
      __syncwarp();
      if (e.resid) stage_load<CP, T>(stg, reinterpret_cast<const T*>(e.resid), e.resid_ld, n0, e.n_valid, rm, lane);
    }
    if (e.resid) {
      __syncwarp();
#pragma unroll
      for (int j = 0; j < NJ; ++j) {
        float r[8];
        row_get8<T>(stg, lane, j, r);
#pragma unroll
        for (int i = 0; i < 8; ++i) v[8 * j + i] += r[i];
      }
      __syncwarp();
      if (e.resid2) stage_load<CP, T>(stg, reinterpret_cast<const T*>(e.resid2), e.resid2_ld, n0, e.n_valid, rm, lane);
    } else if (e.gn_raw && e.resid2) {
      stage_load<CP, T>(stg, reinterpret_cast<const T*>(e.resid2), e.resid2_ld, n0, e.n_valid, rm, lane);
    }
    if (e.resid2) {
      __syncwarp();
#pragma unroll
      for (int j = 0; j < NJ; ++j) {
        float r[8];
        row_get8<T>(stg, lane, j, r);
#pragma unroll
        for (int i = 0; i < 8; ++i) v[8 * j + i] += r[i];
      }
      __syncwarp();
    }
    if (e.scale != 1.f) {
#pragma unroll
      for (int i = 0; i < CP; ++i) v[i] *= e.scale;
    }
    if (e.snake_a) {
#pragma unroll
      for (int i = 0; i < CP; ++i) {
        const float s = __sinf(v[i] * __ldg(e.snake_a + n0 + i));
        v[i] += __ldg(e.snake_invb + n0 + i) * s * s;
      }
    }
    if (e.mask_out) {
#pragma unroll
      for (int i = 0; i < CP; ++i) v[i] *= m;
    }
    if (e.stats_out) {
#pragma unroll
      for (int h = 0; h < CP / 32; ++h) {
        float s = 0.f, ss = 0.f;
        if (valid) {
#pragma unroll
          for (int i = 0; i < 32; ++i) { s += v[32 * h + i]; ss += v[32 * h + i] * v[32 * h + i]; }
        }
        s = warp_sum(s);
        ss = warp_sum(ss);
        if (lane == 0) {
          float* st = e.stats_out + (static_cast<long>(b) * 8 + ((n0 + 32 * h) >> 5)) * 2;
          atomicAdd(st, s);
          atomicAdd(st + 1, ss);
        }
      }
    }
    if (do_ln) {
#pragma unroll
      for (int i = 0; i < CP; ++i) { ln_s += v[i]; ln_ss += v[i] * v[i]; }
      tmem_st_32x32(taddr + c0, v);
      if constexpr (CP == 64) tmem_st_32x32(taddr + c0 + 32, v + (CP == 64 ? 32 : 0));
    }
    if (e.out0) {
#pragma unroll
      for (int j = 0; j < NJ; ++j) {
        float o[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) o[i] = e.out0_slope == 1.f ? v[8 * j + i] : lrelu_f(v[8 * j + i], e.out0_slope);
        row_put8<T>(stg, lane, j, o);
      }
      __syncwarp();
      stage_store<CP, T>(stg, reinterpret_cast<T*>(e.out0), e.out0_ld, n0, e.n_valid, rm, lane);
      __syncwarp();
    }
    if (e.out1) {
#pragma unroll
      for (int j = 0; j < NJ; ++j) {
        float o[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) o[i] = e.out1_slope == 1.f ? v[8 * j + i] : lrelu_f(v[8 * j + i], e.out1_slope);
        row_put8<T>(stg, lane, j, o);
      }
      __syncwarp();
      stage_store<CP, T>(stg, reinterpret_cast<T*>(e.out1), e.out1_ld, n0, e.n_valid, rm, lane);
      __syncwarp();
    }
  }
  if (do_ln) {
    tmem_st_wait();
    const float mean = ln_s * (1.f / 256.f);
    const float rstd = rsqrtf(fmaxf(ln_ss * (1.f / 256.f) - mean * mean, 0.f) + 1e-5f);
#pragma unroll 1
    for (int c0 = 0; c0 < NTILE; c0 += CP) {
      float v[CP];
      tmem_ld_32x32(taddr + c0, v);
      if constexpr (CP == 64) tmem_ld_32x32(taddr + c0 + 32, v + (CP == 64 ? 32 : 0));
      tmem_ld_wait();
#pragma unroll
      for (int j = 0; j < NJ; ++j) {
        float o[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          const int n = c0 + 8 * j + i;
          o[i] = (v[8 * j + i] - mean) * rstd * __ldg(e.ln_g + n) + __ldg(e.ln_b + n);
        }
        row_put8<T>(stg, lane, j, o);
      }
      __syncwarp();
      stage_store<CP, T>(stg, reinterpret_cast<T*>(e.ln_out), e.ln_ld, c0, 256, rm, lane);
      __syncwarp();
    }
  }
}

}  // namespace mb
