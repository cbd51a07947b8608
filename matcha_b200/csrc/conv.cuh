// Shared definitions for the conv-as-GEMM kernels (tcgen05 and SIMT): geometry, fused epilogue.
//
// Activations are "NTC": [B][T][C] with C contiguous, so a conv tap is the same tile shifted along T and
// every Conv1d / ConvTranspose1d / Linear on the path becomes   D[t, n] = sum_{tap, c} A[t*? + off_tap, c] * W[n, tap, c].
// Reference ops covered (citations under /root/reference): Block1D conv (decoder.py:36), res_conv (:54),
// Downsample1D (:67), Upsample1D ConvTranspose1d (:144), final_proj (:313), attention/FF linears
// (transformer.py:48,126,196-204), HiFi-GAN conv_pre/ups/ResBlock1 convs (hifigan/models.py:154-177, 18-87).
#pragma once
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>

namespace mb {

constexpr int kMaxTaps = 11;
constexpr int kMaxPhases = 8;

struct ConvGeom {
  int B;
  int T_in;        // rows per batch in each source view
  int T_out;       // GEMM rows per batch
  int n_src;       // 1 or 2 (channel concat of two tensors: decoder.py:427)
  int src_ch[2];   // channels consumed from each source per tap (multiple of the K chunk)
  int src_ld[2];   // row pitch (elements) of each source view
  int n_taps, n_phases;
  int tap_row[kMaxPhases][kMaxTaps];  // source row = t + tap_row
  int tap_ch[kMaxPhases][kMaxTaps];   // channel base inside the source view (strided conv reads a [T/2][2C] view)
  int N;           // output channels per phase (padded; weights hold n_phases*N rows)
  int K_total;     // n_taps * (src_ch[0] + src_ch[1]) = weight row pitch
};

// Geometry of the persistent kernel (conv_gemm2.cu).  The K loop is organised in A-load GROUPS: one group = one source
// view and channel base; for each K chunk of a group the A tile is loaded ONCE with a halo (128 + 8*extra rows) and
// every tap of the group reuses it through a row-shifted shared-memory descriptor.
constexpr int kMaxGroups = 6;   // 2 sources x 2 channel bases, or 3 split-precision products x 2 bases
constexpr int kMaxSched = 512;
constexpr uint32_t kSchedNull = 1u << 27;    // padding entry: the issuer handles two entries per iteration
constexpr uint32_t kSchedWFirst = 1u << 28;  // first tap of a streamed weight stage: wait for it
constexpr uint32_t kSchedWLast = 1u << 29;   // last tap of a streamed weight stage: release it
constexpr uint32_t kSchedFirst = 1u << 30;   // first tap of an A stage: wait for the halo tile
constexpr uint32_t kSchedLast = 1u << 31;    // last tap of an A stage: release it
struct ConvGeom2 {
  int B, T_in, T_out, N, n_phases, K_total;
  int n_groups;
  int g_src[kMaxGroups], g_ch[kMaxGroups], g_chunks[kMaxGroups], g_ntaps[kMaxGroups];
  int g_rowmin[kMaxPhases][kMaxGroups];            // first halo row relative to the tile's first output row
  int g_extra[kMaxPhases][kMaxGroups];             // extra 8-row boxes after the first 128 rows
  int g_shift[kMaxPhases][kMaxGroups][kMaxTaps];   // row shift of each tap inside the halo tile
  int g_wk[kMaxGroups][kMaxTaps];                  // K coordinate of the tap's weight slice for chunk 0
  int w_resident;                                  // all weight tiles stay in shared memory for the CTA's lifetime
  int a_stages, w_stages;                          // ring depths (w_stages unused when resident)
  int eg;                                          // epilogue groups of four warps (2 or 4)
  int wt;                                          // taps per streamed weight stage (1 or 2): longer steps per handshake
  int m_tiles, n_tiles, total_tiles;
  int mb;                                          // 128-row blocks per tile (1, 2 or 4): narrow-N layers amortise per-tile costs
  int pair;                                        // CTA pairs (cta_group::2): each CTA holds half of the weight tile's rows
  int a_bytes;                                     // bytes of one A ring stage: (128*mb + 64) rows
  // Flat MMA issue schedule, [phase][n_steps]: one entry per (group, K chunk, tap) in issue order.
  //   bits 0-15  tap's row shift inside the halo tile, in 16-byte units (shift * row bytes / 16)
  //   bits 16-26 resident weights: index of the tap's weight tile within the (phase, n_tile) block;
  //              streamed weights: index of the tap's tile within its ring stage (< wt)
  //   bits 27-31 kSchedNull / kSchedWFirst / kSchedWLast / kSchedFirst / kSchedLast
  // Every K chunk's taps are padded to an even count, so n_steps is even and a pair never straddles an A stage.
  int n_steps;
  uint32_t sched[kMaxSched];
  int dbg;   // -DMB200_PROBE builds only: 1 skip weight loads, 2 skip A loads, 4 skip MMAs, 8 skip epilogue work
  long long* dbg_out;   // -DMB200_PROBE builds only: per-CTA cycle counters [grid][16]
};

// Geometry of the fused ResBlock step (resblock_pair.cu): c1 (K taps, dilation dil) and c2 (K taps, dilation 1) on C
// channels, tiles of 128*mb intermediate rows that advance by S output rows.
struct PairGeom {
  int B, T;                      // batch, rows per batch element
  int K, dil;
  int S;                         // output rows per tile = 128*mb - (K-1)
  int tiles_per_b, total_tiles;
  int extra;                     // 8-row halo boxes after the first 128*mb rows of the input tile
  int a_stages, a_bytes;         // input ring
  int mid_bytes, mid_stages;     // intermediate tile (128*mb + 16 rows), single or double buffered
  int w_resident, w_stages;      // both convs' weights resident, or streamed two taps per ring stage
  const float* bias1;            // c1 bias [C]
  float slope;                   // LeakyReLU between the convs
  long long* dbg_out;            // developer probe: per-CTA cycle counters [grid][16] (MB200_PAIR_DBG), else null
};

// Geometry / parameters of the fused feed-forward kernel (ff_fused.cu): rows are the flat [B * L] frame index.
struct FFGeom {
  long rows;                     // B * L
  int total_tiles;               // ceil(rows / 128)
  const float* bias1;            // [1024]
  const float* snake_a;          // [1024] exp(alpha)
  const float* snake_invb;       // [1024] 1 / (exp(beta) + 1e-9)
  const float* bias2;            // [256]
  const float* mask;             // [rows]
  const void* resid;             // [rows][256] operand dtype (the attention sub-block's output)
  void* out;                     // [rows][256] operand dtype
  long long* dbg_out;            // developer probe (MB200_FF_DBG): per-CTA cycle counters [grid][16], else null
  // Optional pre-stage (pre != 0): the attention out-projection, residual and LayerNorm that feed the feed-forward
  // (transformer.py:266-296): hid = att Wo^T + bo + res_in -> hid_out (= resid above); x = LayerNorm(hid) stays on chip.
  int pre;
  const float* bias_o;           // [256]
  const void* res_in;            // [rows][256] operand dtype (the block's input, transformer.py:276)
  void* hid_out;                 // [rows][256] operand dtype
  const float* ln_g;             // [256] norm3 weight / bias
  const float* ln_b;
};

// Geometry / parameters of the fused res_conv + GroupNorm/Mish + LayerNorm1 + QKV kernel (resqkv_fused.cu).
struct RQGeom {
  long rows;                     // B * L (flat frame index)
  int L;                         // frames per utterance (GroupNorm statistics are per utterance)
  int total_tiles;               // ceil(rows / 128)
  int nk0, nk1;                  // 64-channel K chunks of the first / second (channel-concat) source
  const float* bias;             // res_conv bias [256]
  const void* gn_raw;            // block2 conv output [rows][256], operand dtype
  const float* gn_stats;         // [B][8][2] sum, sum of squares
  const float2* gn_part;         // or, when not null: the producing conv's per-32-row partial sums [B][8][gn_slots], added here
  int gn_slots;
  const float* gn_gamma;         // [256]
  const float* gn_beta;
  float gn_inv_count;            // 1 / (32 L)
  const float* mask;             // [rows]
  void* res_out;                 // [rows][256] operand dtype
  const float* ln_g;             // norm1 weight / bias [256]
  const float* ln_b;
  void* qkv_out;                 // [rows][384] operand dtype
  long long* dbg_out;            // developer probe (MB200_RQ_DBG): per-CTA timestamps of the epilogue's phases, else null
};

struct EpiParams {
  int T_o;                 // output rows per batch (T_out * o_mul)
  int o_mul;               // output row = t * o_mul + phase * o_phase_step
  int o_phase_step;
  int n_valid;             // real output channels (<= N), multiple of 8
  const float* bias;       // [N] (zero padded) or null
  // v += mish(groupnorm(raw)) * mask      (Block1D tail fused into the res_conv GEMM: decoder.py:37-43,60)
  const void* gn_raw;  int gn_ld;  const float* gn_stats;  const float* gn_gamma;  const float* gn_beta;  float gn_inv_count;
  // v += resid (+ resid2)
  const void* resid;   int resid_ld;   int resid_f32;
  float resid_inv;         // != 1: the stored residual is LeakyReLU(x, 1/resid_inv); x = min(r, r * resid_inv) is what is added
                           // (HiFi-GAN keeps only lrelu(x): one store per conv instead of two; tcgen05 path only)
  const void* resid2;  int resid2_ld;  int resid2_f32;
  float scale;             // v *= scale (HiFi-GAN xs / num_kernels: hifigan/models.py:192)
  const float* snake_a;    // SnakeBeta exp(alpha), 1/(exp(beta)+1e-9) (transformer.py:70-78); null = off
  const float* snake_invb;
  const float* mask;       // [B][T_o] float 0/1 or null; used by gn_raw and mask_out
  int mask_out;            // v *= mask
  // GroupNorm statistics of v, [B][8][2] = (sum, sum of squares) per 32-channel group over all T_out rows, in a FIXED
  // summation order (no float atomics): every warp stores the partial sums of its 32 rows in slot t/32 of stats_part
  // ([B][8][stats_slots] float2, stats_slots = ceil(T_out / 32)); the last CTA of the launch to finish (stats_cnt, ONE
  // counter per launch) adds the slots in index order into stats_out (gn_stats_store / gn_stats_finish below).
  float* stats_out;
  float2* stats_part;
  int* stats_cnt;
  int stats_slots;
  int stats_nofinish;      // the consumer (gn_mish_kernel) adds the slots itself: no counter, no ordered sum at the end of this launch
  // Euler update fused into final_proj (flow_matching.py:79): x += dt * v ; second output = x * mask
  float* x_state;  int x_ld;  float dt;
  void* out0;  int out0_ld;  int out0_f32;  float out0_slope;   // main store; slope != 1 applies LeakyReLU
  void* out1;  int out1_ld;  float out1_slope;                  // second store (operand dtype), LeakyReLU(slope)
  // LayerNorm of the finished row (needs the whole row in one tile: N == 256) -> ln_out (operand dtype)
  void* ln_out;  int ln_ld;  const float* ln_g;  const float* ln_b;
  // Split-precision fp32 mode (bf16 hi/lo operands, fp32 everything else): EVERY activation stream of the epilogue
  // (gn_raw, resid, resid2, out0, out1, ln_out) is fp32, whatever the kernel's operand type T is.
  int all_f32;
  // ... and an output whose *_lo pointer is set is stored as that split pair instead (bf16 hi at out, bf16 lo at out_lo):
  // the next conv reads it directly (Src::lo) and no fp32 copy of the stream exists.
  void* out0_lo;
  void* out1_lo;
};

template <typename T> struct Elem;
template <> struct Elem<float> {
  static __device__ __forceinline__ float ld(const float* p) { return *p; }
  static __device__ __forceinline__ void ld8(const float* p, float* v) {
    float4 a = *reinterpret_cast<const float4*>(p), b = *reinterpret_cast<const float4*>(p + 4);
    v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w; v[4] = b.x; v[5] = b.y; v[6] = b.z; v[7] = b.w;
  }
  static __device__ __forceinline__ void st8(float* p, const float* v) {
    *reinterpret_cast<float4*>(p) = make_float4(v[0], v[1], v[2], v[3]);
    *reinterpret_cast<float4*>(p + 4) = make_float4(v[4], v[5], v[6], v[7]);
  }
};
template <> struct Elem<__nv_bfloat16> {
  static __device__ __forceinline__ float ld(const __nv_bfloat16* p) { return __bfloat162float(*p); }
  static __device__ __forceinline__ void ld8(const __nv_bfloat16* p, float* v) {
    uint4 u = *reinterpret_cast<const uint4*>(p);
    const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&u);
#pragma unroll
    for (int i = 0; i < 4; ++i) { float2 f = __bfloat1622float2(h[i]); v[2 * i] = f.x; v[2 * i + 1] = f.y; }
  }
  static __device__ __forceinline__ void st8(__nv_bfloat16* p, const float* v) {
    uint4 u; __nv_bfloat162* h = reinterpret_cast<__nv_bfloat162*>(&u);
#pragma unroll
    for (int i = 0; i < 4; ++i) h[i] = __floats2bfloat162_rn(v[2 * i], v[2 * i + 1]);
    *reinterpret_cast<uint4*>(p) = u;
  }
};
template <> struct Elem<__half> {
  static __device__ __forceinline__ float ld(const __half* p) { return __half2float(*p); }
  static __device__ __forceinline__ void ld8(const __half* p, float* v) {
    uint4 u = *reinterpret_cast<const uint4*>(p);
    const __half2* h = reinterpret_cast<const __half2*>(&u);
#pragma unroll
    for (int i = 0; i < 4; ++i) { float2 f = __half22float2(h[i]); v[2 * i] = f.x; v[2 * i + 1] = f.y; }
  }
  static __device__ __forceinline__ void st8(__half* p, const float* v) {
    uint4 u; __half2* h = reinterpret_cast<__half2*>(&u);
#pragma unroll
    for (int i = 0; i < 4; ++i) h[i] = __floats2half2_rn(v[2 * i], v[2 * i + 1]);  // saturates to +-inf only beyond 65504
    *reinterpret_cast<uint4*>(p) = u;
  }
};

// x * tanh(softplus(x)) with torch's softplus threshold 20 (decoder.py:38).  With n = e^x,
// tanh(log(1+n)) = (n^2+2n)/(n^2+2n+2): one exp, one divide, no cancellation for x << 0.
__device__ __forceinline__ float mish_f(float x) {
  if (x > 20.f) return x;
  float n = expf(x);
  float w = n * (n + 2.f);
  return x * (w / (w + 2.f));
}
__device__ __forceinline__ float lrelu_f(float x, float slope) { return x > 0.f ? x : x * slope; }

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// GroupNorm statistics in a FIXED summation order, without float atomics (run-to-run deterministic):
//  * gn_stats_store: one lane of a warp stores the warp's partial sums over rows [32 slot, 32 slot + 32) of (b, group) into
//    its own slot - a plain store, nothing waits for it;
//  * gn_stats_finish: at the end of the kernel every CTA bumps ONE per-launch counter; the CTA that arrives last adds the
//    slots of every (b, group) in index order into stats_out and re-arms the counter.  Which CTA is last changes from run
//    to run, what it computes does not.  (A first version committed per item with fence + atomic-with-return; the fence
//    waited for the thread's output-tile stores each time: +8 us on a 20 us launch, profiles/r02_launch_list_summary.txt.)
__device__ __forceinline__ void gn_stats_store(const EpiParams& e, int b, int group, int slot, float s, float ss) {
  __stcg(e.stats_part + static_cast<size_t>(b * 8 + group) * e.stats_slots + slot, make_float2(s, ss));
}
// Called by ALL threads of the CTA (contains __syncthreads), after its last gn_stats_store.  s_last_p: one int of shared
// memory the caller owns (the persistent kernel has no static shared memory to spare beside its 227 KB).
// Measured per launch of the Block1D conv (ncu launch lists, cfg2): round 1's float atomics 19.7 us; fence + atomic per
// item 28.2; this form 25.4; fencing right after a warp's last tile, or summing with 16-lane teams and a shuffle tree,
// 29-30 (slower: kept simple).  Bit-identical runs cost about 0.7 ms of a 31.7 ms cfg2 step.
__device__ __forceinline__ void gn_stats_publish() {}
__device__ __forceinline__ void gn_stats_finish(const EpiParams& e, int B, unsigned n_ctas, int* s_last_p) {
  volatile int& s_last = *s_last_p;
  __threadfence();                    // every thread publishes its own stores (partial sums among them) device-wide ...
  __syncthreads();                    // ... before the CTA checks in
  if (threadIdx.x == 0) s_last = atomicAdd(e.stats_cnt, 1) == static_cast<int>(n_ctas) - 1;
  __syncthreads();
  if (!s_last) return;
  __threadfence();
  for (int bg = threadIdx.x; bg < B * 8; bg += blockDim.x) {
    const float2* part = e.stats_part + static_cast<size_t>(bg) * e.stats_slots;
    float a = 0.f, q = 0.f;
    for (int i = 0; i < e.stats_slots; ++i) {
      const float2 p = __ldcg(part + i);
      a += p.x; q += p.y;
    }
    e.stats_out[bg * 2] = a;
    e.stats_out[bg * 2 + 1] = q;
  }
  if (threadIdx.x == 0) *e.stats_cnt = 0;
}

// Fused epilogue for one thread's 32-channel slice [n0, n0+32) of output row (b, t, phase).
// Called by ALL lanes of a warp (GroupNorm partial sums are warp-reduced); `valid` gates memory traffic.
// v holds the fp32 accumulators on entry and the finished values on exit (what LayerNorm consumes).
template <typename T>
__device__ __forceinline__ void epilogue_chunk(const EpiParams& e, int b, int t, int phase, int n0, bool valid,
                                               bool do_store, float (&v)[32]) {
  const int to = t * e.o_mul + phase * e.o_phase_step;
  const long row = static_cast<long>(b) * e.T_o + to;
  const int nv = min(32, e.n_valid - n0);  // channels of this chunk that exist (multiple of 8, may be <= 0)
  if (e.bias) {
#pragma unroll
    for (int i = 0; i < 32; ++i) v[i] += __ldg(e.bias + n0 + i);
  }
  const float m = (e.mask && valid) ? __ldg(e.mask + row) : 1.f;
  if (e.gn_raw && valid) {
    const float* st = e.gn_stats + (static_cast<long>(b) * 8 + (n0 >> 5)) * 2;
    const float mean = st[0] * e.gn_inv_count;
    const float var = fmaxf(st[1] * e.gn_inv_count - mean * mean, 0.f);
    const float rstd = rsqrtf(var + 1e-5f);
    const T* raw = reinterpret_cast<const T*>(e.gn_raw) + row * e.gn_ld + n0;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      float r[8];
      Elem<T>::ld8(raw + 8 * j, r);
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        const int n = n0 + 8 * j + i;
        const float y = (r[i] - mean) * rstd * __ldg(e.gn_gamma + n) + __ldg(e.gn_beta + n);
        v[8 * j + i] += mish_f(y) * m;
      }
    }
  }
  if (e.resid && valid) {
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      if (8 * j < nv) {
        float r[8];
        if (e.resid_f32) Elem<float>::ld8(reinterpret_cast<const float*>(e.resid) + row * e.resid_ld + n0 + 8 * j, r);
        else Elem<T>::ld8(reinterpret_cast<const T*>(e.resid) + row * e.resid_ld + n0 + 8 * j, r);
#pragma unroll
        for (int i = 0; i < 8; ++i) v[8 * j + i] += r[i];
      }
    }
  }
  if (e.resid2 && valid) {
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      if (8 * j < nv) {
        float r[8];
        if (e.resid2_f32) Elem<float>::ld8(reinterpret_cast<const float*>(e.resid2) + row * e.resid2_ld + n0 + 8 * j, r);
        else Elem<T>::ld8(reinterpret_cast<const T*>(e.resid2) + row * e.resid2_ld + n0 + 8 * j, r);
#pragma unroll
        for (int i = 0; i < 8; ++i) v[8 * j + i] += r[i];
      }
    }
  }
  if (e.scale != 1.f) {
#pragma unroll
    for (int i = 0; i < 32; ++i) v[i] *= e.scale;
  }
  if (e.snake_a) {
#pragma unroll
    for (int i = 0; i < 32; ++i) {
      const float s = sinf(v[i] * __ldg(e.snake_a + n0 + i));
      v[i] += __ldg(e.snake_invb + n0 + i) * s * s;
    }
  }
  if (e.mask_out) {
#pragma unroll
    for (int i = 0; i < 32; ++i) v[i] *= m;
  }
  if (e.stats_out) {  // warp-collective: every lane takes part, invalid rows contribute zero
    float s = 0.f, ss = 0.f;
    if (valid) {
#pragma unroll
      for (int i = 0; i < 32; ++i) { s += v[i]; ss += v[i] * v[i]; }
    }
    s = warp_sum(s);
    ss = warp_sum(ss);
    // all rows of a warp belong to the same batch element b (tiles never straddle batch elements); lane 0 holds the
    // warp's first row, a multiple of 32
    if ((threadIdx.x & 31) == 0 && do_store && valid) gn_stats_store(e, b, n0 >> 5, t >> 5, s, ss);
  }
  if (!valid) return;
  if (e.x_state) {
    float* xs = e.x_state + row * e.x_ld + n0;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      if (8 * j < nv) {
        float x[8];
        Elem<float>::ld8(xs + 8 * j, x);
#pragma unroll
        for (int i = 0; i < 8; ++i) x[i] += e.dt * v[8 * j + i];
        if (do_store) Elem<float>::st8(xs + 8 * j, x);
#pragma unroll
        for (int i = 0; i < 8; ++i) v[8 * j + i] = x[i] * m;
      }
    }
  }
  if (!do_store) return;
  if (e.out0) {
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      if (8 * j < nv) {
        float o[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) o[i] = e.out0_slope == 1.f ? v[8 * j + i] : lrelu_f(v[8 * j + i], e.out0_slope);
        if (e.out0_f32) Elem<float>::st8(reinterpret_cast<float*>(e.out0) + row * e.out0_ld + n0 + 8 * j, o);
        else Elem<T>::st8(reinterpret_cast<T*>(e.out0) + row * e.out0_ld + n0 + 8 * j, o);
      }
    }
  }
  if (e.out1) {
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      if (8 * j < nv) {
        float o[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) o[i] = e.out1_slope == 1.f ? v[8 * j + i] : lrelu_f(v[8 * j + i], e.out1_slope);
        Elem<T>::st8(reinterpret_cast<T*>(e.out1) + row * e.out1_ld + n0 + 8 * j, o);
      }
    }
  }
}

}  // namespace mb
