// tcgen05 / TMEM / TMA self-attention for the half-precision modes (same contract as attention.cu):
//   out[b, q, h*64:(h+1)*64] = softmax_k(q . k + mask[b, k]) v      2 heads x 64, ALL L keys, additive 0/1 key mask
// (diffusers 0.25.0 Attention / AttnProcessor2_0 as called at matcha/models/components/transformer.py:266-271 with the
// float mask of decoder.py:396-401: the mask is ADDED to the logits, padded keys are attended - SURVEY.md section 0.2).
// W_q carries the 1/sqrt(64) scale (estimator.cu), so the logits are q . k.
//
// Persistent kernel, one CTA per SM.  A work item is one (batch, head, PAIR of 128-query tiles); both tiles of a pair
// share the K/V stream.  Per 128-key tile j and query tile w in {0, 1}:
//     S_w = Q_w K_j^T          tcgen05.mma M=128 N=128 K=64, both operands K-major 128B-swizzled TMA tiles -> TMEM (128 cols)
//     P_w = exp2(S_w log2e + bias_j - m_w)      one softmax warpgroup per query tile, ONE ROW PER THREAD (tcgen05.ld
//                                               32x32b gives exactly that: row max / row sum need no shuffles, no
//                                               barriers), written as the K-major 128B-swizzled A operand into shared memory
//     O_w += P_w V_j           tcgen05.mma M=128 N=64 K=128; V is used as TMA delivers it ([keys][64] rows of 128 B):
//                              an MN-MAJOR B operand (instruction-descriptor bit 16), no transpose anywhere
// The issuer interleaves S_w(j+1) before P_w V(j), so the tensor pipe works on one query tile while the other tile's
// warpgroup is in its exponentials.  The running maximum is only raised when a tile's maximum exceeds it by more than 12
// (a factor 4096: P stays inside the fp16 range, the fp32 row sum and accumulator absorb it), and only then is the O accumulator
// rescaled in TMEM (tcgen05.ld / tcgen05.st) - after the first tiles that is rare.
// The floors per (query tile, key tile) on this part are the TMEM read of S (64 KB at 64 B/cycle = 1024 cycles) and
// the exponentials (16384 at 16 per cycle = 1024 cycles); the MMAs need 512.  The softmax therefore makes ONE streaming
// pass over S with the running maximum of the PREVIOUS tiles: 32-column chunk c+1 is in flight from TMEM while chunk c goes
// through FMA -> ex2 -> pack, and the tile maximum is tracked on the side; only if it exceeds the running maximum by
// more than 12 (or the mask is not uniform over the key tile) is the tile redone in two passes.  A warpgroup signals the
// issuer as soon as S has been read out of TMEM, so the next tile's S is computed under the rest of this tile's work.
// Warps: 0 = TMA producer (Q pair, K/V ring of three 32 KB stages), 1 = MMA issuer + TMEM owner, 2-5 / 6-9 = softmax.
#include "epilogue3.cuh"
#include "launch.h"
#include "ptx.cuh"

namespace mb {

namespace {
constexpr float kLog2e = 1.4426950408889634f;
constexpr float kLn2 = 0.6931471805599453f;
constexpr int kKvStages = 3;
constexpr float kRaise = 12.f;   // the running maximum is raised only when a tile exceeds it by more than this (log2 units)
constexpr int kTile = 16384;                            // 128 rows x 128 bytes
constexpr int kOffQ = 0;                                // 2 query tiles
constexpr int kOffKV = 2 * kTile;                       // ring of (K tile, V tile)
constexpr int kOffP = kOffKV + kKvStages * 2 * kTile;   // 2 query tiles x 2 K chunks of 64 keys
constexpr int kOffBar = kOffP + 4 * kTile;
constexpr int kAttnThreads = 64 + 2 * 128;              // producer, issuer, one softmax warpgroup per query tile
constexpr int kAttnSmem = kOffBar + 256 + 1024;         // + alignment slack

__device__ __forceinline__ float ex2_ftz(float x) {
  float r;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
  return r;
}
// Issued where it is written: the compiler sinks a plain __ldg down to its first use, one tile later (measured).
__device__ __forceinline__ float ldg_now(const float* p) {
  float v;
  asm volatile("ld.global.nc.f32 %0, [%1];" : "=f"(v) : "l"(p));
  return v;
}
__device__ __forceinline__ float max3(float a, float b, float c) {
  float r;
  asm("max.f32 %0, %1, %2, %3;" : "=f"(r) : "f"(a), "f"(b), "f"(c));
  return r;
}
// tcgen05.wait::ld that also names the 32 destination registers of the load it completes: the compiler sees a data
// dependence and cannot schedule their consumers above the wait (loads are issued one chunk ahead of their use here).
__device__ __forceinline__ void tmem_ld_wait_on(float* v) {
  uint32_t* r = reinterpret_cast<uint32_t*>(v);
  asm volatile("tcgen05.wait::ld.sync.aligned;"
               : "+r"(r[0]), "+r"(r[1]), "+r"(r[2]), "+r"(r[3]), "+r"(r[4]), "+r"(r[5]), "+r"(r[6]), "+r"(r[7]), "+r"(r[8]),
                 "+r"(r[9]), "+r"(r[10]), "+r"(r[11]), "+r"(r[12]), "+r"(r[13]), "+r"(r[14]), "+r"(r[15]), "+r"(r[16]),
                 "+r"(r[17]), "+r"(r[18]), "+r"(r[19]), "+r"(r[20]), "+r"(r[21]), "+r"(r[22]), "+r"(r[23]), "+r"(r[24]),
                 "+r"(r[25]), "+r"(r[26]), "+r"(r[27]), "+r"(r[28]), "+r"(r[29]), "+r"(r[30]), "+r"(r[31])
               :
               : "memory");
}
template <typename T> __device__ __forceinline__ uint32_t pack2(float a, float b);
template <> __device__ __forceinline__ uint32_t pack2<__half>(float a, float b) {
  const __half2 h = __floats2half2_rn(a, b);
  return *reinterpret_cast<const uint32_t*>(&h);
}
template <> __device__ __forceinline__ uint32_t pack2<__nv_bfloat16>(float a, float b) {
  const __nv_bfloat162 h = __floats2bfloat162_rn(a, b);
  return *reinterpret_cast<const uint32_t*>(&h);
}
}  // namespace

template <typename T>
__global__ void __launch_bounds__(kAttnThreads, 1)
attention_umma_kernel(const __grid_constant__ CUtensorMap tmQKV, const float* __restrict__ mask, T* __restrict__ out,
                      const int B, const int L, long long* __restrict__ dbg) {
  // developer probe (MB200_ATTN_DBG): cumulative clock64 cycles per phase of softmax warp 2 and of the issuer, CTA 0
#ifdef MB200_PROBE   // MB200_PROBE_BUILD=1 builds only: the counters cost 22 registers and the hot loop has none to spare
#define AT_T0() long long at_last = clock64(), at_acc[10] = {0, 0, 0, 0, 0, 0, 0, 0, 0, 0}
#define AT_MARK(k) do { if (dbg) { const long long at_now = clock64(); at_acc[k] += at_now - at_last; at_last = at_now; } } while (0)
#else
#define AT_T0() do {} while (0)
#define AT_MARK(k) do {} while (0)
#endif
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + kOffBar);
  uint64_t* q_full = bars;                 // [1]
  uint64_t* q_empty = bars + 1;            // [1]
  uint64_t* kv_full = bars + 2;            // [3]
  uint64_t* kv_empty = bars + 5;           // [3]
  uint64_t* s_full = bars + 8;             // [2] S_w ready in TMEM
  uint64_t* p_full = bars + 10;            // [2] P_w in shared memory (and O_w rescaled): four warp arrivals
  uint64_t* pv_done = bars + 12;           // [2] O_w += P_w V complete: P_w reusable, O_w readable
  uint64_t* s_read = bars + 14;            // [2] S_w has been read out of TMEM (four warp arrivals): the next S_w may be issued
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 16);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int n_pairs = (L + 255) / 256, nkt = (L + 127) / 128;
  const int items = B * 2 * n_pairs;

  if (threadIdx.x == 0) {
    tma_prefetch_desc(&tmQKV);
    mbar_init(q_full, 1); mbar_init(q_empty, 1);
    for (int s = 0; s < kKvStages; ++s) { mbar_init(&kv_full[s], 1); mbar_init(&kv_empty[s], 1); }
    for (int w = 0; w < 2; ++w) { mbar_init(&s_full[w], 1); mbar_init(&p_full[w], 4); mbar_init(&pv_done[w], 1); mbar_init(&s_read[w], 4); }
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc<512>(tmem_slot);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    // ================================================================= TMA producer
    int kc = 0, it = 0;
    for (int item = blockIdx.x; item < items; item += gridDim.x, ++it) {
      const int pair = item % n_pairs, bh = item / n_pairs, h = bh & 1, b = bh >> 1;
      const int q0 = pair * 256;
      const bool two = q0 + 128 < L;
      mbar_wait(q_empty, (it & 1) ^ 1);
      if (elect_one()) {
        mbar_expect_tx(q_full, (two ? 2 : 1) * kTile);
        tma_load_3d(smem + kOffQ, &tmQKV, q_full, h * 64, q0, b);
        if (two) tma_load_3d(smem + kOffQ + kTile, &tmQKV, q_full, h * 64, q0 + 128, b);
      }
      __syncwarp();
      for (int j = 0; j < nkt; ++j, ++kc) {
        const int st = kc % kKvStages;
        mbar_wait(&kv_empty[st], ((kc / kKvStages) & 1) ^ 1);
        if (elect_one()) {
          uint8_t* dst = smem + kOffKV + st * 2 * kTile;
          mbar_expect_tx(&kv_full[st], 2 * kTile);
          tma_load_3d(dst, &tmQKV, &kv_full[st], 128 + h * 64, j * 128, b);
          tma_load_3d(dst + kTile, &tmQKV, &kv_full[st], 256 + h * 64, j * 128, b);
        }
        __syncwarp();
      }
    }
  } else if (warp == 1) {
    // ================================================================= MMA issuer
    constexpr int fmt = std::is_same<T, __nv_bfloat16>::value ? 1 : 0;
    constexpr uint32_t idesc_s = umma_idesc_f16(fmt, 128, 128);
    constexpr uint32_t idesc_o = umma_idesc_f16_bmn(fmt, 128, 64);
    const uint64_t q_desc0 = umma_smem_desc<128>(smem_u32(smem + kOffQ));
    const uint64_t p_desc0 = umma_smem_desc<128>(smem_u32(smem + kOffP));
    int kc = 0, it = 0;
    uint32_t base[2] = {0, 0};   // key tiles of earlier items per query tile: parities of its barriers
    AT_T0();
    for (int item = blockIdx.x; item < items; item += gridDim.x, ++it) {
      const int pair = item % n_pairs;
      const int q0 = pair * 256;
      const int nw = q0 + 128 < L ? 2 : 1;
      const int kc0 = kc;         // ring position of this item's first key tile
      auto issue_s = [&](int w, int st) {   // S_w = Q_w K^T for the K tile in ring stage st
        const uint64_t k_desc = umma_smem_desc<128>(smem_u32(smem + kOffKV + st * 2 * kTile));
        const uint64_t q_desc = q_desc0 + static_cast<uint64_t>(w * (kTile >> 4));
#pragma unroll
        for (int k = 0; k < 4; ++k) umma_f16(tmem_base + w * 128, q_desc + 2 * k, k_desc + 2 * k, idesc_s, k != 0);
        umma_commit(&s_full[w]);
      };
      mbar_wait(q_full, it & 1);
      mbar_wait(&kv_full[kc0 % kKvStages], (kc0 / kKvStages) & 1);
      tc_fence_after();
      if (elect_one()) {
        for (int w = 0; w < nw; ++w) issue_s(w, kc0 % kKvStages);
      }
      __syncwarp();
      AT_MARK(0);
      // Polling state machine over both query tiles.  Per tile t of query tile w:
      //   S_w(t)   may be issued once S_w(t-1) has been read out of TMEM (s_read) and K(t) has landed;
      //   P_w V(t) once P_w(t) is in shared memory (p_full).
      // Neither query tile ever blocks the other: the two softmax warpgroups drift into a ping-pong by themselves.
      int js[2] = {1 < nkt ? 1 : nkt, nw == 2 ? (1 < nkt ? 1 : nkt) : nkt};   // next S tile per query tile
      int jp[2] = {0, nw == 2 ? 0 : nkt};                                      // next P V tile
      int kv_seen = 1, released = 0;
      uint32_t spins = 0;
      while (jp[0] < nkt || jp[1] < nkt) {
#pragma unroll
        for (int w = 0; w < 2; ++w) {
          if (js[w] < nkt) {
            if (kv_seen <= js[w]) {
              const int t = kc0 + kv_seen;
              if (mbar_test(&kv_full[t % kKvStages], (t / kKvStages) & 1)) ++kv_seen;
            }
            if (kv_seen > js[w] && mbar_test(&s_read[w], (base[w] + js[w] - 1) & 1)) {
              tc_fence_after();
              if (elect_one()) issue_s(w, (kc0 + js[w]) % kKvStages);
              __syncwarp();
              ++js[w];
              spins = 0;
            }
          }
          if (jp[w] < nkt && mbar_test(&p_full[w], (base[w] + jp[w]) & 1)) {
            tc_fence_after();
            const int j = jp[w], st = (kc0 + j) % kKvStages;
            ++jp[w];
            const int done = jp[0] < jp[1] ? jp[0] : jp[1];   // key tiles whose P V products have all been issued
            if (elect_one()) {
              // O_w (+)= P_w V_j : 8 K steps of 16 keys; a P chunk holds 64 keys, V advances 16 key rows = 2048 bytes
              const uint32_t v_addr = smem_u32(smem + kOffKV + st * 2 * kTile + kTile);
#pragma unroll
              for (int k = 0; k < 8; ++k) {
                const uint64_t a = p_desc0 + static_cast<uint64_t>((w * 2 * kTile + (k >> 2) * kTile) >> 4) + 2 * (k & 3);
                const uint64_t bd = umma_smem_desc_mn128(v_addr + k * 2048, 8192);
                umma_f16(tmem_base + 256 + w * 64, a, bd, idesc_o, (j | k) != 0);
              }
              umma_commit(&pv_done[w]);
              for (int t = released; t < done; ++t) umma_commit(&kv_empty[(kc0 + t) % kKvStages]);   // ring stage free
              if (done == nkt && released < nkt) umma_commit(q_empty);                               // and Q at the end
            }
            __syncwarp();
            released = done;
            spins = 0;
          }
        }
        if (++spins > (1u << 24)) __trap();   // a pipeline bug must surface as a failed launch, never as a hang
        if (spins > 1) __nanosleep(20);       // nothing was ready: leave the scheduler's issue slots to the softmax warps
      }
      AT_MARK(1);
      base[0] += nkt;
      if (nw == 2) base[1] += nkt;
      kc += nkt;
    }
#ifdef MB200_PROBE
    if (dbg && blockIdx.x == 0 && lane == 0) { for (int k = 0; k < 10; ++k) dbg[16 + k] = at_acc[k]; }
#endif
  } else {
    // ================================================================= softmax warpgroups (one query row per thread)
    const int w = (warp - 2) >> 2;
    const int r = (warp & 3) * 32 + lane;            // TMEM lane = query row inside the tile
    const uint32_t t_s = tmem_base + (static_cast<uint32_t>((warp & 3) * 32) << 16) + w * 128;
    const uint32_t t_o = tmem_base + (static_cast<uint32_t>((warp & 3) * 32) << 16) + 256 + w * 64;
    uint8_t* p_row = smem + kOffP + w * 2 * kTile + r * 128;   // + chunk * kTile + swizzled 16-byte unit
    uint32_t nt = 0;                                  // key tiles processed so far (all items): barrier parities
    AT_T0();
    for (int item = blockIdx.x; item < items; item += gridDim.x) {
      const int pair = item % n_pairs, bh = item / n_pairs, h = bh & 1, b = bh >> 1;
      const int q_first = pair * 256 + w * 128;
      if (q_first >= L) continue;                     // this pair has one query tile only
      const float* mrow = mask + static_cast<size_t>(b) * L;
      float m_used = 0.f, l2[2] = {0.f, 0.f};
      // Additive bias (log2 domain) of keys lane, lane + 32, lane + 64, lane + 96 of the tile, fetched ONE TILE AHEAD; keys
      // past the end of the sequence do not exist (-inf).  Key k of the tile lives in lane k % 32, slot k / 32.
      float mkb[4];
      auto fetch_bias = [&](int jt) {
#pragma unroll
        for (int c = 0; c < 4; ++c) {
          const int key = jt * 128 + c * 32 + lane;
          mkb[c] = key < L ? ldg_now(mrow + key) : -INFINITY;     // scaled by log2e at its use
        }
      };
      fetch_bias(0);
      for (int j = 0; j < nkt; ++j, ++nt) {
        AT_MARK(9);
        // uniform = every key of the tile exists and carries the same bias (always, except at most two tiles of a
        // prefix-masked utterance): the bias then folds into the exponent's additive constant.  Any other tile is MADE
        // uniform first: its scores are rewritten in TMEM with each key's bias difference folded in (see below).
        float bcur[4] = {mkb[0] * kLog2e, mkb[1] * kLog2e, mkb[2] * kLog2e, mkb[3] * kLog2e};
        const float bias_u = __shfl_sync(0xffffffffu, bcur[0], 0);
        const bool uni = __all_sync(0xffffffffu, bcur[0] == bias_u && bcur[1] == bias_u && bcur[2] == bias_u && bcur[3] == bias_u);
        if (j + 1 < nkt) fetch_bias(j + 1);
        AT_MARK(0);
        mbar_wait(&s_full[w], nt & 1);
        AT_MARK(1);
        tc_fence_after();
        uint32_t pk[64];
        float part[2] = {0.f, 0.f};                   // this tile's contribution to the row sum
        if (!uni) {
          // Irregular tile (a mask edge, or the keys run out): S[k] += (bias[k] - bias[key 0]) in place, -inf for a key that
          // does not exist, 32-key chunk by chunk and only where a chunk needs it.  One tcgen05.ld / st round trip per
          // chunk (~150 cycles) instead of a separate slow path through the whole tile.
          float sv[32];
          for (int ch = 0; ch < 4; ++ch) {
            const float mine = ch == 0 ? bcur[0] : ch == 1 ? bcur[1] : ch == 2 ? bcur[2] : bcur[3];   // log2 domain
            if (__all_sync(0xffffffffu, mine == bias_u)) continue;
            tmem_ld_32x32(t_s + 32 * ch, sv);
            tmem_ld_wait_on(sv);
#pragma unroll
            for (int i = 0; i < 32; ++i) sv[i] = fmaf(__shfl_sync(0xffffffffu, mine, i) - bias_u, kLn2, sv[i]);
            tmem_st_32x32(t_s + 32 * ch, sv);
          }
          tmem_st_wait();
        }
        bool redo = false;
        bool waited = j == 0;                         // nothing of this item is in flight on P_w / O_w at its first tile
        auto signal_s_read = [&]() {                  // S_w is out of TMEM: the issuer may overwrite it with the next tile's
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(&s_read[w]);
        };
        {
          // ---- ONE streaming pass with the running maximum of the previous tiles (first tile: of its first 32 keys)
          float m0 = -INFINITY, m1 = -INFINITY;
          float sa[32], sb[32];
          auto track = [&](const float* v) {
#pragma unroll
            for (int i = 0; i < 8; ++i) { m0 = max3(m0, v[4 * i], v[4 * i + 1]); m1 = max3(m1, v[4 * i + 2], v[4 * i + 3]); }
          };
          tmem_ld_32x32(t_s, sa);
          tmem_ld_wait_on(sa);
          tmem_ld_32x32(t_s + 32, sb);
          track(sa);
          if (j == 0) m_used = fmaf(fmaxf(m0, m1), kLog2e, bias_u);
          const float c0 = bias_u - m_used;
          auto expo = [&](const float* v, int ch) {
#pragma unroll
            for (int i = 0; i < 16; ++i) {
              const float p0 = ex2_ftz(fmaf(v[2 * i], kLog2e, c0)), p1 = ex2_ftz(fmaf(v[2 * i + 1], kLog2e, c0));
              part[i & 1] += p0 + p1;
              pk[ch * 16 + i] = pack2<T>(p0, p1);
            }
          };
          expo(sa, 0);
          tmem_ld_wait_on(sb);
          tmem_ld_32x32(t_s + 64, sa);
          track(sb);
          expo(sb, 1);
          tmem_ld_wait_on(sa);
          tmem_ld_32x32(t_s + 96, sb);
          track(sa);
          expo(sa, 2);
          tmem_ld_wait_on(sb);
          track(sb);
          const float mt = fmaf(fmaxf(m0, m1), kLog2e, bias_u);
          redo = __any_sync(0xffffffffu, mt > m_used + kRaise);   // rare: the running maximum has to be raised
          if (!redo) {
            signal_s_read();
            expo(sb, 3);
          }
        }
        AT_MARK(2);
        if (redo) {
          // ---- two passes over S (it stays in TMEM): the running maximum has to be raised before this tile's exponentials
          {
            float sv[32];
            float m0 = -INFINITY, m1 = -INFINITY;
#pragma unroll
            for (int ch = 0; ch < 4; ++ch) {
              tmem_ld_32x32(t_s + 32 * ch, sv);
              tmem_ld_wait_on(sv);
#pragma unroll
              for (int i = 0; i < 32; ++i) {
                const float x = fmaf(sv[i], kLog2e, bias_u);
                if (i & 1) m1 = fmaxf(m1, x); else m0 = fmaxf(m0, x);
              }
            }
            const float mt = fmaxf(m0, m1);
            if (j == 0) {
              m_used = mt;
            } else {
              const bool raise = mt > m_used + kRaise;
              if (__any_sync(0xffffffffu, raise)) {   // tcgen05.ld / st are warp-collective: the whole warp takes part
                mbar_wait(&pv_done[w], (nt - 1) & 1); // O_w stable
                waited = true;
                tc_fence_after();
                const float sc = raise ? ex2_ftz(m_used - mt) : 1.f;
                if (raise) m_used = mt;
                l2[0] *= sc; l2[1] *= sc;
#pragma unroll
                for (int c = 0; c < 2; ++c) {
                  tmem_ld_32x32(t_o + c * 32, sv);
                  tmem_ld_wait_on(sv);
#pragma unroll
                  for (int i = 0; i < 32; ++i) sv[i] *= sc;
                  tmem_st_32x32(t_o + c * 32, sv);
                }
                tmem_st_wait();
              }
            }
            part[0] = 0.f; part[1] = 0.f;
#pragma unroll
            for (int ch = 0; ch < 4; ++ch) {
              tmem_ld_32x32(t_s + 32 * ch, sv);
              tmem_ld_wait_on(sv);
              if (ch == 3) signal_s_read();
#pragma unroll
              for (int i = 0; i < 16; ++i) {
                const float cr = bias_u - m_used;
                const float p0 = ex2_ftz(fmaf(sv[2 * i], kLog2e, cr)), p1 = ex2_ftz(fmaf(sv[2 * i + 1], kLog2e, cr));
                part[i & 1] += p0 + p1;
                pk[ch * 16 + i] = pack2<T>(p0, p1);
              }
            }
          }
        }
        AT_MARK(3);
        l2[0] += part[0]; l2[1] += part[1];
        if (!waited) mbar_wait(&pv_done[w], (nt - 1) & 1);   // P_w free again
        AT_MARK(4);
#pragma unroll
        for (int u = 0; u < 16; ++u)                  // 16-byte unit u = keys [8u, 8u + 8)
          *reinterpret_cast<uint4*>(p_row + (u >> 3) * kTile + (((u & 7) ^ (r & 7)) << 4)) =
              make_uint4(pk[4 * u], pk[4 * u + 1], pk[4 * u + 2], pk[4 * u + 3]);
        tc_fence_before();
        fence_proxy_async();          // generic-proxy stores -> visible to the tensor core's async-proxy reads
        __syncwarp();
        if (lane == 0) mbar_arrive(&p_full[w]);
        AT_MARK(5);
      }
      // ---- O_w / l -> out[b, q, h*64 ...]
      const float l_sum = l2[0] + l2[1];
      mbar_wait(&pv_done[w], (nt - 1) & 1);
      tc_fence_after();
      const float inv = 1.f / l_sum;
      const int q = q_first + r;
      T* orow = out + (static_cast<size_t>(b) * L + q) * 128 + h * 64;
#pragma unroll
      for (int c = 0; c < 2; ++c) {
        float o[32];
        tmem_ld_32x32(t_o + c * 32, o);
        tmem_ld_wait_on(o);
#pragma unroll
        for (int i = 0; i < 32; ++i) o[i] *= inv;
        if (q < L) {
#pragma unroll
          for (int u = 0; u < 4; ++u) *reinterpret_cast<uint4*>(orow + c * 32 + 8 * u) = pack8<T>(o + 8 * u);
        }
      }
      tc_fence_before();
    }
    AT_MARK(9);
#ifdef MB200_PROBE
    if (dbg && blockIdx.x == 0 && warp == 2 && lane == 0) { for (int k = 0; k < 10; ++k) dbg[k] = at_acc[k]; }
#endif
  }
#undef AT_T0
#undef AT_MARK
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc<512>(tmem_base);
}

template <typename T>
cudaError_t launch_attention_umma(const AttnLaunch& A, cudaStream_t st) {
  auto kern = attention_umma_kernel<T>;
  static PerDeviceOnce configured;
  if (configured.needed()) {
    cudaError_t err = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, kAttnSmem);
    if (err != cudaSuccess) return err;
    configured.mark();
  }
  kern<<<A.grid, kAttnThreads, kAttnSmem, st>>>(A.tmQKV, A.mask, static_cast<T*>(A.out), A.B, A.L, A.dbg);
  return cudaGetLastError();
}
template cudaError_t launch_attention_umma<__half>(const AttnLaunch&, cudaStream_t);
template cudaError_t launch_attention_umma<__nv_bfloat16>(const AttnLaunch&, cudaStream_t);

}  // namespace mb
