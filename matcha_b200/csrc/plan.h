// Host-side plan: packed weights, workspaces, TMA descriptors, launch programs and their CUDA graphs.
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>

#include <functional>
#include <map>
#include <memory>
#include <mutex>
#include <stdexcept>
#include <string>
#include <tuple>
#include <vector>

#include "../../include/matcha_b200.h"
#include "launch.h"

namespace mb {

struct Error : std::runtime_error {
  using std::runtime_error::runtime_error;
};
void set_last_error(const std::string& s);

#define MB_CUDA(expr)                                                                                       \
  do {                                                                                                      \
    cudaError_t _e = (expr);                                                                                \
    if (_e != cudaSuccess)                                                                                  \
      throw ::mb::Error(std::string(#expr) + ": " + cudaGetErrorString(_e) + " (" __FILE__ ":" + std::to_string(__LINE__) + ")"); \
  } while (0)

inline size_t esize(int dtype) { return dtype == MB200_F32 ? 4 : 2; }

// Device allocation owned by the plan (freed with it).
struct DevBuf {
  void* p = nullptr;
  size_t bytes = 0;
};

// A reference state_dict tensor on the host.
struct HostTensor {
  const float* data;
  std::vector<long long> shape;
  long long numel() const { long long n = 1; for (auto s : shape) n *= s; return n; }
};
using StateDict = std::map<std::string, HostTensor>;

// One conv / linear repacked as the GEMM B operand: [n_phases * N][K_total] in the plan dtype, K ordered
// (tap, source, channel) with every source's channel count padded to the K chunk.
struct PackedW {
  void* w = nullptr;
  float* bias = nullptr;     // [N] zero padded, may be null
  int N = 0, n_valid = 0, K_total = 0, k_real = 0, n_taps = 0, n_phases = 1;
  int ntile = 0, kc = 0;
  CUtensorMap tm{};          // valid in tensor-core modes
  // split-precision fp32 mode: [n_phases * N][2 * K_total] bf16 = (hi | lo) halves of every row, and its tensor map
  void* w3 = nullptr;
  CUtensorMap tm3{};
  // tap table (source row offset, channel base in the source view) per phase
  int tap_row[kMaxPhases][kMaxTaps] = {};
  int tap_ch[kMaxPhases][kMaxTaps] = {};
  int o_mul = 1;             // output rows per GEMM row (ConvTranspose stride)
};

struct Src {
  const void* p;
  int ld;       // row pitch of the view, elements
  int ch;       // channels consumed (padded)
  int rows;     // rows per batch element of the view
  const void* lo = nullptr;   // split-precision fp32 mode: p / lo are the bf16 hi / lo halves a producer already wrote
};

// One enqueued operation of a program plus the bookkeeping the built-in profiler needs.
struct Launch {
  std::function<cudaError_t(cudaStream_t)> fn;
  int kind = 2;     // 0 conv-GEMM, 1 attention, 2 elementwise / small kernels, 3 memset / memcpy (not a kernel)
  int region = 0;   // 0 estimator, 1..4 vocoder stage (conv_pre counted in 1), 5 boundary
  double flops = 0; // algorithmic FLOPs (2 per MAC, real channels only)
  Launch() = default;
  template <typename F, typename = std::enable_if_t<!std::is_same<std::decay_t<F>, Launch>::value>>
  Launch(F&& f) : fn(std::forward<F>(f)) {}
  cudaError_t operator()(cudaStream_t s) const { return fn(s); }
};
constexpr int kProfRegions = 6, kProfKinds = 4;
// Cached launch programs (one CUDA graph per shape) a plan keeps per cache: serving sees a bounded set of T_pad buckets,
// and an unbounded map of instantiated graphs is a slow leak when it does not.
constexpr size_t kMaxCachedPrograms = 32;
struct Program {
  std::vector<Launch> ops;
  cudaGraphExec_t graph = nullptr;
  long long kernels = 0;  // kernel launches per run (memsets not counted)
  unsigned long long last_use = 0;   // plan-wide use counter at the last run: the program caches are bounded LRUs
  std::vector<float> host_f32;  // small per-program host constants (Euler t schedule)
};

struct ResnetW {
  PackedW conv1, conv2, res;
  float *gn1_g, *gn1_b, *gn2_g, *gn2_b;
};
struct TransformerW {
  PackedW qkv, out, ff1, ff2;
  float *ln1_g, *ln1_b, *ln3_g, *ln3_b, *snake_a, *snake_invb;
};
struct EstimatorW {
  bool loaded = false;
  float *t_w1, *t_b1, *t_w2, *t_b2;   // time MLP fp32
  float *t_wp, *t_bp;                 // six mlp.1 projections stacked [6*256][1024], [6*256]
  ResnetW res[6];                     // down0, down1, mid0, mid1, up0, up1
  TransformerW tr[6];
  PackedW down0, down1, up0, up1, final_conv, final_proj;
  float *fgn_g, *fgn_b;
};
struct GeneratorW {
  bool loaded = false;
  PackedW conv_pre, ups[4], c1[12][3], c2[12][3];
  float* post_w = nullptr;  // [7][32] fp32
  float post_b = 0.f;
};

// Text encoder + duration predictor (text_encoder.cu); fp32 plans only.
struct TextEncW {
  bool loaded = false;
  int n_vocab = 0, C0 = 0, C = 0, H = 0, Dh = 0, n_layers = 0, filt = 0, filt_dp = 0, spk_dim = 0, n_feats = 80;
  float* emb = nullptr;
  PackedW pre_conv[3], pre_proj;
  float *pre_g[3], *pre_b[3];
  struct Layer {
    PackedW qkv, o, ffn1, ffn2;
    float *n1_g, *n1_b, *n2_g, *n2_b;
  } layer[8];
  PackedW proj_m, dp_conv1, dp_conv2, dp_proj;
  float *dp_g1, *dp_b1, *dp_g2, *dp_b2;
  float* theta = nullptr;   // [Dh / 4] rotary frequencies
};

class Plan {
 public:
  explicit Plan(const mb200_cfg& cfg);
  ~Plan();
  void load_estimator(const StateDict& sd);
  void load_generator(const StateDict& sd);

  void estimator_forward(const float* x, const float* mask, const float* mu, const float* t, bool batched,
                         const float* spks, int B, int T, float* out, cudaStream_t st);
  void cfm_decode(const float* mu, const float* mask, const float* z, const float* spks, int B, int T, int n_steps,
                  float* out, cudaStream_t st);
  void hifigan_forward(const float* mel, int Tsrc, float scale, float shift, int B, int L, float* wav, cudaStream_t st);
  void synthesise(const float* mu, const float* mask, const float* z, const float* spks, int B, int T, int n_steps, int L,
                  float mel_mean, float mel_std, float* dec_out, float* mel_out, float* wav_out, cudaStream_t st);
  void synthesise_host(const float* mu, const float* mask, const float* z, const float* spks, int B, int T, int n_steps,
                       int L, float mel_mean, float mel_std, float* mel_out, float* wav_out);
  // The same call split in two for double buffering: submit() queues upload, hot path and read-backs of one batch on the
  // plan's private streams and returns; wait() blocks until that batch's outputs are in host memory.  Two slots: the upload
  // of batch i+1 and the read-back of batch i overlap the compute of the other.  Host buffers must stay valid until wait().
  void synthesise_host_submit(const float* mu, const float* mask, const float* z, const float* spks, int B, int T, int n_steps,
                              int L, float mel_mean, float mel_std, float* mel_out, float* wav_out, int slot);
  void synthesise_host_wait(int slot);
  // BASECFM.compute_loss forward (+ backward) on the fp32 training kernels (train.cu).  params: name -> (parameter, gradient)
  // device pointers in the reference layout; gradients are ACCUMULATED (+=) into non-null gradient pointers.
  void compute_loss_grad(const std::map<std::string, std::pair<const float*, float*>>& params, const float* x1,
                         const float* mask, const float* mu, const float* z, const float* t, const float* spks, int B, int T,
                         float sigma_min, float* loss_out, float* y_out, float* v_out, float* dmu, float* dspks, bool backward,
                         cudaStream_t st);
  // TextEncoder.forward (text_encoder.py:378-410): ids (B, Lx) int64, lengths (B,) int64, spks (B, spk_dim) or null ->
  // mu_x (B, 80, Lx), logw (B, 1, Lx), x_mask (B, 1, Lx); all device pointers.  fp32 plans only.
  void load_text_encoder(const StateDict& sd, int n_heads);
  void text_encoder_forward(const long long* ids, const long long* lengths, const float* spks, int B, int Lx, float* mu_x,
                            float* logw, float* x_mask, cudaStream_t st);
  // Dropout of the transformer blocks for the following compute_loss_grad calls (0 = eval mode, the default)
  void set_train_dropout(float p, unsigned long long seed) {
    if (!(p >= 0.f && p < 1.f)) throw Error("dropout rate must be in [0, 1)");
    train_dropout_p_ = p; train_dropout_seed_ = seed;
  }
  // HiFi-GAN bias denoiser (hifigan/denoiser.py:53-64): bias = |STFT(audio)|[:, 0] (513 bins); out (B, 256*(L/256))
  void denoiser_bias(const float* audio, int L, float* bias, cudaStream_t st);
  void denoise(const float* audio, int B, int L, const float* bias, float strength, float* out, cudaStream_t st);
  void test_conv1d(const float* x, const float* w, const float* bias, int B, int Cin, int Cout, int T, int K, int dil,
                   int stride, bool transposed, float* y, cudaStream_t st);
  void test_attention(const float* qkv, const float* mask, int B, int T, float* out, cudaStream_t st);

  long long launches = 0;
  long long bytes_allocated = 0;
  // built-in profiler: when on, programs run eagerly with an event pair around every op
  // fp32 plans: convolutions as three bf16 tcgen05 products (hi/lo split operands, fp32 accumulation and epilogue)
  // instead of the FFMA kernel.  MB200_FP32=simt keeps the FFMA kernel (A/B and debugging).
  bool fp32_split = true;
  bool ff_pair = true;      // CTA pairs in the fused post-attention kernel (MB200_FF_PAIR=0 turns them off)
  bool res_pair = true;     // CTA pairs in the fused ResBlock kernel (MB200_RES_PAIR=0 turns them off)
  bool rq_pair = false;     // CTA pairs in the fused res_conv + GroupNorm + LayerNorm + QKV kernel (MB200_RQ_PAIR=1)
  bool pair_resident = false;   // CTA pairs also for long layers with resident weights (MB200_PAIR_RESIDENT=1)
  int use_pair = 2;         // CTA pairs (cta_group::2) in the persistent conv-GEMM: 1 = N = 256 layers, 2 = also N = 128 (HiFi-GAN stage 2); MB200_PAIR=0 turns them off
  int kernel_version = 2;   // 2 = persistent conv_gemm2.cu (default), 1 = first non-persistent kernel (MB200_KERNEL=1)
  long long* probe_out = nullptr;   // developer probe (MB200_DBG): cycle counters of the last conv_op built
  int num_sms = 148;
  int tile_policy = 0;      // 1: 256-row x 128-channel tiles for wide layers (halves the weight stream; measured neutral at cfg2); 0: 128 x 256 (MB200_TILE_POLICY)
  int use_pdl = 0;          // programmatic dependent launch for the persistent conv kernel (MB200_PDL=1 enables; measured neutral inside CUDA graphs)
  int fuse_ff = 2;          // transformer feed-forward as one kernel, hidden activation kept on chip; 2 = also the attention
                            // out-projection + residual + LayerNorm as its pre-stage (MB200_FUSE_FF=1: FF only, 0: conv launches)
  int fuse_rq = 1;          // res_conv + GroupNorm/Mish + LayerNorm1 + QKV as one kernel (MB200_FUSE_RQ=0: two conv launches)
  int attn_umma = 1;        // tcgen05 attention (attention_umma.cu); MB200_ATTN=mma selects the mma.sync kernel for A/B runs
  int fuse_pairs = 1;       // HiFi-GAN stages 3-4: one kernel per ResBlock step (MB200_FUSE_PAIRS=0: two conv launches)
  bool profiling = false;
  double prof[kProfRegions][kProfKinds][3] = {};   // [region][kind] = {launches, milliseconds, flops}
  mb200_cfg cfg;

  // ---- helpers shared by the translation units
  bool tensor_core() const { return cfg.dtype != MB200_F32 && !cfg.force_simt; }
  void* dalloc(size_t bytes);                       // plan-owned, zero-initialised
  float* upload_f32(const float* host, size_t n);   // plan-owned
  void* ensure(DevBuf& b, size_t bytes);            // grow-only workspace
  void* upload_operand(const std::vector<float>& host);  // fp32 host -> plan dtype device
  CUtensorMap tmap_act(const void* p, int ld, int rows, int B, int kc, int box_rows = 128, int dtype = -1) const;   // dtype -1: the plan's
  CUtensorMap tmap_w(const void* p, int K_total, int rows, int kc, int ntile, int dtype = -1) const;
  bool split_mode() const { return cfg.dtype == MB200_F32 && !cfg.force_simt && fp32_split; }
  void reserve_split(size_t elems);                 // scratch for the hi / lo operand copies of the largest conv input
  // Generic weight repack. weight(n, c, j) with c the concatenated real input channel; segs = (real, padded) channels of
  // each source; taps_j[phase][tap] = kernel index j.
  PackedW pack_conv(const std::function<float(int, int, int)>& weight, const float* bias, int n_real, int n_pad,
                    const std::vector<std::pair<int, int>>& segs, const std::vector<std::vector<int>>& taps_j);
  PackedW pack_conv1d(const HostTensor& w, const float* bias, int dil, const std::vector<std::pair<int, int>>& segs,
                      int n_pad = 0, float row_scale_first = 1.f, int scaled_rows = 0);
  PackedW pack_conv_stride2(const HostTensor& w, const float* bias, int C);
  PackedW pack_convT(const HostTensor& w, const float* bias, int stride);
  Launch conv_op(const PackedW& W, Src s0, const Src* s1, int B, int T_out, EpiParams e);
  // x + c2(lrelu(c1(in))) as one kernel (resblock_pair.cu); `in` is [B][T][C] dense, e carries c2's residual / outputs
  Launch pair_op(const PackedW& W1, const PackedW& W2, int dil, const void* in, int C, int B, int T, EpiParams e);
  // (SnakeBeta(x W1^T + b1) W2^T + b2 + resid) * mask as one kernel (ff_fused.cu); x, resid, out are [B*L][256] dense
  // With Wo != nullptr the attention out-projection, its residual and LayerNorm3 run as the kernel's pre-stage:
  // x is then the attention output [B*L][128], res_in the block input, resid (written here) the post-attention hidden.
  Launch ff_op(const PackedW& W1, const PackedW& W2, const float* snake_a, const float* snake_invb, const void* x,
               void* resid, const float* mask, void* out, long rows, const PackedW* Wo = nullptr,
               const void* res_in = nullptr, const float* ln_g = nullptr, const float* ln_b = nullptr);
  // res = x Wres^T + b + mish(GN(raw2)) * mask ; qkv = LayerNorm1(res) Wqkv^T as one kernel (resqkv_fused.cu);
  // sources are dense [B][L][ld] buffers tiled over the flat B*L rows
  Launch resqkv_op(const PackedW& Wres, const PackedW& Wqkv, Src s0, const Src* s1, const void* gn_raw, const EpiParams& gn_producer,
                   const float* gn_gamma, const float* gn_beta, const float* mask, void* res_out, const float* ln_g,
                   const float* ln_b, void* qkv_out, int B, int L);
  // softmax(q k^T + mask) v on [B][L][384] qkv -> [B][L][128] (operand dtype; fp32 mode: SIMT kernel)
  Launch attention_op(const void* qkv, const float* mask, void* out, int B, int L);
  bool rq_supported() const { return fuse_rq && tensor_core() && kernel_version >= 2; }
  bool ff_supported() const { return fuse_ff && tensor_core() && kernel_version >= 2; }
  bool pair_supported(int C) const { return fuse_pairs && tensor_core() && kernel_version >= 2 && (C == 64 || C == 32); }
  void run(Program& prog, cudaStream_t st);
  // makes room for one more entry: evicts the least recently used program (its graph is freed once it has drained)
  template <class Map>
  void make_room(Map& cache) {
    while (cache.size() >= kMaxCachedPrograms) {
      auto victim = cache.begin();
      for (auto it = cache.begin(); it != cache.end(); ++it)
        if (it->second->last_use < victim->second->last_use) victim = it;
      if (victim->second->graph) cudaGraphExecDestroy(victim->second->graph);
      cache.erase(victim);
    }
  }
  size_t cached_programs() const { return decode_cache_.size() + vocoder_cache_.size() + encoder_cache_.size(); }

 private:
  std::vector<void*> owned_;
  void* train_arena_ = nullptr;       // activations saved for the backward pass + gradient scratch (train.cu)
  size_t train_arena_bytes_ = 0;
  std::recursive_mutex mu_;
  EstimatorW est_;
  GeneratorW gen_;
  TextEncW enc_;
  DevBuf ws_enc_[16];
  // workspaces
  DevBuf ws_est_[28];
  DevBuf ws_voc_[8];
  DevBuf ws_split_[2];      // split-precision fp32 mode: bf16 hi / lo copies of the current conv's input
  DevBuf ws_io_[8];
  unsigned long long use_clock_ = 0;
  float train_dropout_p_ = 0.f;
  unsigned long long train_dropout_seed_ = 0;
  cudaStream_t host_stream_ = nullptr;   // synthesise_host: private non-blocking compute stream (never the legacy default stream)
  cudaEvent_t busy_ = nullptr;           // recorded after every asynchronous call on the caller's stream: the workspaces'
                                         // last user, which the private stream waits for
  void mark_busy(cudaStream_t st);
  struct HostSlot {                      // staging of one in-flight synthesise_host_submit
    DevBuf mu, z, mask, spk, mel, wav;
    cudaEvent_t uploaded = nullptr, computed = nullptr, copied = nullptr;
    bool busy = false;
  } slots_[2];
  cudaStream_t upload_stream_ = nullptr; // host -> device copies of the next batch, overlapping the current batch's compute
  cudaStream_t copy_stream_ = nullptr;   // synthesise_host: mel read-back overlapping the vocoder
  cudaEvent_t mel_ready_ = nullptr;
  cudaEvent_t wav_ready_[4] = {nullptr, nullptr, nullptr, nullptr};
  float* host_wav_dst_ = nullptr;        // set only while synthesise_host runs the vocoder
  std::map<std::tuple<int, int, int, int, long long>, std::unique_ptr<Program>> decode_cache_;   // (B,T,steps,spk,ptr-key)
  std::map<std::tuple<int, int>, std::unique_ptr<Program>> vocoder_cache_;
  std::map<std::tuple<int, int>, std::unique_ptr<Program>> encoder_cache_;   // text encoder programs by (B, Lx)

  void build_estimator_step(Program& prog, int B, int T, const float* tproj, int tproj_bstride, bool euler, float dt,
                            float* v_out_ntc);
  void build_time_embed(Program& prog, const float* t_dev, int n_t);
  void build_vocoder(Program& prog, int B, int L, float* wav);
  void estimator_buffers(int B, int T);
  void vocoder_buffers(int B, int L);
  friend struct PlanAccess;
};

// Length regulation (matcha_tts.py:118-131, utils/model.py:29-41), length_regulate.cu.  Plan-independent.
void length_regulate(const float* mu_x, const float* w, const float* x_mask, const long long* y_len, int B, int C, int Lx,
                     int T, float* cum, float* mu_y, float* attn, float* y_mask, cudaStream_t st);

// Monotonic alignment search (monotonic_align/core.pyx:11-47), mas.cu.  Plan-independent.
size_t maximum_path_workspace_bytes(int B, int Tx, int Ty);
void maximum_path(const float* value, const float* mask, int B, int Tx, int Ty, float* path, void* workspace, cudaStream_t st);
// Same, scratch in a file-scope __device__ array of MB200_MAS_STATIC_SCRATCH_BYTES (mas.cu): allocates nothing.
void maximum_path_static(const float* value, const float* mask, int B, int Tx, int Ty, float* path, cudaStream_t st);

}  // namespace mb
