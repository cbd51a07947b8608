/* matcha_b200 -- C ABI of the B200-native Matcha-TTS synthesis hot path.
 *
 * The reference (Mocifiel/Matcha, a pure-PyTorch project) has no FFI layer; its boundary for this path is the Python
 * class API.  Each entry point below names the reference interface it stands behind (path:line under the reference
 * tree).  All tensor arguments are plain pointers in the REFERENCE layouts: (B, C, T) contiguous fp32; any internal
 * relayout ([B][T][C], bf16/f16 operands) is private.  Unless stated otherwise pointers are DEVICE pointers and work is
 * enqueued on `stream` (a cudaStream_t passed as void*).  Every function returns 0 on success; otherwise
 * mb200_last_error() describes the failure (thread-local string).  There is no CPU fallback.
 */
#ifndef MATCHA_B200_H_
#define MATCHA_B200_H_

#ifdef __cplusplus
extern "C" {
#endif

typedef struct mb200_plan mb200_plan;

enum { MB200_F32 = 0, MB200_BF16 = 1, MB200_F16 = 2 };

typedef struct {
  int dtype;       /* MB200_F32: fp32 activations and epilogues; convolutions as split-precision tcgen05 products (bf16 hi/lo
                      operands, ~2^-16 relative per term, fp32 accumulation), attention / norms in fp32 FFMA.
                      MB200_BF16 / MB200_F16: tcgen05 tensor-core kernels, bf16 / f16 operands and activations, fp32
                      accumulation and statistics. */
  int c_in;        /* estimator input channels: 160 (LJSpeech) or 224 (VCTK: + 64 speaker dims), flow_matching.py:130 */
  int force_simt;  /* debugging: run every mode's convolutions on the FFMA kernel instead of tcgen05 */
  int use_graph;   /* capture decode / vocoder launches into a CUDA graph per (B, T, n_steps) shape */
} mb200_cfg;

/* One parameter tensor of a reference state_dict: HOST fp32, contiguous. */
typedef struct {
  const char* name;
  const float* data;
  int ndim;
  long long shape[4];
} mb200_tensor;

const char* mb200_last_error(void);
const char* mb200_version(void);

int mb200_plan_create(const mb200_cfg* cfg, mb200_plan** out);
void mb200_plan_destroy(mb200_plan* plan);

/* Decoder.__init__ parameter tree (matcha/models/components/decoder.py:201-316); `prefix` is stripped from names
 * (e.g. "estimator." or "decoder.estimator.").  Repacks to the kernels' operand layouts and uploads. */
int mb200_load_estimator(mb200_plan* plan, const mb200_tensor* tensors, int n, const char* prefix);
/* HiFi-GAN Generator parameters (matcha/hifigan/models.py:149-179), either weight_g/weight_v (checkpoint layout,
 * matcha/cli.py:87) or plain weight (after remove_weight_norm, models.py:199-206). */
int mb200_load_generator(mb200_plan* plan, const mb200_tensor* tensors, int n);

/* TextEncoder parameters (matcha/models/components/text_encoder.py:328-376: emb, prenet.*, encoder.*, proj_m, proj_w.*);
 * `prefix` ("encoder." in a MatchaTTS state_dict) is stripped.  n_heads = encoder_params.n_heads (2 in the shipped
 * configs; head dims 96 = LJSpeech and 128 = VCTK have kernels).  fp32 plans only: see mb200_text_encoder_forward. */
int mb200_load_text_encoder(mb200_plan* plan, const mb200_tensor* tensors, int n, const char* prefix, int n_heads);
/* TextEncoder.forward(x, x_lengths, spks) (text_encoder.py:378-410) including the DurationPredictor (:70-94).  DEVICE
 * pointers: ids (B, Lx) int64 phoneme ids, lengths (B,) int64, spks (B, spk_emb_dim) fp32 speaker vectors or NULL for a
 * single-speaker model; outputs mu_x (B, 80, Lx), logw (B, 1, Lx), x_mask (B, 1, Lx), fp32.  Runs in fp32 whatever the
 * decoder's mode: the durations go through ceil(exp(logw)) (matcha_tts.py:118-119).  Inference only (no dropout). */
int mb200_text_encoder_forward(mb200_plan* plan, const long long* ids, const long long* lengths, const float* spks, int B, int Lx,
                               float* mu_x, float* logw, float* x_mask, void* stream);

/* The (t, dt) pairs BASECFM.solve_euler uses (matcha/models/components/flow_matching.py:52,70-83), fp32. Host only. */
int mb200_euler_schedule(int n_steps, float* t_out, float* dt_out);

/* Decoder.forward(x, mask, mu, t, spks) -> velocity (B,80,T)   (decoder.py:363-443).
 * t: device fp32, 1 value (t_is_batched = 0) or B values (t_is_batched = 1, the compute_loss case). spks: (B, c_in-160) or NULL. */
int mb200_estimator_forward(mb200_plan* plan, const float* x, const float* mask, const float* mu, const float* t,
                            int t_is_batched, const float* spks, int B, int T, float* out, void* stream);

/* BASECFM.solve_euler(x=z, t_span=linspace(0,1,n_steps+1), mu, mask, spks) -> (B,80,T)   (flow_matching.py:55-85).
 * z is the already temperature-scaled noise (flow_matching.py:51). */
int mb200_cfm_decode(mb200_plan* plan, const float* mu, const float* mask, const float* z, const float* spks, int B,
                     int T, int n_steps, float* out, void* stream);

/* Generator.forward(mel (B,80,L)) -> wav (B,1,256*L)   (matcha/hifigan/models.py:181-197). */
int mb200_hifigan_forward(mb200_plan* plan, const float* mel, int B, int L, float* wav, void* stream);

/* The whole hot path of MatchaTTS.synthesise + cli.to_waveform: decode, crop to L frames, denormalise
 * (matcha_tts.py:135-145, utils/model.py:71-90), vocode, clamp (cli.py:117-118).
 * dec_out (B,80,T), mel_out (B,80,L), wav_out (B,1,256*L); any of the three may be NULL except wav_out. */
int mb200_synthesise(mb200_plan* plan, const float* mu, const float* mask, const float* z, const float* spks, int B,
                     int T, int n_steps, int L, float mel_mean, float mel_std, float* dec_out, float* mel_out,
                     float* wav_out, void* stream);

/* Same, with HOST buffers (pinned or pageable): copies inputs in, runs, copies mel_out / wav_out back and synchronises.
 * This is the end-to-end call the benchmark's `e2e` figure times. */
int mb200_synthesise_host(mb200_plan* plan, const float* mu, const float* mask, const float* z, const float* spks, int B,
                          int T, int n_steps, int L, float mel_mean, float mel_std, float* mel_out, float* wav_out);

/* mb200_synthesise_host split in two, for double buffering by a serving loop (no counterpart in the reference, whose
 * synthesise is one synchronous call): submit queues the upload, the hot path and the read-backs of one batch on the
 * plan's private streams and returns at once; wait blocks until that batch's mel_out / wav_out are complete in host
 * memory.  slot is 0 or 1: with two batches in flight the upload of batch i+1 and the read-back of batch i overlap the
 * other's compute.  Host buffers must stay valid (and untouched) until wait, and must be pinned for the copies to overlap.
 * A slot must be waited for before it is submitted again.  mb200_synthesise_host = submit(slot 0) + wait(0). */
int mb200_synthesise_host_submit(mb200_plan* plan, const float* mu, const float* mask, const float* z, const float* spks, int B,
                                 int T, int n_steps, int L, float mel_mean, float mel_std, float* mel_out, float* wav_out,
                                 int slot);
int mb200_synthesise_host_wait(mb200_plan* plan, int slot);

/* Denoiser.__init__ (matcha/hifigan/denoiser.py:53-57): bias (513,) = |STFT(bias_audio (L,))|[:, frame 0] with
 * n_fft = win = 1024, hop 256, periodic Hann, center / reflect padding (denoiser.py:26-44).  L > 512. */
int mb200_denoiser_bias(mb200_plan* plan, const float* bias_audio, int L, float* bias, void* stream);
/* Denoiser.forward(audio (B, L), strength) (denoiser.py:59-64; called on the vocoder output by cli.py:119-120):
 * STFT -> max(|S| - strength * bias, 0) with the phase kept -> iSTFT.  out (B, 256 * (L / 256)).  L > 512. */
int mb200_denoise(mb200_plan* plan, const float* audio, int B, int L, const float* bias, float strength, float* out,
                  void* stream);

/* Length regulation on the synthesis path (matcha/models/matcha_tts.py:118-131): replaces
 * `attn = generate_path(w_ceil, x_mask (x) y_mask)` (matcha/utils/model.py:29-41) and the dense
 * `mu_y = matmul(attn^T, mu_x^T)^T` by a prefix sum + gather.  All DEVICE pointers, fp32, reference layouts:
 * mu_x (B, C, Lx), w (B, Lx) frames per token (w_ceil * length_scale), x_mask (B, Lx) 0/1, y_lengths (B,) int64,
 * outputs mu_y (B, C, T), attn (B, Lx, T) (may be NULL), y_mask (B, T) (may be NULL); cum (B, Lx) is workspace that
 * returns the fp32 prefix sums.  Frames at or past y_lengths[b] and frames no token owns are zero.  No plan needed. */
int mb200_length_regulate(const float* mu_x, const float* w, const float* x_mask, const long long* y_lengths, int B, int C,
                          int Lx, int T, float* cum, float* mu_y, float* attn, float* y_mask, void* stream);

/* Monotonic alignment search: `maximum_path(value, mask)` of matcha/utils/monotonic_align/__init__.py:7-22, i.e. the
 * Cython `maximum_path_c` (monotonic_align/core.pyx:11-47) without the device -> host -> device round trip.  DEVICE
 * pointers, fp32: value (B, Tx, Ty) log-prior, mask (B, Tx, Ty) 0/1 (token count = nonzeros of mask[b, :, 0], frame count
 * = nonzeros of mask[b, 0, :], as the reference derives them); path (B, Tx, Ty) receives 0/1.  workspace: at least
 * mb200_maximum_path_workspace_bytes(B, Tx, Ty) bytes.  Defined for 1 <= tokens <= frames per utterance. */
long long mb200_maximum_path_workspace_bytes(int B, int Tx, int Ty);
int mb200_maximum_path(const float* value, const float* mask, int B, int Tx, int Ty, float* path, void* workspace,
                       void* stream);

/* Allocation-free variant of mb200_maximum_path.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that
 * forbid it).  Same kernel, inputs, outputs and results; the scratch (B * Ty * ceil(Tx / 32) * 4 bytes of decision bits)
 * lives in a file-scope __device__ array of MB200_MAS_STATIC_SCRATCH_BYTES inside the library, which every GPU that loads
 * the library holds until the process exits.  Shapes that need more are rejected with an error (use mb200_maximum_path).
 * The call performs no cudaMalloc / cudaFree and calls no allocating library routine.  The array is shared by all calls
 * on a device and keeps the last call's contents (nothing needs clearing: every word read is written earlier in the same
 * launch); calls that overlap on one GPU race on it - serialise them on one stream.  The first call on a device resolves
 * the array's device address (cudaGetSymbolAddress) and configures the kernel: make it outside stream capture. */
#define MB200_MAS_STATIC_SCRATCH_BYTES (4u * 1024u * 1024u)
int mb200_maximum_path_static(const float* value, const float* mask, int B, int Tx, int Ty, float* path, void* stream);

/* One estimator parameter for the training entry point: DEVICE fp32 tensors in the reference layout (SURVEY.md
 * Appendix A), `grad` may be NULL (that gradient is then not computed). */
typedef struct {
  const char* name;
  const float* param;
  float* grad;
} mb200_param;

/* Dropout of the estimator's transformer blocks for the mb200_compute_loss calls that follow (decoder.py:206 ->
 * transformer.py:124 after the feed-forward activation, :200 -> Attention.to_out[1] after the attention out-projection;
 * rate 0.05 in configs/model/decoder/default.yaml:2).  p = 0 (the default) is eval mode.  The keep mask of element i at site
 * s = 2 * block + {0 attention, 1 feed-forward} (blocks in the order down 0, down 1, mid 0, mid 1, up 0, up 1; i indexes
 * the (B, L, C) activation) is splitmix64(seed + 0x632BE59BD9B4E019 (s + 1) + i) >> 32 >= p * 2^32: reproducible from
 * the seed, regenerated in the backward pass, and injectable into a reference run (tests/test_train_gpu.py).  It is not
 * torch's Philox stream: draws differ from the reference's, their distribution does not. */
int mb200_set_train_dropout(mb200_plan* plan, float p, unsigned long long seed);

/* BASECFM.compute_loss(x1, mask, mu, spks) (matcha/models/components/flow_matching.py:87-118) with the random draws
 * injected: t (B,) ~ U(0,1) and z (B,80,T) ~ N(0,1) are device inputs.  Computes y = (1-(1-sigma_min) t) z + t x1,
 * u = x1 - (1-sigma_min) z, v = Decoder.forward(y, mask, mu, t, spks) (decoder.py:363-443, per-sample t) and
 * loss = sum((v-u)^2) / (sum(mask) * 80).  With backward != 0 the whole estimator backward pass runs as well:
 * d loss / d parameter is ACCUMULATED (+=) into every non-NULL `grad` (zero-fill them for plain gradients),
 * d loss / d mu -> dmu (B,80,T) and d loss / d spks -> dspks (B, c_in-160), both optional (NULL = skip).
 * Parameters are read in place (no repacking between optimiser steps).  fp32 data in every plan dtype; the GEMMs are error-compensated 3xTF32 tensor-core kernels (mma.sync, fp32-level accuracy; csrc/train_kernels.cuh), the element-wise kernels fp32.
 * loss: device scalar; y, v: optional (B,80,T) outputs.  `prefix` is stripped from parameter names. */
int mb200_compute_loss(mb200_plan* plan, const mb200_param* params, int n, const char* prefix, const float* x1,
                       const float* mask, const float* mu, const float* z, const float* t, const float* spks, int B, int T,
                       float sigma_min, int backward, float* loss, float* y, float* v, float* dmu, float* dspks,
                       void* stream);

/* Bookkeeping for the benchmark: kernels launched (or replayed inside graphs) by this plan since creation. */
long long mb200_launch_count(const mb200_plan* plan);
/* Device-memory footprint of the plan's workspaces and weights, bytes. */
long long mb200_workspace_bytes(const mb200_plan* plan);

/* Built-in profiler.  While on, launch programs run eagerly (no CUDA graph) with a CUDA-event pair around every
 * operation on the launching stream.  mb200_get_profile fills out[6][4][3]: [region][kind] = {launches, milliseconds,
 * algorithmic FLOPs}; region 0 = estimator, 1..4 = vocoder stages (conv_pre in 1), 5 = other; kind 0 = conv-GEMM,
 * 1 = attention, 2 = elementwise / small kernels, 3 = memset / memcpy.  Turning it on or off zeroes the counters. */
int mb200_set_profiling(mb200_plan* plan, int on);
int mb200_get_profile(const mb200_plan* plan, double* out);

/* Developer hook: copies the cycle counters of the last instrumented kernel (MB200_DBG / MB200_PAIR_DBG / MB200_FF_DBG
 * environment switches) to HOST memory: out[n_ctas * 16] for the first n_ctas CTAs.  Fails if nothing was instrumented. */
int mb200_debug_probe(mb200_plan* plan, long long* out, int n_ctas);

/* Unit-test hook: one Conv1d / ConvTranspose1d (+ bias) through the conv-GEMM kernel of the plan's mode.
 * x (B,Cin,T) fp32 device, w reference layout HOST fp32 ((Cout,Cin,K) or (Cin,Cout,K) when transposed), bias HOST (Cout)
 * or NULL, y (B,Cout,T_out) fp32 device.  stride applies to Conv1d (1|2) or ConvTranspose1d (K == 2*stride). */
int mb200_test_conv1d(mb200_plan* plan, const float* x, const float* w, const float* bias, int B, int Cin, int Cout, int T,
                      int K, int dilation, int stride, int transposed, float* y, void* stream);
/* Unit-test hook: masked-additive self-attention, qkv (B,T,384) fp32 device (q pre-scaled), mask (B,T) -> out (B,T,128). */
int mb200_test_attention(mb200_plan* plan, const float* qkv, const float* mask, int B, int T, float* out, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* MATCHA_B200_H_ */
