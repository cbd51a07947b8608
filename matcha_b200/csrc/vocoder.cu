// HiFi-GAN v1 generator (matcha/hifigan/models.py:149-197) as a launch program: every Conv1d / ConvTranspose1d is
// a conv-GEMM launch; LeakyReLU, residual adds, the 3-branch mean and the stage hand-over live in the epilogues.
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <vector>
#include <cuda_fp16.h>

#include "plan.h"

namespace mb {

namespace {
enum VocBuf { V_MEL, V_A, V_X, V_XL, V_R, V_RL, V_C1, V_SUM };
const int kRates[4] = {8, 8, 2, 2};
const int kKernels[3] = {3, 7, 11};
const int kDils[3] = {1, 3, 5};

const HostTensor& need(const StateDict& sd, const std::string& k) {
  auto it = sd.find(k);
  if (it == sd.end()) throw Error("generator state_dict is missing '" + k + "'");
  return it->second;
}
// Effective weight for either key layout: plain `weight`, or weight_norm's g * v / ||v|| with the norm over all dims
// but 0 (old-style torch.nn.utils.weight_norm, dim=0; matcha/hifigan/models.py:20-87,154-177,199-206).
HostTensor effective_weight(const StateDict& sd, const std::string& name, std::vector<std::vector<float>>& keep) {
  auto it = sd.find(name + ".weight");
  if (it != sd.end()) return it->second;
  const HostTensor& g = need(sd, name + ".weight_g");
  const HostTensor& v = need(sd, name + ".weight_v");
  const long long d0 = v.shape[0], inner = v.numel() / d0;
  if (g.numel() != d0) throw Error("weight_g of '" + name + "' does not match weight_v");
  keep.emplace_back(static_cast<size_t>(v.numel()));
  std::vector<float>& w = keep.back();
  for (long long i = 0; i < d0; ++i) {
    double ss = 0.0;
    for (long long j = 0; j < inner; ++j) { const double x = v.data[i * inner + j]; ss += x * x; }
    const float nrm = static_cast<float>(std::sqrt(ss));
    const float s = g.data[i] / nrm;
    for (long long j = 0; j < inner; ++j) w[i * inner + j] = v.data[i * inner + j] * s;
  }
  return HostTensor{w.data(), v.shape};
}
EpiParams epi() {
  EpiParams e{};
  e.scale = 1.f; e.out0_slope = 1.f; e.out1_slope = 1.f; e.resid_inv = 1.f; e.o_mul = 1;
  return e;
}
}  // namespace

void Plan::load_generator(const StateDict& sd) {
  std::lock_guard<std::recursive_mutex> lk(mu_);
  std::vector<std::vector<float>> keep;
  GeneratorW& G = gen_;
  {
    HostTensor w = effective_weight(sd, "conv_pre", keep);
    if (w.shape != std::vector<long long>{512, 80, 7}) throw Error("conv_pre.weight must be (512,80,7): only HiFi-GAN v1 is supported");
    G.conv_pre = pack_conv1d(w, need(sd, "conv_pre.bias").data, 1, {{80, 128}});
  }
  int ch = 512;
  for (int i = 0; i < 4; ++i) {
    HostTensor w = effective_weight(sd, "ups." + std::to_string(i), keep);
    if (w.shape != std::vector<long long>{ch, ch / 2, 2 * kRates[i]}) throw Error("ups." + std::to_string(i) + " has an unexpected shape");
    G.ups[i] = pack_convT(w, need(sd, "ups." + std::to_string(i) + ".bias").data, kRates[i]);
    ch /= 2;
    for (int j = 0; j < 3; ++j)
      for (int m = 0; m < 3; ++m) {
        const std::string p = "resblocks." + std::to_string(3 * i + j) + ".";
        HostTensor w1 = effective_weight(sd, p + "convs1." + std::to_string(m), keep);
        HostTensor w2 = effective_weight(sd, p + "convs2." + std::to_string(m), keep);
        if (w1.shape != std::vector<long long>{ch, ch, kKernels[j]}) throw Error(p + "convs1 has an unexpected shape");
        G.c1[3 * i + j][m] = pack_conv1d(w1, need(sd, p + "convs1." + std::to_string(m) + ".bias").data, kDils[m], {{ch, ch}});
        G.c2[3 * i + j][m] = pack_conv1d(w2, need(sd, p + "convs2." + std::to_string(m) + ".bias").data, 1, {{ch, ch}});
      }
  }
  {
    HostTensor w = effective_weight(sd, "conv_post", keep);   // (1, 32, 7) -> [7][32]
    std::vector<float> pw(7 * 32);
    for (int c = 0; c < 32; ++c)
      for (int j = 0; j < 7; ++j) pw[j * 32 + c] = w.data[c * 7 + j];
    G.post_w = upload_f32(pw.data(), pw.size());
    G.post_b = need(sd, "conv_post.bias").data[0];
  }
  G.loaded = true;
  vocoder_cache_.clear();
}

void Plan::vocoder_buffers(int B, int L) {
  const size_t es = esize(cfg.dtype);
  const size_t bl = static_cast<size_t>(B) * L;
  ensure(ws_voc_[V_MEL], bl * 128 * es);
  reserve_split(bl * 8192);            // every stage's activations are 8192 values per mel frame
  const bool single = tensor_core() && kernel_version >= 2 && cfg.dtype != MB200_F32;   // see build_vocoder
  for (int b : {V_A, V_X, V_XL, V_R, V_RL, V_C1, V_SUM})
    if (!(single && (b == V_X || b == V_R))) ensure(ws_voc_[b], bl * 8192 * es);
}

// Everything between the packed mel and the conv_post input; conv_post itself writes to the caller's buffer and is
// launched outside the cached program.
void Plan::build_vocoder(Program& prog, int B, int L, float* /*wav*/) {
  const int f32 = cfg.dtype == MB200_F32;
  // Tensor-core path: a ResBlock input is kept ONLY as lrelu(x) (what the convs read); the residual read inverts it
  // (x = min(r, r / slope), slope = the f16/bf16-rounded 0.1 the store used), so every conv writes one stream instead
  // of two.  The fp32 SIMT path keeps both x and lrelu(x).
  const bool single = tensor_core() && kernel_version >= 2 && !f32;
  float slope_stored = 0.1f;
  if (cfg.dtype == MB200_F16) slope_stored = __half2float(__float2half_rn(0.1f));
  const float inv_slope = 1.f / slope_stored;
  auto buf = [&](int i) { return ws_voc_[i].p; };
  // Split-precision fp32 mode: the streams that only convolutions read (A, XL, RL, C1) exist ONLY as bf16 hi / lo pairs
  // (the two halves of the fp32-sized buffer), written by the producing epilogue and read by the next conv as they are;
  // the residual streams (X, R, SUM) stay fp32.  No separate split pass, no fp32 copy of a conv input.
  const bool sp = split_mode();
  const size_t lo_off = static_cast<size_t>(B) * L * 8192 * 2;
  auto lo_of = [&](int i) { return static_cast<void*>(static_cast<char*>(buf(i)) + lo_off); };
  auto conv_in = [&](int i, int ld, int ch, int rows) { Src s{buf(i), ld, ch, rows}; if (sp) s.lo = lo_of(i); return s; };
  auto to_conv0 = [&](EpiParams& e, int i, int ld, float slope) {   // out0 = a conv-input stream
    e.out0 = buf(i); e.out0_ld = ld; e.out0_slope = slope;
    if (sp) e.out0_lo = lo_of(i); else e.out0_f32 = f32;
  };
  auto to_conv1 = [&](EpiParams& e, int i, int ld, float slope) {   // out1 = a conv-input stream
    e.out1 = buf(i); e.out1_ld = ld; e.out1_slope = slope;
    if (sp) e.out1_lo = lo_of(i);
  };
  int region = 1;
  auto add = [&](Launch l) { l.region = region; prog.ops.push_back(std::move(l)); prog.kernels++; };
  const GeneratorW& G = gen_;
  {  // conv_pre, then the first stage's leaky_relu(0.1) (models.py:182,184)
    EpiParams e = epi();
    to_conv0(e, V_A, 512, 0.1f);
    add(conv_op(G.conv_pre, Src{buf(V_MEL), 128, 128, L}, nullptr, B, L, e));
  }
  int ch = 512, len = L;
  for (int i = 0; i < 4; ++i) {
    const int C = ch / 2;
    region = i + 1;
    {  // ups[i]: x and leaky_relu(x) for the three ResBlocks (models.py:185, :92)
      EpiParams e = epi();
      if (single) {
        e.out0 = buf(V_XL); e.out0_ld = C; e.out0_slope = 0.1f;
      } else {
        e.out0 = buf(V_X); e.out0_ld = C; e.out0_f32 = f32;
        to_conv1(e, V_XL, C, 0.1f);
      }
      add(conv_op(G.ups[i], conv_in(V_A, ch, ch, len), nullptr, B, len, e));
    }
    len *= kRates[i];
    ch = C;
    for (int j = 0; j < 3; ++j) {
      for (int m = 0; m < 3; ++m) {
        const PackedW& W1 = G.c1[3 * i + j][m];
        const PackedW& W2 = G.c2[3 * i + j][m];
        const bool fused = single && pair_supported(C);
        if (fused) {
          // One kernel per ResBlock step (resblock_pair.cu).  Tiles read a halo of their input, so the step cannot run
          // in place: lrelu(x) ping-pongs XL -> RL -> C1 -> (branch sum | next stage input).
          const void* in = m == 0 ? buf(V_XL) : (m == 1 ? buf(V_RL) : buf(V_C1));
          EpiParams e = epi();
          e.resid = in; e.resid_ld = C; e.resid_inv = inv_slope;
          if (m < 2) {
            e.out0 = m == 0 ? buf(V_RL) : buf(V_C1); e.out0_ld = C; e.out0_slope = 0.1f;
          } else if (j == 0) {
            e.out0 = buf(V_SUM); e.out0_ld = C;
          } else {
            e.resid2 = buf(V_SUM); e.resid2_ld = C;
            if (j == 1) { e.out0 = buf(V_SUM); e.out0_ld = C; }
            else { e.scale = 1.f / 3.f; e.out0 = buf(V_A); e.out0_ld = C; e.out0_slope = i == 3 ? 0.01f : 0.1f; }
          }
          add(pair_op(W1, W2, kDils[m], in, C, B, len, e));
          continue;
        }
        {  // xt = lrelu(c1(lrelu(r)))   (models.py:92-94)
          EpiParams e = epi();
          to_conv0(e, V_C1, C, 0.1f);
          add(conv_op(W1, conv_in(m == 0 ? V_XL : V_RL, C, C, len), nullptr, B, len, e));
        }
        EpiParams e = epi();  // r = c2(xt) + r   (models.py:95-96)
        if (single) {
          e.resid = m == 0 ? buf(V_XL) : buf(V_RL); e.resid_ld = C; e.resid_inv = inv_slope;
        } else {
          e.resid = m == 0 ? buf(V_X) : buf(V_R); e.resid_ld = C; e.resid_f32 = f32;
        }
        if (m < 2 && single) {
          // in place: every thread reads exactly the residual elements it overwrites, and the conv input is V_C1
          e.out0 = buf(V_RL); e.out0_ld = C; e.out0_slope = 0.1f;
        } else if (m < 2) {
          e.out0 = buf(V_R); e.out0_ld = C; e.out0_f32 = f32;
          to_conv1(e, V_RL, C, 0.1f);
        } else if (j == 0) {
          e.out0 = buf(V_SUM); e.out0_ld = C; e.out0_f32 = f32;
        } else if (j == 1) {
          e.resid2 = buf(V_SUM); e.resid2_ld = C; e.resid2_f32 = f32;
          e.out0 = buf(V_SUM); e.out0_ld = C; e.out0_f32 = f32;
        } else {  // x = xs / 3, then the next consumer's leaky_relu: 0.1 inside the loop, 0.01 before conv_post (:192-193)
          e.resid2 = buf(V_SUM); e.resid2_ld = C; e.resid2_f32 = f32;
          e.scale = 1.f / 3.f;
          if (i == 3) { e.out0 = buf(V_A); e.out0_ld = C; e.out0_f32 = f32; e.out0_slope = 0.01f; }   // conv_post reads fp32
          else to_conv0(e, V_A, C, 0.1f);
        }
        add(conv_op(W2, conv_in(V_C1, C, C, len), nullptr, B, len, e));
      }
    }
  }
}

}  // namespace mb

namespace mb {

void Plan::hifigan_forward(const float* mel, int Tsrc, float scale, float shift, int B, int L, float* wav, cudaStream_t st) {
  std::lock_guard<std::recursive_mutex> lk(mu_);
  if (!gen_.loaded) throw Error("generator weights are not loaded");
  if (B < 1 || L < 1) throw Error("empty vocoder input");
  vocoder_buffers(B, L);
  const int dtype = cfg.dtype;
  void* melp = ws_voc_[V_MEL].p;
  if (dtype == MB200_F32) MB_CUDA(launch_pack_mel<float>(mel, Tsrc, scale, shift, (float*)melp, 128, B, L, st));
  else if (dtype == MB200_BF16) MB_CUDA(launch_pack_mel<__nv_bfloat16>(mel, Tsrc, scale, shift, (__nv_bfloat16*)melp, 128, B, L, st));
  else MB_CUDA(launch_pack_mel<__half>(mel, Tsrc, scale, shift, (__half*)melp, 128, B, L, st));
  auto key = std::make_tuple(B, L);
  auto it = vocoder_cache_.find(key);
  if (it == vocoder_cache_.end()) {
    auto prog = std::make_unique<Program>();
    build_vocoder(*prog, B, L, nullptr);
    make_room(vocoder_cache_);
    it = vocoder_cache_.emplace(key, std::move(prog)).first;
  }
  run(*it->second, st);
  if (probe_out && std::getenv("MB200_PAIR_DBG")) {   // developer probe: per-role cycle counters of the instrumented pair op
    MB_CUDA(cudaStreamSynchronize(st));
    std::vector<long long> h(148 * 16);
    MB_CUDA(cudaMemcpy(h.data(), probe_out, h.size() * sizeof(long long), cudaMemcpyDeviceToHost));
    static const char* names[15] = {"mma.total", "mma.w_t1_empty", "mma.w_a_full", "mma.w_w_full", "mma.w_t2_empty", "mma.w_mid_full",
                                    "e1.total", "e1.w_t1_full", "e1.w_mid_empty", "e2.total", "e2.w_t2_full", "prod.total",
                                    "prod.w_a_empty", "prod.w_w_empty", "tiles"};
    for (int k = 0; k < 15; ++k) {
      double s = 0; for (int c = 0; c < 148; ++c) s += h[c * 16 + k];
      std::fprintf(stderr, "[pair probe] %-16s %12.0f\n", names[k], s / 148);
    }
  }
  const long Lw = static_cast<long>(L) * 256;
  const char* a = static_cast<const char*>(ws_voc_[V_A].p);
  // conv_post + tanh.  On the host-buffer path (synthesise_host) it runs in four batch slices and each slice's waveform
  // starts its device -> host copy on the copy stream while the next slice is computed: the read-back (29.5 MB at cfg2)
  // is the longest single item outside the kernels.
  const int parts = (host_wav_dst_ && B >= 4) ? 4 : 1;
  for (int q = 0; q < parts; ++q) {
    const int b0 = static_cast<int>(static_cast<long>(B) * q / parts), b1 = static_cast<int>(static_cast<long>(B) * (q + 1) / parts);
    const char* aq = a + static_cast<size_t>(b0) * Lw * 32 * esize(dtype);
    float* wq = wav + static_cast<long>(b0) * Lw;
    if (dtype == MB200_F32) MB_CUDA(launch_conv_post<float>((const float*)aq, gen_.post_w, gen_.post_b, wq, b1 - b0, Lw, st));
    else if (dtype == MB200_BF16) MB_CUDA(launch_conv_post<__nv_bfloat16>((const __nv_bfloat16*)aq, gen_.post_w, gen_.post_b, wq, b1 - b0, Lw, st));
    else MB_CUDA(launch_conv_post<__half>((const __half*)aq, gen_.post_w, gen_.post_b, wq, b1 - b0, Lw, st));
    launches += 1;
    if (host_wav_dst_) {
      MB_CUDA(cudaEventRecord(wav_ready_[q], st));
      MB_CUDA(cudaStreamWaitEvent(copy_stream_, wav_ready_[q], 0));
      MB_CUDA(cudaMemcpyAsync(host_wav_dst_ + static_cast<long>(b0) * Lw, wq, static_cast<size_t>(b1 - b0) * Lw * 4,
                              cudaMemcpyDeviceToHost, copy_stream_));
    }
  }
  launches += 1;
  mark_busy(st);
}

void Plan::synthesise(const float* mu, const float* mask, const float* z, const float* spks, int B, int T, int n_steps,
                      int L, float mel_mean, float mel_std, float* dec_out, float* mel_out, float* wav_out, cudaStream_t st) {
  std::lock_guard<std::recursive_mutex> lk(mu_);
  if (L < 1 || L > T) throw Error("vocoder length L must be in [1, T]");
  if (!wav_out) throw Error("wav_out must not be NULL");
  float* dec = dec_out;
  if (!dec) dec = static_cast<float*>(ensure(ws_io_[1], static_cast<size_t>(B) * 80 * T * 4));
  cfm_decode(mu, mask, z, spks, B, T, n_steps, dec, st);
  if (mel_out) { MB_CUDA(launch_affine_crop(dec, T, mel_std, mel_mean, mel_out, B, 80, L, st)); launches += 1; }
  // tanh already bounds the waveform to [-1, 1]: cli.py:118's clamp is the identity here
  hifigan_forward(dec, T, mel_std, mel_mean, B, L, wav_out, st);
}

void Plan::synthesise_host(const float* mu, const float* mask, const float* z, const float* spks, int B, int T, int n_steps,
                           int L, float mel_mean, float mel_std, float* mel_out, float* wav_out) {
  synthesise_host_submit(mu, mask, z, spks, B, T, n_steps, L, mel_mean, mel_std, mel_out, wav_out, 0);
  synthesise_host_wait(0);
}

void Plan::synthesise_host_submit(const float* mu, const float* mask, const float* z, const float* spks, int B, int T,
                                  int n_steps, int L, float mel_mean, float mel_std, float* mel_out, float* wav_out, int slot) {
  std::lock_guard<std::recursive_mutex> lk(mu_);
  if (slot < 0 || slot > 1) throw Error("synthesise_host_submit: slot must be 0 or 1");
  HostSlot& S = slots_[slot];
  if (S.busy) throw Error("synthesise_host_submit: this slot still holds a batch; call mb200_synthesise_host_wait first");
  if (L < 1 || L > T) throw Error("vocoder length L must be in [1, T]");
  const size_t n_mel = static_cast<size_t>(B) * 80 * T, n_mask = static_cast<size_t>(B) * T;
  const size_t n_spk = spks ? static_cast<size_t>(B) * (cfg.c_in - 160) : 0;
  const size_t n_out_mel = static_cast<size_t>(B) * 80 * L, n_wav = static_cast<size_t>(B) * 256 * L;
  // Private NON-BLOCKING streams: the call neither serialises against the legacy default stream's implicit barriers nor
  // stalls other streams of the process.  upload -> compute -> copy, chained by events; the compute stream is also ordered
  // after the plan's last asynchronous call on a caller's stream (the shared workspaces' last user).
  if (!host_stream_) {
    MB_CUDA(cudaStreamCreateWithFlags(&host_stream_, cudaStreamNonBlocking));
    MB_CUDA(cudaStreamCreateWithFlags(&upload_stream_, cudaStreamNonBlocking));
    MB_CUDA(cudaStreamCreateWithFlags(&copy_stream_, cudaStreamNonBlocking));
    MB_CUDA(cudaEventCreateWithFlags(&mel_ready_, cudaEventDisableTiming));
    for (auto& e : wav_ready_) MB_CUDA(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
    for (auto& sl : slots_)
      for (cudaEvent_t* e : {&sl.uploaded, &sl.computed, &sl.copied}) MB_CUDA(cudaEventCreateWithFlags(e, cudaEventDisableTiming));
  }
  // slot staging is private to the slot (not part of any cached program): plain grow-only device buffers
  auto grow = [&](DevBuf& b, size_t bytes) -> float* {
    if (b.bytes < bytes) {
      if (b.p) { MB_CUDA(cudaDeviceSynchronize()); cudaFree(b.p); bytes_allocated -= static_cast<long long>(b.bytes); }
      MB_CUDA(cudaMalloc(&b.p, bytes));
      b.bytes = bytes;
      bytes_allocated += static_cast<long long>(bytes);
    }
    return static_cast<float*>(b.p);
  };
  float* d_mu = grow(S.mu, n_mel * 4);
  float* d_z = grow(S.z, n_mel * 4);
  float* d_mask = grow(S.mask, n_mask * 4);
  float* d_spk = spks ? grow(S.spk, n_spk * 4) : nullptr;
  float* d_melo = grow(S.mel, n_out_mel * 4);
  float* d_wav = grow(S.wav, n_wav * 4);
  float* dec = static_cast<float*>(ensure(ws_io_[1], n_mel * 4));
  MB_CUDA(cudaMemcpyAsync(d_mu, mu, n_mel * 4, cudaMemcpyHostToDevice, upload_stream_));
  MB_CUDA(cudaMemcpyAsync(d_z, z, n_mel * 4, cudaMemcpyHostToDevice, upload_stream_));
  MB_CUDA(cudaMemcpyAsync(d_mask, mask, n_mask * 4, cudaMemcpyHostToDevice, upload_stream_));
  if (spks) MB_CUDA(cudaMemcpyAsync(d_spk, spks, n_spk * 4, cudaMemcpyHostToDevice, upload_stream_));
  MB_CUDA(cudaEventRecord(S.uploaded, upload_stream_));
  cudaStream_t st = host_stream_;
  MB_CUDA(cudaStreamWaitEvent(st, S.uploaded, 0));
  if (busy_) MB_CUDA(cudaStreamWaitEvent(st, busy_, 0));
  cfm_decode(d_mu, d_mask, d_z, d_spk, B, T, n_steps, dec, st);
  // The mel read-back (9 MB at cfg2) overlaps the vocoder: it is issued on the copy stream as soon as decode + crop are done.
  if (mel_out) {
    MB_CUDA(launch_affine_crop(dec, T, mel_std, mel_mean, d_melo, B, 80, L, st));
    launches += 1;
    MB_CUDA(cudaEventRecord(mel_ready_, st));
    MB_CUDA(cudaStreamWaitEvent(copy_stream_, mel_ready_, 0));
    MB_CUDA(cudaMemcpyAsync(mel_out, d_melo, n_out_mel * 4, cudaMemcpyDeviceToHost, copy_stream_));
  }
  host_wav_dst_ = wav_out;                 // hifigan_forward slices conv_post and queues the waveform copies itself
  try {
    hifigan_forward(dec, T, mel_std, mel_mean, B, L, d_wav, st);
  } catch (...) {
    host_wav_dst_ = nullptr;
    throw;
  }
  host_wav_dst_ = nullptr;
  MB_CUDA(cudaEventRecord(S.computed, st));
  MB_CUDA(cudaEventRecord(S.copied, copy_stream_));
  S.busy = true;
}

void Plan::synthesise_host_wait(int slot) {
  if (slot < 0 || slot > 1) throw Error("synthesise_host_wait: slot must be 0 or 1");
  HostSlot& S = slots_[slot];
  if (!S.busy) throw Error("synthesise_host_wait: nothing was submitted in this slot");
  MB_CUDA(cudaEventSynchronize(S.computed));
  MB_CUDA(cudaEventSynchronize(S.copied));
  S.busy = false;
}

}  // namespace mb
