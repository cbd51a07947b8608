// extern "C" boundary (include/matcha_b200.h) and the unit-test hooks.
#include <cstdio>
#include <cstdlib>
#include <algorithm>
#include <cstring>

#include "plan.h"

namespace mb {
const char* last_error_cstr();
void euler_schedule_host(int n_steps, float* ts, float* dts);

void Plan::test_conv1d(const float* x, const float* w, const float* bias, int B, int Cin, int Cout, int T, int K, int dil,
                       int stride, bool transposed, float* y, cudaStream_t st) {
  std::lock_guard<std::recursive_mutex> lk(mu_);
  const int dtype = cfg.dtype;
  const size_t es = esize(dtype);
  const int cpad = Cin % 64 == 0 ? Cin : (Cin == 32 ? 32 : (Cin + 63) / 64 * 64);
  PackedW W;
  int T_out = T, T_o = T, view_rows = T, view_ld = cpad;
  if (transposed) {
    HostTensor wt{w, {Cin, Cout, K}};
    W = pack_convT(wt, bias, stride);
    T_o = T * stride;
  } else if (stride == 2) {
    if (K != 3 || T % 2 || cpad != Cin) throw Error("test_conv1d: stride 2 needs K == 3, even T and Cin % 64 == 0");
    HostTensor wt{w, {Cout, Cin, K}};
    W = pack_conv_stride2(wt, bias, Cin);
    T_out = T / 2; T_o = T / 2; view_rows = T / 2; view_ld = 2 * Cin;
  } else {
    HostTensor wt{w, {Cout, Cin, K}};
    W = pack_conv1d(wt, bias, dil, {{Cin, cpad}});
  }
  // stage: x (B,Cin,T) fp32 -> [B][T][cpad] operand dtype, zero padded channels
  float* x_ntc = static_cast<float*>(dalloc(static_cast<size_t>(B) * T * Cin * 4));
  void* xa = dalloc(static_cast<size_t>(B) * T * cpad * es);
  float* y_ntc = static_cast<float*>(dalloc(static_cast<size_t>(B) * T_o * W.N * 4));
  MB_CUDA(launch_bct_to_btc(x, x_ntc, B, Cin, T, st));
  {
    // reuse pack_mel's "first C channels, zero pad" only fits C == 80; do a generic strided convert instead
    std::vector<float> tmp(static_cast<size_t>(B) * T * Cin);
    MB_CUDA(cudaStreamSynchronize(st));
    MB_CUDA(cudaMemcpy(tmp.data(), x_ntc, tmp.size() * 4, cudaMemcpyDeviceToHost));
    std::vector<float> padded(static_cast<size_t>(B) * T * cpad, 0.f);
    for (size_t r = 0; r < static_cast<size_t>(B) * T; ++r)
      for (int c = 0; c < Cin; ++c) padded[r * cpad + c] = tmp[r * Cin + c];
    void* up = upload_operand(padded);
    MB_CUDA(cudaMemcpy(xa, up, padded.size() * es, cudaMemcpyDeviceToDevice));
  }
  EpiParams e{};
  e.scale = 1.f; e.out0_slope = 1.f; e.out1_slope = 1.f; e.resid_inv = 1.f;
  e.out0 = y_ntc; e.out0_ld = W.N; e.out0_f32 = 1;
  if (const char* v = std::getenv("MB200_PROBE_EPI")) {   // developer probe: HiFi-GAN ResBlock c2-style epilogue
    if (std::atoi(v) && Cin == Cout && cpad == Cin && !transposed && stride == 1 && dtype != MB200_F32) {
      e.resid = xa; e.resid_ld = cpad;
      e.out1 = dalloc(static_cast<size_t>(B) * T_o * W.N * es); e.out1_ld = W.N; e.out1_slope = 0.1f;
      if (std::atoi(v) > 1) { e.out0 = dalloc(static_cast<size_t>(B) * T_o * W.N * es); e.out0_f32 = 0; }
    }
  }
  reserve_split(static_cast<size_t>(B) * T * cpad);
  Launch op = conv_op(W, Src{xa, view_ld, transposed || stride == 1 ? cpad : Cin, view_rows}, nullptr, B, T_out, e);
  MB_CUDA(op(st));
  if (const char* v = std::getenv("MB200_PROBE")) {       // developer probe: time repeated launches of this op
    const int reps = std::max(1, std::atoi(v));
    cudaEvent_t e0, e1;
    MB_CUDA(cudaEventCreate(&e0)); MB_CUDA(cudaEventCreate(&e1));
    MB_CUDA(cudaEventRecord(e0, st));
    for (int r = 0; r < reps; ++r) MB_CUDA(op(st));
    MB_CUDA(cudaEventRecord(e1, st));
    MB_CUDA(cudaStreamSynchronize(st));
    float ms = 0.f;
    MB_CUDA(cudaEventElapsedTime(&ms, e0, e1));
    std::fprintf(stderr, "[probe] B=%d Cin=%d Cout=%d T=%d K=%d dil=%d: %.1f us/launch, %.1f TFLOP/s\n", B, Cin, Cout, T, K, dil,
                 1e3 * ms / reps, op.flops * reps / (ms * 1e9));
    cudaEventDestroy(e0); cudaEventDestroy(e1);
    if (probe_out) {
      std::vector<long long> h(148 * 16);
      MB_CUDA(cudaMemcpy(h.data(), probe_out, h.size() * 8, cudaMemcpyDeviceToHost));
      const char* names[5] = {"producer", "mma", "epi-w2", "epi-w6", "kernel"};
      for (int cta : {0, 147}) {
        std::fprintf(stderr, "  [cta %3d]", cta);
        for (int r = 0; r < 5; ++r) std::fprintf(stderr, " %s %lld/%lld", names[r], h[cta * 16 + 2 * r], h[cta * 16 + 2 * r + 1]);
        std::fprintf(stderr, "  (role: cycles total/in waits; kernel: total/setup)\n      epilogue w2 phases: header %lld, tmem %lld, staging %lld, arithmetic %lld, stores %lld\n", h[cta * 16 + 10],
                     h[cta * 16 + 11], h[cta * 16 + 12], h[cta * 16 + 13], h[cta * 16 + 14]);
      }
    }
  }
  // [B][T_o][N] -> (B, Cout, T_o): transpose all N channels, then copy the first Cout rows per batch element
  float* y_full = static_cast<float*>(dalloc(static_cast<size_t>(B) * T_o * W.N * 4));
  MB_CUDA(launch_btc_to_bct(y_ntc, y_full, B, W.N, T_o, st));
  for (int b = 0; b < B; ++b)
    MB_CUDA(cudaMemcpyAsync(y + static_cast<size_t>(b) * Cout * T_o, y_full + static_cast<size_t>(b) * W.N * T_o,
                            static_cast<size_t>(Cout) * T_o * 4, cudaMemcpyDeviceToDevice, st));
  MB_CUDA(cudaStreamSynchronize(st));
  launches += 3;
}

void Plan::test_attention(const float* qkv, const float* mask, int B, int T, float* out, cudaStream_t st) {
  std::lock_guard<std::recursive_mutex> lk(mu_);
  const int dtype = cfg.dtype;
  const size_t n_in = static_cast<size_t>(B) * T * 384, n_out = static_cast<size_t>(B) * T * 128;
  std::vector<float> h(n_in);
  MB_CUDA(cudaStreamSynchronize(st));
  MB_CUDA(cudaMemcpy(h.data(), qkv, n_in * 4, cudaMemcpyDeviceToHost));
  void* q = upload_operand(h);
  void* o = dalloc(n_out * esize(dtype));
  Launch aop = attention_op(q, mask, o, B, T);
  MB_CUDA(aop(st));
  MB_CUDA(cudaStreamSynchronize(st));
  if (const char* v = std::getenv("MB200_PROBE")) {       // developer probe: time repeated launches of this op
    const int reps = std::max(1, std::atoi(v));
    cudaEvent_t e0, e1;
    MB_CUDA(cudaEventCreate(&e0)); MB_CUDA(cudaEventCreate(&e1));
    MB_CUDA(cudaEventRecord(e0, st));
    for (int r = 0; r < reps; ++r) MB_CUDA(aop(st));
    MB_CUDA(cudaEventRecord(e1, st));
    MB_CUDA(cudaStreamSynchronize(st));
    float ms = 0.f;
    MB_CUDA(cudaEventElapsedTime(&ms, e0, e1));
    std::fprintf(stderr, "[probe] attention B=%d L=%d: %.1f us/launch, %.1f TFLOP/s\n", B, T, 1e3 * ms / reps,
                 aop.flops * reps / (ms * 1e9));
    cudaEventDestroy(e0); cudaEventDestroy(e1);
  }
  if (dtype == MB200_F32) {
    MB_CUDA(cudaMemcpy(out, o, n_out * 4, cudaMemcpyDeviceToDevice));
  } else {
    std::vector<uint16_t> raw(n_out);
    MB_CUDA(cudaMemcpy(raw.data(), o, n_out * 2, cudaMemcpyDeviceToHost));
    std::vector<float> f(n_out);
    for (size_t i = 0; i < n_out; ++i) {
      if (dtype == MB200_BF16) { __nv_bfloat16 v; std::memcpy(&v, &raw[i], 2); f[i] = __bfloat162float(v); }
      else { __half v; std::memcpy(&v, &raw[i], 2); f[i] = __half2float(v); }
    }
    MB_CUDA(cudaMemcpy(out, f.data(), n_out * 4, cudaMemcpyHostToDevice));
  }
  launches += 1;
}
}  // namespace mb

using mb::Plan;

namespace {
template <typename F>
int guarded(F&& f) {
  try {
    f();
    return 0;
  } catch (const std::exception& e) {
    mb::set_last_error(e.what());
    return 1;
  } catch (...) {
    mb::set_last_error("unknown error");
    return 1;
  }
}
mb::StateDict to_state_dict(const mb200_tensor* t, int n, const char* prefix) {
  mb::StateDict sd;
  const std::string pre = prefix ? prefix : "";
  for (int i = 0; i < n; ++i) {
    std::string name = t[i].name;
    if (!pre.empty()) {
      if (name.compare(0, pre.size(), pre) != 0) continue;
      name = name.substr(pre.size());
    }
    mb::HostTensor h{t[i].data, {}};
    for (int d = 0; d < t[i].ndim; ++d) h.shape.push_back(t[i].shape[d]);
    sd.emplace(name, std::move(h));
  }
  return sd;
}
Plan* P(mb200_plan* p) {
  if (!p) throw mb::Error("null plan");
  return reinterpret_cast<Plan*>(p);
}
}  // namespace

extern "C" {

const char* mb200_last_error(void) { return mb::last_error_cstr(); }
const char* mb200_version(void) { return "matcha_b200 0.1 (sm_100a)"; }

int mb200_plan_create(const mb200_cfg* cfg, mb200_plan** out) {
  return guarded([&] {
    if (!cfg || !out) throw mb::Error("null argument");
    *out = reinterpret_cast<mb200_plan*>(new Plan(*cfg));
  });
}
void mb200_plan_destroy(mb200_plan* plan) { delete reinterpret_cast<Plan*>(plan); }

int mb200_load_estimator(mb200_plan* plan, const mb200_tensor* tensors, int n, const char* prefix) {
  return guarded([&] { P(plan)->load_estimator(to_state_dict(tensors, n, prefix)); });
}
int mb200_load_generator(mb200_plan* plan, const mb200_tensor* tensors, int n) {
  return guarded([&] { P(plan)->load_generator(to_state_dict(tensors, n, nullptr)); });
}
int mb200_euler_schedule(int n_steps, float* t_out, float* dt_out) {
  return guarded([&] {
    if (n_steps < 1 || !t_out || !dt_out) throw mb::Error("bad arguments");
    mb::euler_schedule_host(n_steps, t_out, dt_out);
  });
}
int mb200_estimator_forward(mb200_plan* plan, const float* x, const float* mask, const float* mu, const float* t,
                            int t_is_batched, const float* spks, int B, int T, float* out, void* stream) {
  return guarded([&] { P(plan)->estimator_forward(x, mask, mu, t, t_is_batched != 0, spks, B, T, out, static_cast<cudaStream_t>(stream)); });
}
int mb200_cfm_decode(mb200_plan* plan, const float* mu, const float* mask, const float* z, const float* spks, int B, int T,
                     int n_steps, float* out, void* stream) {
  return guarded([&] { P(plan)->cfm_decode(mu, mask, z, spks, B, T, n_steps, out, static_cast<cudaStream_t>(stream)); });
}
int mb200_hifigan_forward(mb200_plan* plan, const float* mel, int B, int L, float* wav, void* stream) {
  return guarded([&] { P(plan)->hifigan_forward(mel, L, 1.f, 0.f, B, L, wav, static_cast<cudaStream_t>(stream)); });
}
int mb200_synthesise(mb200_plan* plan, const float* mu, const float* mask, const float* z, const float* spks, int B, int T,
                     int n_steps, int L, float mel_mean, float mel_std, float* dec_out, float* mel_out, float* wav_out,
                     void* stream) {
  return guarded([&] {
    P(plan)->synthesise(mu, mask, z, spks, B, T, n_steps, L, mel_mean, mel_std, dec_out, mel_out, wav_out, static_cast<cudaStream_t>(stream));
  });
}
int mb200_synthesise_host(mb200_plan* plan, const float* mu, const float* mask, const float* z, const float* spks, int B,
                          int T, int n_steps, int L, float mel_mean, float mel_std, float* mel_out, float* wav_out) {
  return guarded([&] { P(plan)->synthesise_host(mu, mask, z, spks, B, T, n_steps, L, mel_mean, mel_std, mel_out, wav_out); });
}
int mb200_denoiser_bias(mb200_plan* plan, const float* bias_audio, int L, float* bias, void* stream) {
  return guarded([&] {
    if (!bias_audio || !bias) throw mb::Error("null argument");
    P(plan)->denoiser_bias(bias_audio, L, bias, static_cast<cudaStream_t>(stream));
  });
}
int mb200_denoise(mb200_plan* plan, const float* audio, int B, int L, const float* bias, float strength, float* out,
                  void* stream) {
  return guarded([&] {
    if (!audio || !bias || !out) throw mb::Error("null argument");
    P(plan)->denoise(audio, B, L, bias, strength, out, static_cast<cudaStream_t>(stream));
  });
}
int mb200_length_regulate(const float* mu_x, const float* w, const float* x_mask, const long long* y_lengths, int B, int C,
                          int Lx, int T, float* cum, float* mu_y, float* attn, float* y_mask, void* stream) {
  return guarded([&] {
    if (!mu_x || !w || !x_mask || !y_lengths || !cum || !mu_y) throw mb::Error("null argument");
    mb::length_regulate(mu_x, w, x_mask, y_lengths, B, C, Lx, T, cum, mu_y, attn, y_mask, static_cast<cudaStream_t>(stream));
  });
}
long long mb200_maximum_path_workspace_bytes(int B, int Tx, int Ty) {
  if (B <= 0 || Tx <= 0 || Ty <= 0) return 0;
  return static_cast<long long>(mb::maximum_path_workspace_bytes(B, Tx, Ty));
}
int mb200_maximum_path(const float* value, const float* mask, int B, int Tx, int Ty, float* path, void* workspace,
                       void* stream) {
  return guarded([&] {
    if (!value || !mask || !path || !workspace) throw mb::Error("null argument");
    mb::maximum_path(value, mask, B, Tx, Ty, path, workspace, static_cast<cudaStream_t>(stream));
  });
}
int mb200_synthesise_host_submit(mb200_plan* plan, const float* mu, const float* mask, const float* z, const float* spks, int B, int T,
                                 int n_steps, int L, float mel_mean, float mel_std, float* mel_out, float* wav_out, int slot) {
  return guarded([&] {
    if (!mu || !mask || !z || !wav_out) throw mb::Error("null argument");
    P(plan)->synthesise_host_submit(mu, mask, z, spks, B, T, n_steps, L, mel_mean, mel_std, mel_out, wav_out, slot);
  });
}
int mb200_synthesise_host_wait(mb200_plan* plan, int slot) {
  return guarded([&] { P(plan)->synthesise_host_wait(slot); });
}
int mb200_load_text_encoder(mb200_plan* plan, const mb200_tensor* tensors, int n, const char* prefix, int n_heads) {
  return guarded([&] { P(plan)->load_text_encoder(to_state_dict(tensors, n, prefix ? prefix : ""), n_heads); });
}
int mb200_text_encoder_forward(mb200_plan* plan, const long long* ids, const long long* lengths, const float* spks, int B, int Lx,
                               float* mu_x, float* logw, float* x_mask, void* stream) {
  return guarded([&] {
    if (!ids || !lengths || !mu_x || !logw || !x_mask) throw mb::Error("null argument");
    P(plan)->text_encoder_forward(ids, lengths, spks, B, Lx, mu_x, logw, x_mask, static_cast<cudaStream_t>(stream));
  });
}
int mb200_maximum_path_static(const float* value, const float* mask, int B, int Tx, int Ty, float* path, void* stream) {
  return guarded([&] {
    if (!value || !mask || !path) throw mb::Error("null argument");
    mb::maximum_path_static(value, mask, B, Tx, Ty, path, static_cast<cudaStream_t>(stream));
  });
}
int mb200_set_train_dropout(mb200_plan* plan, float p, unsigned long long seed) {
  return guarded([&] { P(plan)->set_train_dropout(p, seed); });
}
int mb200_compute_loss(mb200_plan* plan, const mb200_param* params, int n, const char* prefix, const float* x1,
                       const float* mask, const float* mu, const float* z, const float* t, const float* spks, int B, int T,
                       float sigma_min, int backward, float* loss, float* y, float* v, float* dmu, float* dspks,
                       void* stream) {
  return guarded([&] {
    if (!params || !x1 || !mask || !mu || !z || !t || !loss) throw mb::Error("null argument");
    std::map<std::string, std::pair<const float*, float*>> m;
    const std::string pre = prefix ? prefix : "";
    for (int i = 0; i < n; ++i) {
      std::string name = params[i].name;
      if (!pre.empty()) {
        if (name.compare(0, pre.size(), pre) != 0) continue;
        name = name.substr(pre.size());
      }
      m[name] = {params[i].param, params[i].grad};
    }
    P(plan)->compute_loss_grad(m, x1, mask, mu, z, t, spks, B, T, sigma_min, loss, y, v, dmu, dspks, backward != 0,
                               static_cast<cudaStream_t>(stream));
  });
}
long long mb200_launch_count(const mb200_plan* plan) { return plan ? reinterpret_cast<const Plan*>(plan)->launches : 0; }
int mb200_set_profiling(mb200_plan* plan, int on) {
  return guarded([&] {
    Plan* p = P(plan);
    p->profiling = on != 0;
    std::memset(p->prof, 0, sizeof(p->prof));
  });
}
int mb200_get_profile(const mb200_plan* plan, double* out) {
  return guarded([&] {
    if (!plan || !out) throw mb::Error("null argument");
    std::memcpy(out, reinterpret_cast<const Plan*>(plan)->prof, sizeof(double) * mb::kProfRegions * mb::kProfKinds * 3);
  });
}
long long mb200_workspace_bytes(const mb200_plan* plan) { return plan ? reinterpret_cast<const Plan*>(plan)->bytes_allocated : 0; }

int mb200_debug_probe(mb200_plan* plan, long long* out, int n_ctas) {
  return guarded([&] {
    Plan* p = P(plan);
    if (!p->probe_out || !out || n_ctas < 1 || n_ctas > 148) throw mb::Error("no instrumented kernel (or bad arguments)");
    MB_CUDA(cudaDeviceSynchronize());
    MB_CUDA(cudaMemcpy(out, p->probe_out, static_cast<size_t>(n_ctas) * 16 * sizeof(long long), cudaMemcpyDeviceToHost));
  });
}
int mb200_test_conv1d(mb200_plan* plan, const float* x, const float* w, const float* bias, int B, int Cin, int Cout, int T,
                      int K, int dilation, int stride, int transposed, float* y, void* stream) {
  return guarded([&] { P(plan)->test_conv1d(x, w, bias, B, Cin, Cout, T, K, dilation, stride, transposed != 0, y, static_cast<cudaStream_t>(stream)); });
}
int mb200_test_attention(mb200_plan* plan, const float* qkv, const float* mask, int B, int T, float* out, void* stream) {
  return guarded([&] { P(plan)->test_attention(qkv, mask, B, T, out, static_cast<cudaStream_t>(stream)); });
}

}  // extern "C"
