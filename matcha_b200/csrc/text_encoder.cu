// The text encoder and duration predictor as CUDA kernels:
//   TextEncoder.forward            matcha/models/components/text_encoder.py:378-410
//   ConvReluNorm (prenet)          :36-67      conv k5 -> channel LayerNorm -> ReLU (x3), 1x1 proj, residual, mask
//   Encoder                        :276-325    6 x [RoPE multi-head attention -> +res -> LayerNorm -> conv-k3 FFN -> +res -> LayerNorm]
//   MultiHeadAttention / RoPE      :96-166, :175-252   rotary embedding on the first half of every head, scores / sqrt(Dh),
//                                              masked_fill(mask == 0, -1e4) (REPLACES the logit, unlike the decoder's additive mask)
//   DurationPredictor              :70-94      conv k3 -> ReLU -> LayerNorm (x2), 1x1 proj, mask
//   LayerNorm (over channels)      :16-33      eps 1e-4, parameters gamma / beta
// Everything runs in fp32: the durations go through ceil(exp(logw)) downstream (matcha_tts.py:118-119), so a half-precision
// encoder would change frame counts.  Convolutions and 1x1 projections are the plan's conv-as-GEMM kernels on [B][Lx][C]
// activations (conv_op, fp32 mode); the kernels below are what the decoder path does not have.  Values at padded positions
// are not the reference's (it carries LayerNorm outputs there): they never reach a valid position - every consumer
// masks its input or its keys - and the three outputs are masked, so mu_x / logw / x_mask agree everywhere.
#include <cmath>

#include "plan.h"

namespace mb {

namespace {

// ---------------------------------------------------------------- x_mask, emb(x) * sqrt(C) * mask   (text_encoder.py:395-397)
__global__ void __launch_bounds__(256)
embed_kernel(const long long* __restrict__ ids, const long long* __restrict__ lens, const float* __restrict__ emb, int n_vocab,
             int C, float scale, int Lx, long rows, float* __restrict__ out, float* __restrict__ mask) {
  const long row = static_cast<long>(blockIdx.x) * 8 + (threadIdx.x >> 5);
  if (row >= rows) return;
  const int lane = threadIdx.x & 31;
  const int b = static_cast<int>(row / Lx), l = static_cast<int>(row - static_cast<long>(b) * Lx);
  const float m = l < lens[b] ? 1.f : 0.f;
  long long id = ids[row];
  id = id < 0 ? 0 : (id >= n_vocab ? n_vocab - 1 : id);
  for (int c = lane; c < C; c += 32) out[row * C + c] = emb[id * C + c] * scale * m;
  if (lane == 0) mask[row] = m;
}

// channels [c0, c0 + n) of every row <- spk[b][0..n)   (text_encoder.py:401-402: the speaker vector is concatenated over time)
__global__ void __launch_bounds__(256)
fill_spk_kernel(const float* __restrict__ spk, const float* __restrict__ mask, int n, int c0, int C, int Lx, long rows,
                float* __restrict__ x) {
  const long i = static_cast<long>(blockIdx.x) * 256 + threadIdx.x;
  if (i >= rows * n) return;
  const long row = i / n;
  const int c = static_cast<int>(i - row * n);
  x[row * C + c0 + c] = spk[(row / Lx) * n + c] * mask[row];
}

// ---------------------------------------------------------------- channel LayerNorm, eps 1e-4 (text_encoder.py:16-33)
// y = LN(pre_relu ? relu(x) : x); y = post_relu ? relu(y) : y; y *= mask (if given).  Warp per row, two-pass variance as
// the reference computes it.  C <= 256.
__global__ void __launch_bounds__(256)
channel_ln_kernel(const float* __restrict__ in, const float* __restrict__ gamma, const float* __restrict__ beta,
                  const float* __restrict__ mask, int C, int pre_relu, int post_relu, long rows, float* __restrict__ out) {
  const long row = static_cast<long>(blockIdx.x) * 8 + (threadIdx.x >> 5);
  if (row >= rows) return;
  const int lane = threadIdx.x & 31;
  float v[8];
  float s = 0.f;
#pragma unroll
  for (int k = 0; k < 8; ++k) {
    const int c = lane + 32 * k;
    v[k] = c < C ? in[row * C + c] : 0.f;
    if (pre_relu) v[k] = fmaxf(v[k], 0.f);
    s += v[k];
  }
  const float mean = warp_sum(s) / static_cast<float>(C);
  float ss = 0.f;
#pragma unroll
  for (int k = 0; k < 8; ++k) {
    const float d = (lane + 32 * k < C) ? v[k] - mean : 0.f;
    ss += d * d;
  }
  const float rstd = rsqrtf(warp_sum(ss) / static_cast<float>(C) + 1e-4f);
  const float m = mask ? mask[row] : 1.f;
#pragma unroll
  for (int k = 0; k < 8; ++k) {
    const int c = lane + 32 * k;
    if (c < C) {
      float y = (v[k] - mean) * rstd * __ldg(gamma + c) + __ldg(beta + c);
      if (post_relu) y = fmaxf(y, 0.f);
      out[row * C + c] = y * m;
    }
  }
}

// ---------------------------------------------------------------- rotary embedding, in place on q and k of [rows][3C]
// First d = Dh / 2 features of every head (text_encoder.py:197-198: RotaryPositionalEmbeddings(k_channels * 0.5)); feature i
// pairs with i + d/2; angle = position * theta[i], theta[i] = 10000^(-2i/d) (table built on the host in fp32, as torch
// builds it).  q is also scaled by 1 / sqrt(Dh) (scores / sqrt(k_channels), :219).
__global__ void __launch_bounds__(256)
rope_kernel(float* __restrict__ qkv, const float* __restrict__ theta, int C, int H, int Dh, int Lx, long rows, float q_scale) {
  const int half = Dh / 4;                     // d / 2 rotation pairs per head
  const long i = static_cast<long>(blockIdx.x) * 256 + threadIdx.x;
  const long total = rows * 2 * H * half;
  if (i >= total) return;
  const int pr = static_cast<int>(i % half);
  const int hh = static_cast<int>((i / half) % H);
  const int which = static_cast<int>((i / (static_cast<long>(half) * H)) % 2);      // 0 = q, 1 = k
  const long row = i / (static_cast<long>(half) * H * 2);
  const int pos = static_cast<int>(row % Lx);
  float sn, cs;
  sincosf(static_cast<float>(pos) * theta[pr], &sn, &cs);
  float* p = qkv + row * 3 * C + which * C + hh * Dh;
  const float a = p[pr], bq = p[pr + half];
  const float sc = which == 0 ? q_scale : 1.f;
  p[pr] = (a * cs - bq * sn) * sc;
  p[pr + half] = (bq * cs + a * sn) * sc;
}
// the un-rotated second half of q still needs the 1 / sqrt(Dh) scale
__global__ void __launch_bounds__(256)
scale_q_tail_kernel(float* __restrict__ qkv, int C, int H, int Dh, long rows, float q_scale) {
  const int tail = Dh / 2;
  const long i = static_cast<long>(blockIdx.x) * 256 + threadIdx.x;
  if (i >= rows * H * tail) return;
  const int d = static_cast<int>(i % tail);
  const int hh = static_cast<int>((i / tail) % H);
  const long row = i / (static_cast<long>(tail) * H);
  qkv[row * 3 * C + hh * Dh + Dh / 2 + d] *= q_scale;
}

// ---------------------------------------------------------------- multi-head attention, head dim D in {96, 128}
// out[b, q, h*D ...] = softmax_k(mask[b,k] != 0 ? q . k : -1e4) v.  One thread per query row (q and the output row in
// registers), K / V tiles of 32 keys broadcast from shared memory, online softmax in fp32 (as attention_simt_kernel).
// A single long utterance has too few query rows to fill the GPU (2 heads x 2001 rows = 64 CTAs), so the KEYS can be split
// over ks CTAs (blockIdx.x = query block * ks + split): each writes its unnormalised row, running maximum and sum to
// `part` ([b][h][q][split][D + 2]) and text_attention_merge_kernel combines them.  ks == 1 writes `out` directly.
template <int D>
__global__ void __launch_bounds__(64)
text_attention_kernel(const float* __restrict__ qkv, const float* __restrict__ mask, float* __restrict__ out, int L, int C,
                      int ks, float* __restrict__ part) {
  __shared__ __align__(16) float Ks[32][D];
  __shared__ __align__(16) float Vs[32][D];
  __shared__ float Ms[32];
  const int b = blockIdx.z, h = blockIdx.y;
  const int tid = threadIdx.x;
  const int split = blockIdx.x % ks;
  const int qi = (blockIdx.x / ks) * 64 + tid;
  const bool qvalid = qi < L;
  const float* base = qkv + static_cast<long>(b) * L * 3 * C;
  const int n_kt = (L + 31) / 32;                      // key tiles; this CTA takes [kt0, kt1)
  const int kt0 = static_cast<int>(static_cast<long>(n_kt) * split / ks), kt1 = static_cast<int>(static_cast<long>(n_kt) * (split + 1) / ks);
  float q[D], o[D];
#pragma unroll
  for (int d = 0; d < D; ++d) { o[d] = 0.f; q[d] = 0.f; }
  if (qvalid) {
    const float4* qp = reinterpret_cast<const float4*>(base + static_cast<long>(qi) * 3 * C + h * D);
#pragma unroll
    for (int d4 = 0; d4 < D / 4; ++d4) { const float4 t = qp[d4]; q[4 * d4] = t.x; q[4 * d4 + 1] = t.y; q[4 * d4 + 2] = t.z; q[4 * d4 + 3] = t.w; }
  }
  float mrun = -INFINITY, lrun = 0.f;
  for (int k0 = kt0 * 32; k0 < min(L, kt1 * 32); k0 += 32) {
    __syncthreads();
    for (int i = tid; i < 32 * (D / 4); i += 64) {     // 32 keys x D/4 float4 of K and of V
      const int kj = i / (D / 4), d4 = i - kj * (D / 4);
      float4 kk = make_float4(0.f, 0.f, 0.f, 0.f), vv = kk;
      if (k0 + kj < L) {
        const float* kp = base + static_cast<long>(k0 + kj) * 3 * C + C + h * D + 4 * d4;
        kk = *reinterpret_cast<const float4*>(kp);
        vv = *reinterpret_cast<const float4*>(kp + C);
      }
      *reinterpret_cast<float4*>(&Ks[kj][4 * d4]) = kk;
      *reinterpret_cast<float4*>(&Vs[kj][4 * d4]) = vv;
    }
    if (tid < 32) Ms[tid] = (k0 + tid < L) ? mask[static_cast<long>(b) * L + k0 + tid] : 0.f;
    __syncthreads();
    const int nk = min(32, L - k0);
    for (int j0 = 0; j0 < nk; j0 += 4) {
      float s[4];
      float mx = -INFINITY;
#pragma unroll
      for (int jj = 0; jj < 4; ++jj) {
        const int j = j0 + jj;
        float acc = -INFINITY;
        if (j < nk) {
          acc = 0.f;
          const float4* kr = reinterpret_cast<const float4*>(Ks[j]);
#pragma unroll
          for (int d4 = 0; d4 < D / 4; ++d4) {
            const float4 kk = kr[d4];
            acc = fmaf(q[4 * d4], kk.x, acc); acc = fmaf(q[4 * d4 + 1], kk.y, acc);
            acc = fmaf(q[4 * d4 + 2], kk.z, acc); acc = fmaf(q[4 * d4 + 3], kk.w, acc);
          }
          if (Ms[j] == 0.f) acc = -1e4f;             // masked_fill(mask == 0, -1e4): replaces the logit
        }
        s[jj] = acc;
        mx = fmaxf(mx, acc);
      }
      if (mx > mrun) {
        const float sc = expf(mrun - mx);            // exp(-inf) = 0 on the first chunk
        lrun *= sc;
#pragma unroll
        for (int d = 0; d < D; ++d) o[d] *= sc;
        mrun = mx;
      }
#pragma unroll
      for (int jj = 0; jj < 4; ++jj) {
        const int j = j0 + jj;
        if (j < nk) {
          const float p = expf(s[jj] - mrun);
          lrun += p;
          const float4* vr = reinterpret_cast<const float4*>(Vs[j]);
#pragma unroll
          for (int d4 = 0; d4 < D / 4; ++d4) {
            const float4 vv = vr[d4];
            o[4 * d4] = fmaf(p, vv.x, o[4 * d4]); o[4 * d4 + 1] = fmaf(p, vv.y, o[4 * d4 + 1]);
            o[4 * d4 + 2] = fmaf(p, vv.z, o[4 * d4 + 2]); o[4 * d4 + 3] = fmaf(p, vv.w, o[4 * d4 + 3]);
          }
        }
      }
    }
  }
  if (qvalid && ks == 1) {
    const float inv = 1.f / lrun;
    float4* op = reinterpret_cast<float4*>(out + (static_cast<long>(b) * L + qi) * C + h * D);
#pragma unroll
    for (int d4 = 0; d4 < D / 4; ++d4) op[d4] = make_float4(o[4 * d4] * inv, o[4 * d4 + 1] * inv, o[4 * d4 + 2] * inv, o[4 * d4 + 3] * inv);
  } else if (qvalid) {
    float* pp = part + (((static_cast<long>(b) * gridDim.y + h) * L + qi) * ks + split) * (D + 2);
#pragma unroll
    for (int d = 0; d < D; ++d) pp[d] = o[d];
    pp[D] = mrun;
    pp[D + 1] = lrun;
  }
}

// out[b, q, h*D + d] = sum_s e^(m_s - m) o_s[d] / sum_s e^(m_s - m) l_s over the key splits, m = max_s m_s.  One thread per
// (b, h, q, d); a split that saw no key has m_s = -inf and drops out.
template <int D>
__global__ void __launch_bounds__(256)
text_attention_merge_kernel(const float* __restrict__ part, float* __restrict__ out, int L, int C, int H, int ks, long total) {
  const long i = static_cast<long>(blockIdx.x) * 256 + threadIdx.x;
  if (i >= total) return;
  const int d = static_cast<int>(i % D);
  const long row = i / D;                               // (b * H + h) * L + q
  const int q = static_cast<int>(row % L);
  const long bh = row / L;
  const int h = static_cast<int>(bh % H);
  const long b = bh / H;
  const float* pp = part + row * ks * (D + 2);
  float m = -INFINITY;
  for (int s = 0; s < ks; ++s) m = fmaxf(m, pp[s * (D + 2) + D]);
  float num = 0.f, den = 0.f;
  for (int s = 0; s < ks; ++s) {
    const float ms = pp[s * (D + 2) + D];
    if (ms == -INFINITY) continue;
    const float w = expf(ms - m);
    num = fmaf(w, pp[s * (D + 2) + d], num);
    den = fmaf(w, pp[s * (D + 2) + D + 1], den);
  }
  out[(b * L + q) * C + h * D + d] = num / den;
}

// (rows, 1) column 0 of a [rows][ld] buffer -> dense, times mask   (DurationPredictor.proj output, text_encoder.py:93-94)
__global__ void __launch_bounds__(256)
take_col0_kernel(const float* __restrict__ in, int ld, const float* __restrict__ mask, long rows, float* __restrict__ out) {
  const long i = static_cast<long>(blockIdx.x) * 256 + threadIdx.x;
  if (i < rows) out[i] = in[i * ld] * mask[i];
}

}  // namespace

namespace {
const HostTensor& need_t(const StateDict& sd, const std::string& k) {
  auto it = sd.find(k);
  if (it == sd.end()) throw Error("state_dict is missing '" + k + "'");
  return it->second;
}
EpiParams epi_f32() {
  EpiParams e{};
  e.scale = 1.f; e.out0_slope = 1.f; e.out1_slope = 1.f; e.resid_inv = 1.f; e.o_mul = 1; e.out0_f32 = 1;
  return e;
}
}  // namespace

void Plan::load_text_encoder(const StateDict& sd, int n_heads) {
  std::lock_guard<std::recursive_mutex> lk(mu_);
  if (cfg.dtype != MB200_F32) throw Error("the text encoder runs on fp32 plans only (durations go through ceil(exp(logw)))");
  TextEncW& E = enc_;
  auto up = [&](const std::string& k) { const HostTensor& t = need_t(sd, k); return upload_f32(t.data, t.numel()); };
  const HostTensor& emb = need_t(sd, "emb.weight");
  E.n_vocab = static_cast<int>(emb.shape[0]);
  E.C0 = static_cast<int>(emb.shape[1]);
  E.C = static_cast<int>(need_t(sd, "encoder.attn_layers.0.conv_q.weight").shape[0]);
  E.spk_dim = E.C - E.C0;
  E.H = n_heads;
  if (E.H < 1 || E.C % E.H) throw Error("text encoder: channels must divide into heads");
  E.Dh = E.C / E.H;
  if (E.Dh != 96 && E.Dh != 128) throw Error("text encoder: kernels exist for head dims 96 (LJSpeech) and 128 (VCTK), got " + std::to_string(E.Dh));
  if (E.C0 % 64 || E.C % 64 || E.C > 256) throw Error("text encoder: channel counts must be multiples of 64, at most 256");
  E.n_layers = 0;
  while (sd.count("encoder.attn_layers." + std::to_string(E.n_layers) + ".conv_q.weight")) ++E.n_layers;
  if (E.n_layers < 1 || E.n_layers > 8) throw Error("text encoder: 1..8 layers supported");
  E.filt = static_cast<int>(need_t(sd, "encoder.ffn_layers.0.conv_1.weight").shape[0]);
  E.filt_dp = static_cast<int>(need_t(sd, "proj_w.conv_1.weight").shape[0]);
  E.n_feats = static_cast<int>(need_t(sd, "proj_m.weight").shape[0]);
  if (E.filt % 64 || E.filt_dp % 64 || E.filt_dp > 256 || E.n_feats != 80) throw Error("text encoder: unsupported filter / feature sizes");
  E.emb = upload_f32(emb.data, emb.numel());
  for (int i = 0; i < 3; ++i) {
    const std::string p = "prenet.conv_layers." + std::to_string(i);
    if (!sd.count(p + ".weight")) throw Error("text encoder: the prenet (3 conv layers) is required");
    E.pre_conv[i] = pack_conv1d(need_t(sd, p + ".weight"), need_t(sd, p + ".bias").data, 1, {{E.C0, E.C0}});
    E.pre_g[i] = up("prenet.norm_layers." + std::to_string(i) + ".gamma");
    E.pre_b[i] = up("prenet.norm_layers." + std::to_string(i) + ".beta");
  }
  E.pre_proj = pack_conv1d(need_t(sd, "prenet.proj.weight"), need_t(sd, "prenet.proj.bias").data, 1, {{E.C0, E.C0}});
  const int C = E.C;
  for (int l = 0; l < E.n_layers; ++l) {
    const std::string a = "encoder.attn_layers." + std::to_string(l) + ".";
    // fused q | k | v projection: rows [conv_q ; conv_k ; conv_v], biases alike
    std::vector<float> w(static_cast<size_t>(3) * C * C), bb(static_cast<size_t>(3) * C);
    const char* nm[3] = {"conv_q", "conv_k", "conv_v"};
    for (int i = 0; i < 3; ++i) {
      const HostTensor& wt = need_t(sd, a + nm[i] + ".weight");
      if (wt.numel() != static_cast<long long>(C) * C) throw Error("text encoder: attention projections must be (C, C, 1)");
      std::copy(wt.data, wt.data + static_cast<size_t>(C) * C, w.begin() + static_cast<size_t>(i) * C * C);
      std::copy(need_t(sd, a + nm[i] + ".bias").data, need_t(sd, a + nm[i] + ".bias").data + C, bb.begin() + static_cast<size_t>(i) * C);
    }
    HostTensor wq{w.data(), {3 * C, C, 1}};
    TextEncW::Layer& Ly = E.layer[l];
    Ly.qkv = pack_conv1d(wq, bb.data(), 1, {{C, C}});
    Ly.o = pack_conv1d(need_t(sd, a + "conv_o.weight"), need_t(sd, a + "conv_o.bias").data, 1, {{C, C}});
    const std::string f = "encoder.ffn_layers." + std::to_string(l) + ".";
    Ly.ffn1 = pack_conv1d(need_t(sd, f + "conv_1.weight"), need_t(sd, f + "conv_1.bias").data, 1, {{C, C}});
    Ly.ffn2 = pack_conv1d(need_t(sd, f + "conv_2.weight"), need_t(sd, f + "conv_2.bias").data, 1, {{E.filt, E.filt}});
    Ly.n1_g = up("encoder.norm_layers_1." + std::to_string(l) + ".gamma");
    Ly.n1_b = up("encoder.norm_layers_1." + std::to_string(l) + ".beta");
    Ly.n2_g = up("encoder.norm_layers_2." + std::to_string(l) + ".gamma");
    Ly.n2_b = up("encoder.norm_layers_2." + std::to_string(l) + ".beta");
  }
  E.proj_m = pack_conv1d(need_t(sd, "proj_m.weight"), need_t(sd, "proj_m.bias").data, 1, {{C, C}});
  E.dp_conv1 = pack_conv1d(need_t(sd, "proj_w.conv_1.weight"), need_t(sd, "proj_w.conv_1.bias").data, 1, {{C, C}});
  E.dp_conv2 = pack_conv1d(need_t(sd, "proj_w.conv_2.weight"), need_t(sd, "proj_w.conv_2.bias").data, 1, {{E.filt_dp, E.filt_dp}});
  E.dp_proj = pack_conv1d(need_t(sd, "proj_w.proj.weight"), need_t(sd, "proj_w.proj.bias").data, 1, {{E.filt_dp, E.filt_dp}});
  E.dp_g1 = up("proj_w.norm_1.gamma"); E.dp_b1 = up("proj_w.norm_1.beta");
  E.dp_g2 = up("proj_w.norm_2.gamma"); E.dp_b2 = up("proj_w.norm_2.beta");
  // rotary frequencies: theta_i = 1 / 10000^(2 i / d), d = Dh / 2, evaluated in fp32 as torch evaluates it on the host
  const int d = E.Dh / 2;
  std::vector<float> th(d / 2);
  for (int i = 0; i < d / 2; ++i) th[i] = 1.0f / std::pow(10000.0f, static_cast<float>(2 * i) / static_cast<float>(d));
  E.theta = upload_f32(th.data(), th.size());
  E.loaded = true;
}

void Plan::text_encoder_forward(const long long* ids_in, const long long* lengths_in, const float* spks_in, int B, int Lx, float* mu_x,
                                float* logw, float* x_mask, cudaStream_t st) {
  std::lock_guard<std::recursive_mutex> lk(mu_);
  const TextEncW& E = enc_;
  if (!E.loaded) throw Error("text encoder weights are not loaded");
  if (B < 1 || Lx < 1) throw Error("text encoder: empty batch");
  if ((E.spk_dim > 0) != (spks_in != nullptr)) throw Error("spks must be given exactly when the encoder has speaker channels");
  const long rows = static_cast<long>(B) * Lx;
  const int C = E.C, C0 = E.C0;
  const int wide = std::max(E.filt, 3 * C);
  enum { X0, H0, H1, X, Y, Z, A, M, P, D1, APART, IDS, LENS, SPK, MUX, LOGW };
  reserve_split(rows * static_cast<size_t>(std::max(std::max(wide, C0), 2 * E.filt_dp)));
  float* x0 = static_cast<float*>(ensure(ws_enc_[X0], rows * C0 * 4));
  float* h0 = static_cast<float*>(ensure(ws_enc_[H0], rows * C0 * 4));
  float* h1 = static_cast<float*>(ensure(ws_enc_[H1], rows * C0 * 4));
  float* x = static_cast<float*>(ensure(ws_enc_[X], rows * C * 4));
  float* y = static_cast<float*>(ensure(ws_enc_[Y], rows * C * 4));
  float* z = static_cast<float*>(ensure(ws_enc_[Z], rows * wide * 4));
  float* a = static_cast<float*>(ensure(ws_enc_[A], rows * C * 4));
  float* m = static_cast<float*>(ensure(ws_enc_[M], rows * 4));
  float* p = static_cast<float*>(ensure(ws_enc_[P], rows * 80 * 4));
  float* d1 = static_cast<float*>(ensure(ws_enc_[D1], rows * 2 * E.filt_dp * 4));
  float* d2 = d1 + rows * E.filt_dp;
  // key splits of the attention: enough CTAs for two per SM, at most 8 and at most one split per 4 key tiles
  const int att_ctas = ((Lx + 63) / 64) * E.H * B;
  int att_ks = std::min(8, (2 * num_sms + att_ctas - 1) / att_ctas);
  att_ks = std::max(1, std::min(att_ks, ((Lx + 31) / 32) / 4));
  // The launch sequence is built once per (B, Lx), captured into a CUDA graph (run()) and replayed: at B = 1 the ~90 launches
  // are otherwise bound by the host's launch rate.  Inputs and outputs go through plan-owned buffers so that the recorded
  // pointers stay valid whatever the caller passes next time.
  long long* ids_s = static_cast<long long*>(ensure(ws_enc_[IDS], rows * 8));
  long long* lens_s = static_cast<long long*>(ensure(ws_enc_[LENS], static_cast<size_t>(B) * 8));
  float* spk_s = E.spk_dim > 0 ? static_cast<float*>(ensure(ws_enc_[SPK], static_cast<size_t>(B) * E.spk_dim * 4)) : nullptr;
  float* mux_s = static_cast<float*>(ensure(ws_enc_[MUX], rows * 80 * 4));
  float* logw_s = static_cast<float*>(ensure(ws_enc_[LOGW], rows * 4));
  float* att_part = att_ks > 1 ? static_cast<float*>(ensure(ws_enc_[APART], rows * E.H * att_ks * (E.Dh + 2) * 4)) : nullptr;
  const unsigned row_blocks = static_cast<unsigned>((rows + 7) / 8);
  auto grid1 = [](long n) { return static_cast<unsigned>((n + 255) / 256); };
  auto key = std::make_tuple(B, Lx);
  auto it = encoder_cache_.find(key);
  if (it == encoder_cache_.end()) {
  auto prog = std::make_unique<Program>();
  auto emit = [&](std::function<cudaError_t(cudaStream_t)> fn) { prog->ops.emplace_back(std::move(fn)); prog->kernels++; };
  auto conv = [&](const PackedW& W, const float* in, int in_ch, EpiParams e) {
    prog->ops.push_back(conv_op(W, Src{in, in_ch, in_ch, Lx}, nullptr, B, Lx, e));
    prog->kernels++;
  };
  auto ln = [&](const float* in, const float* g, const float* b, const float* mask, int ch, int pre_relu, int post_relu, float* out) {
    emit([=](cudaStream_t s) { channel_ln_kernel<<<row_blocks, 256, 0, s>>>(in, g, b, mask, ch, pre_relu, post_relu, rows, out); return cudaGetLastError(); });
  };
  const long long* ids = ids_s;
  const long long* lengths = lens_s;
  const float* spks = spk_s;

  // ---- emb(x) * sqrt(C0) * x_mask  (text_encoder.py:395-397)
  {
    const float* emb = E.emb; const int n_vocab = E.n_vocab; const float esc = std::sqrt(static_cast<float>(C0));
    emit([=](cudaStream_t s) { embed_kernel<<<row_blocks, 256, 0, s>>>(ids, lengths, emb, n_vocab, C0, esc, Lx, rows, x0, m); return cudaGetLastError(); });
  }
  // ---- prenet (:36-67): x = x_org; 3 x [conv k5 (x * mask) -> LayerNorm -> ReLU]; (x_org + proj(x)) * mask
  const float* h = x0;
  for (int i = 0; i < 3; ++i) {
    EpiParams e = epi_f32();
    e.out0 = h0; e.out0_ld = C0;
    conv(E.pre_conv[i], h, C0, e);
    ln(h0, E.pre_g[i], E.pre_b[i], m, C0, 0, 1, h1);
    h = h1;
  }
  {
    EpiParams e = epi_f32();
    e.resid = x0; e.resid_ld = C0; e.resid_f32 = 1;
    e.mask = m; e.mask_out = 1;
    e.out0 = x; e.out0_ld = C;          // lands in the first C0 channels of the encoder's C-wide rows
    conv(E.pre_proj, h1, C0, e);
  }
  if (E.spk_dim > 0) {                   // :401-402, times the mask the encoder applies first thing (:313)
    const int sd = E.spk_dim; const unsigned gsp = grid1(rows * E.spk_dim);
    emit([=](cudaStream_t s) { fill_spk_kernel<<<gsp, 256, 0, s>>>(spks, m, sd, C0, C, Lx, rows, x); return cudaGetLastError(); });
  }
  // ---- encoder (:276-325)
  const float q_scale = 1.0f / std::sqrt(static_cast<float>(E.Dh));
  for (int l = 0; l < E.n_layers; ++l) {
    const TextEncW::Layer& Ly = E.layer[l];
    EpiParams e = epi_f32();
    e.out0 = z; e.out0_ld = 3 * C;
    conv(Ly.qkv, x, C, e);
    const float* theta = E.theta; const int H = E.H, Dh = E.Dh;
    const unsigned g_rope = grid1(rows * 2 * E.H * (E.Dh / 4)), g_tail = grid1(rows * E.H * (E.Dh / 2));
    emit([=](cudaStream_t s) { rope_kernel<<<g_rope, 256, 0, s>>>(z, theta, C, H, Dh, Lx, rows, q_scale); return cudaGetLastError(); });
    emit([=](cudaStream_t s) { scale_q_tail_kernel<<<g_tail, 256, 0, s>>>(z, C, H, Dh, rows, q_scale); return cudaGetLastError(); });
    const dim3 ag(((Lx + 63) / 64) * att_ks, E.H, B);
    emit([=](cudaStream_t s) {
      if (Dh == 96) text_attention_kernel<96><<<ag, 64, 0, s>>>(z, m, a, Lx, C, att_ks, att_part);
      else text_attention_kernel<128><<<ag, 64, 0, s>>>(z, m, a, Lx, C, att_ks, att_part);
      return cudaGetLastError();
    });
    if (att_ks > 1) {
      const long total = rows * E.H * E.Dh;
      const unsigned gm = grid1(total);
      emit([=](cudaStream_t s) {
        if (Dh == 96) text_attention_merge_kernel<96><<<gm, 256, 0, s>>>(att_part, a, Lx, C, H, att_ks, total);
        else text_attention_merge_kernel<128><<<gm, 256, 0, s>>>(att_part, a, Lx, C, H, att_ks, total);
        return cudaGetLastError();
      });
    }
    e = epi_f32();                       // x = LayerNorm(x + conv_o(attn))
    e.resid = x; e.resid_ld = C; e.resid_f32 = 1;
    e.out0 = y; e.out0_ld = C;
    conv(Ly.o, a, C, e);
    ln(y, Ly.n1_g, Ly.n1_b, m, C, 0, 0, x);
    e = epi_f32();                       // FFN: conv_2(relu(conv_1(x * mask)) * mask) * mask  (:240-257)
    e.mask = m; e.mask_out = 1; e.out0_slope = 0.f;   // LeakyReLU with slope 0 = ReLU
    e.out0 = z; e.out0_ld = E.filt;
    conv(Ly.ffn1, x, C, e);
    e = epi_f32();
    e.resid = x; e.resid_ld = C; e.resid_f32 = 1;
    e.out0 = y; e.out0_ld = C;
    conv(Ly.ffn2, z, E.filt, e);
    ln(y, Ly.n2_g, Ly.n2_b, m, C, 0, 0, x);   // masked: the next layer's (and the heads') x * mask
  }
  // ---- mu = proj_m(x) * mask  (:407)
  {
    EpiParams e = epi_f32();
    e.mask = m; e.mask_out = 1;
    e.out0 = p; e.out0_ld = 80;
    conv(E.proj_m, x, C, e);
    emit([=](cudaStream_t s) { return launch_btc_to_bct(p, mux_s, B, 80, Lx, s); });
  }
  // ---- logw = proj_w(x.detach(), mask)  (:70-94, :409): conv -> ReLU -> LayerNorm, twice; 1x1 proj; mask
  {
    EpiParams e = epi_f32();
    e.out0 = d1; e.out0_ld = E.filt_dp;
    conv(E.dp_conv1, x, C, e);
    ln(d1, E.dp_g1, E.dp_b1, m, E.filt_dp, 1, 0, d2);
    e = epi_f32();
    e.out0 = d1; e.out0_ld = E.filt_dp;
    conv(E.dp_conv2, d2, E.filt_dp, e);
    ln(d1, E.dp_g2, E.dp_b2, m, E.filt_dp, 1, 0, d2);
    e = epi_f32();
    e.out0 = d1; e.out0_ld = 8;          // one real output channel; the epilogue stores groups of eight
    conv(E.dp_proj, d2, E.filt_dp, e);
    const unsigned g0 = grid1(rows);
    emit([=](cudaStream_t s) { take_col0_kernel<<<g0, 256, 0, s>>>(d1, 8, m, rows, logw_s); return cudaGetLastError(); });
  }
  make_room(encoder_cache_);
  it = encoder_cache_.emplace(key, std::move(prog)).first;
  }
  MB_CUDA(cudaMemcpyAsync(ids_s, ids_in, rows * 8, cudaMemcpyDeviceToDevice, st));
  MB_CUDA(cudaMemcpyAsync(lens_s, lengths_in, static_cast<size_t>(B) * 8, cudaMemcpyDeviceToDevice, st));
  if (spk_s) MB_CUDA(cudaMemcpyAsync(spk_s, spks_in, static_cast<size_t>(B) * E.spk_dim * 4, cudaMemcpyDeviceToDevice, st));
  run(*it->second, st);
  MB_CUDA(cudaMemcpyAsync(mu_x, mux_s, rows * 80 * 4, cudaMemcpyDeviceToDevice, st));
  MB_CUDA(cudaMemcpyAsync(logw, logw_s, rows * 4, cudaMemcpyDeviceToDevice, st));
  MB_CUDA(cudaMemcpyAsync(x_mask, m, rows * 4, cudaMemcpyDeviceToDevice, st));
  mark_busy(st);
}

}  // namespace mb
