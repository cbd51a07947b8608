// BASECFM.compute_loss (matcha/models/components/flow_matching.py:87-118): y/u construction, Decoder.forward with
// per-sample t (decoder.py:363-443), the masked-sum MSE, and the full backward pass of the estimator, as an fp32
// launch sequence over the kernels of train_kernels.cuh.  Parameters and their gradients are device tensors in the
// reference layout, addressed in place.
#include <cmath>
#include <cstring>

#include "plan.h"
#include "train_kernels.cuh"

namespace mb {

namespace {
using trn::Gemm;

struct Ctx {
  cudaStream_t st;
  long long launches = 0;
  void check() { cudaError_t e = cudaGetLastError(); if (e != cudaSuccess) throw Error(std::string("training kernel launch failed: ") + cudaGetErrorString(e)); }
};
inline unsigned blocks256(long n) { return static_cast<unsigned>((n + 255) / 256); }

// MB200_TRAIN_GEMM = tf32x3 (default: tensor cores, error-compensated, fp32-level accuracy) | tf32 | ffma
int train_gemm_mode() {
  static const int mode = [] {
    const char* e = std::getenv("MB200_TRAIN_GEMM");
    if (!e) return 3;
    const std::string v = e;
    if (v == "ffma") return 0;
    if (v == "tf32") return 1;
    if (v == "tf32x3") return 3;
    throw Error("MB200_TRAIN_GEMM must be ffma, tf32 or tf32x3");
  }();
  return mode;
}

void gemm(Ctx& c, Gemm p, int nz) {
  if (p.M <= 0 || p.N <= 0 || p.K <= 0 || nz <= 0) return;
  if (p.nz2 <= 0) p.nz2 = 1;
  const int mode = train_gemm_mode();
  if (mode != 0 && p.M >= 32 && p.N >= 32) {
    static PerDeviceOnce configured;
    if (configured.needed()) {
      MB_CUDA(cudaFuncSetAttribute(trn::tgemm_kernel<3, 128>, cudaFuncAttributeMaxDynamicSharedMemorySize, trn::tgemm_smem_bytes(128)));
      MB_CUDA(cudaFuncSetAttribute(trn::tgemm_kernel<1, 128>, cudaFuncAttributeMaxDynamicSharedMemorySize, trn::tgemm_smem_bytes(128)));
      MB_CUDA(cudaFuncSetAttribute(trn::tgemm_kernel<3, 64>, cudaFuncAttributeMaxDynamicSharedMemorySize, trn::tgemm_smem_bytes(64)));
      MB_CUDA(cudaFuncSetAttribute(trn::tgemm_kernel<1, 64>, cudaFuncAttributeMaxDynamicSharedMemorySize, trn::tgemm_smem_bytes(64)));
      configured.mark();
    }
    // 64-row tiles only for short M: measured at cfg5's shape, trading 128-row tiles for twice the blocks on the
    // ~128-block GEMMs is slower (28.9 against 28.2 ms per step: B fragments are re-read per flop twice as often)
    const int tm = p.M <= 64 ? 64 : 128;
    dim3 grid((p.M + tm - 1) / tm, (p.N + trn::TN - 1) / trn::TN, nz);
    const int smem = trn::tgemm_smem_bytes(tm);
    if (tm == 128) {
      if (mode == 3) trn::tgemm_kernel<3, 128><<<grid, 256, smem, c.st>>>(p);
      else trn::tgemm_kernel<1, 128><<<grid, 256, smem, c.st>>>(p);
    } else {
      if (mode == 3) trn::tgemm_kernel<3, 64><<<grid, 256, smem, c.st>>>(p);
      else trn::tgemm_kernel<1, 64><<<grid, 256, smem, c.st>>>(p);
    }
  } else {
    dim3 grid((p.M + trn::BM - 1) / trn::BM, (p.N + trn::BN - 1) / trn::BN, nz);
    trn::sgemm_kernel<<<grid, 256, 0, c.st>>>(p);
  }
  c.check();
  c.launches++;
}

// A conv / linear layer in the reference layout.  GEMM row m of tap j reads input row in_step*m + in_off[j] and
// writes output row out_step*m + out_off[j]; weight element (n, c, j) sits at n*w_sn + c*w_sc + j.
struct ConvSpec {
  int Cin = 0, Cout = 0, K = 1;
  int in_step = 1, out_step = 1;
  int in_off[kMaxTaps] = {}, out_off[kMaxTaps] = {};
  long w_sn = 0, w_sc = 0;
  const float* w = nullptr; const float* b = nullptr;   // parameters (device)
  float* dw = nullptr; float* db = nullptr;             // gradients (device, may be null)
  float scale = 1.f;                                    // forward output scale folded nowhere: kept 1 (q scale is applied in the logits GEMM)
};
ConvSpec conv1d_spec(int Cin, int Cout, int K, int dil = 1) {   // Conv1d "same" padding (decoder.py:36,54,253)
  ConvSpec s; s.Cin = Cin; s.Cout = Cout; s.K = K;
  for (int j = 0; j < K; ++j) s.in_off[j] = (j - (K - 1) / 2) * dil;
  s.w_sn = static_cast<long>(Cin) * K; s.w_sc = K;
  return s;
}
ConvSpec conv_s2_spec(int C) {   // Conv1d(C, C, 3, stride 2, padding 1) (decoder.py:67)
  ConvSpec s = conv1d_spec(C, C, 3);
  s.in_step = 2;
  for (int j = 0; j < 3; ++j) s.in_off[j] = j - 1;
  return s;
}
ConvSpec convT_spec(int C) {     // ConvTranspose1d(C, C, 4, 2, 1) (decoder.py:144): weight (Cin, Cout, K)
  ConvSpec s; s.Cin = C; s.Cout = C; s.K = 4; s.out_step = 2;
  for (int j = 0; j < 4; ++j) s.out_off[j] = j - 1;
  s.w_sn = 4; s.w_sc = static_cast<long>(C) * 4;
  return s;
}
struct View { float* p; int ld; int rows; };   // [B][rows][ld]
struct Part { View x; int c0; int ch; };       // input source: channels [c0, c0+ch) of the layer's Cin

void tap_range(const ConvSpec& s, int j, int M, int T_in, int T_out, int& lo, int& hi) {
  lo = 0; hi = M;
  auto in = [&](int m) { return s.in_step * m + s.in_off[j]; };
  auto out = [&](int m) { return s.out_step * m + s.out_off[j]; };
  while (lo < hi && (in(lo) < 0 || out(lo) < 0)) ++lo;
  while (hi > lo && (in(hi - 1) >= T_in || out(hi - 1) >= T_out)) --hi;
}

// Y = conv(X) + bias.  M = GEMM rows per batch element.
void conv_forward(Ctx& c, const ConvSpec& s, const Part* parts, int n_parts, View y, int B, int M) {
  trn::fill_rows_kernel<<<blocks256(static_cast<long>(B) * y.rows * s.Cout), 256, 0, c.st>>>(y.p, s.b, static_cast<long>(B) * y.rows, s.Cout, y.ld);
  c.check(); c.launches++;
  for (int q = 0; q < n_parts; ++q) {
    const Part& pt = parts[q];
    for (int j = 0; j < s.K; ++j) {
      int lo, hi;
      tap_range(s, j, M, pt.x.rows, y.rows, lo, hi);
      Gemm g{};
      g.M = hi - lo; g.N = s.Cout; g.K = pt.ch;
      g.A = pt.x.p + static_cast<long>(s.in_step * lo + s.in_off[j]) * pt.x.ld; g.sAm = static_cast<long>(s.in_step) * pt.x.ld; g.sAk = 1;
      g.bA1 = static_cast<long>(pt.x.rows) * pt.x.ld;
      g.B = s.w + pt.c0 * s.w_sc + j; g.sBk = s.w_sc; g.sBn = s.w_sn;
      g.C = y.p + static_cast<long>(s.out_step * lo + s.out_off[j]) * y.ld; g.sCm = static_cast<long>(s.out_step) * y.ld; g.sCn = 1;
      g.bC1 = static_cast<long>(y.rows) * y.ld;
      g.nz2 = 1; g.alpha = 1.f; g.mode = 1;
      gemm(c, g, B);
    }
  }
}
// dW += dY^T X (per tap, reduced over the batch with atomics); db += colsum(dY)
void conv_wgrad(Ctx& c, const ConvSpec& s, const Part* parts, int n_parts, View dy, int B, int M) {
  if (s.db) {
    const long rows = static_cast<long>(B) * dy.rows;
    trn::colsum_kernel<<<static_cast<unsigned>((rows + 127) / 128), 256, 0, c.st>>>(dy.p, rows, s.Cout, dy.ld, s.db);
    c.check(); c.launches++;
  }
  if (!s.dw) return;
  for (int q = 0; q < n_parts; ++q) {
    const Part& pt = parts[q];
    for (int j = 0; j < s.K; ++j) {
      int lo, hi;
      tap_range(s, j, M, pt.x.rows, dy.rows, lo, hi);
      Gemm g{};
      g.M = s.Cout; g.N = pt.ch; g.K = hi - lo;
      g.A = dy.p + static_cast<long>(s.out_step * lo + s.out_off[j]) * dy.ld; g.sAm = 1; g.sAk = static_cast<long>(s.out_step) * dy.ld;
      g.bA1 = static_cast<long>(dy.rows) * dy.ld;
      g.B = pt.x.p + static_cast<long>(s.in_step * lo + s.in_off[j]) * pt.x.ld; g.sBk = static_cast<long>(s.in_step) * pt.x.ld; g.sBn = 1;
      g.bB1 = static_cast<long>(pt.x.rows) * pt.x.ld;
      g.C = s.dw + pt.c0 * s.w_sc + j; g.sCm = s.w_sn; g.sCn = s.w_sc;
      g.nz2 = 1; g.alpha = 1.f; g.mode = 2;
      gemm(c, g, B);
    }
  }
}
// dX[part] += dY W  (dx views must be zeroed by the caller unless accumulation is wanted)
void conv_dgrad(Ctx& c, const ConvSpec& s, const Part* dparts, int n_parts, View dy, int B, int M) {
  for (int q = 0; q < n_parts; ++q) {
    const Part& pt = dparts[q];
    if (!pt.x.p) continue;
    for (int j = 0; j < s.K; ++j) {
      int lo, hi;
      tap_range(s, j, M, pt.x.rows, dy.rows, lo, hi);
      Gemm g{};
      g.M = hi - lo; g.N = pt.ch; g.K = s.Cout;
      g.A = dy.p + static_cast<long>(s.out_step * lo + s.out_off[j]) * dy.ld; g.sAm = static_cast<long>(s.out_step) * dy.ld; g.sAk = 1;
      g.bA1 = static_cast<long>(dy.rows) * dy.ld;
      g.B = s.w + pt.c0 * s.w_sc + j; g.sBk = s.w_sn; g.sBn = s.w_sc;
      g.C = pt.x.p + static_cast<long>(s.in_step * lo + s.in_off[j]) * pt.x.ld; g.sCm = static_cast<long>(s.in_step) * pt.x.ld; g.sCn = 1;
      g.bC1 = static_cast<long>(pt.x.rows) * pt.x.ld;
      g.nz2 = 1; g.alpha = 1.f; g.mode = 1;
      gemm(c, g, B);
    }
  }
}

struct Arena {
  char* base = nullptr;
  size_t off = 0;
  float* take(size_t n_floats) {
    const size_t o = off;
    off += (n_floats * 4 + 255) & ~size_t(255);
    return base ? reinterpret_cast<float*>(base + o) : nullptr;
  }
};

struct ResSave { float *raw1, *st1, *h1, *raw2, *st2, *res; };
struct TrSave { float *ln1, *qkv, *P, *att, *hid, *ln3, *pre, *ff, *out; };
struct WS {
  float *x0, *yb, *emb, *pre1, *act1, *temb, *mtemb, *tproj, *maskH, *den;
  ResSave rs[6];
  TrSave ts[6];
  float *d0, *d1, *u0, *u1, *rawf, *stf, *hf, *v;
  // backward scratch
  float *G, *IA, *TA, *TB, *GFF, *GQKV, *GATT, *GP, *DSK0, *DSK1, *gsum, *dtproj, *dmt, *dx0;
};
void layout(Arena& a, WS& w, int B, int T, int cin) {
  const size_t rT = static_cast<size_t>(B) * T, rH = rT / 2;
  const int Th = T / 2;
  w.x0 = a.take(rT * cin); w.yb = a.take(rT * 80);
  w.emb = a.take(static_cast<size_t>(B) * cin); w.pre1 = a.take(B * 1024); w.act1 = a.take(B * 1024);
  w.temb = a.take(B * 1024); w.mtemb = a.take(B * 1024); w.tproj = a.take(B * 1536);
  w.maskH = a.take(rH); w.den = a.take(64);
  for (int i = 0; i < 6; ++i) {
    const bool full = i == 0 || i == 5;
    const size_t r = full ? rT : rH;
    const size_t L = full ? T : Th;
    ResSave& R = w.rs[i];
    R.raw1 = a.take(r * 256); R.st1 = a.take(B * 16); R.h1 = a.take(r * 256);
    R.raw2 = a.take(r * 256); R.st2 = a.take(B * 16); R.res = a.take(r * 256);
    TrSave& X = w.ts[i];
    X.ln1 = a.take(r * 256); X.qkv = a.take(r * 384); X.P = a.take(static_cast<size_t>(B) * 2 * L * L); X.att = a.take(r * 128);
    X.hid = a.take(r * 256); X.ln3 = a.take(r * 256); X.pre = a.take(r * 1024); X.ff = a.take(r * 1024); X.out = a.take(r * 256);
  }
  w.d0 = a.take(rH * 256); w.d1 = a.take(rH * 256); w.u0 = a.take(rT * 256); w.u1 = a.take(rT * 256);
  w.rawf = a.take(rT * 256); w.stf = a.take(B * 16); w.hf = a.take(rT * 256); w.v = a.take(rT * 80);
  w.G = a.take(rT * 256); w.IA = a.take(rT * 256); w.TA = a.take(rT * 256); w.TB = a.take(rT * 256);
  w.GFF = a.take(rT * 1024); w.GQKV = a.take(rT * 384); w.GATT = a.take(rT * 128);
  w.GP = a.take(static_cast<size_t>(B) * 2 * T * T); w.DSK0 = a.take(rT * 256); w.DSK1 = a.take(rH * 256);
  w.gsum = a.take(B * 16); w.dtproj = a.take(B * 1536); w.dmt = a.take(B * 1024); w.dx0 = a.take(rT * cin);
}

struct Params {
  std::map<std::string, std::pair<const float*, float*>> m;
  const float* p(const std::string& k) const {
    auto it = m.find(k);
    if (it == m.end() || !it->second.first) throw Error("compute_loss: parameter '" + k + "' was not given");
    return it->second.first;
  }
  float* g(const std::string& k) const { auto it = m.find(k); return it == m.end() ? nullptr : it->second.second; }
  void bind(ConvSpec& s, const std::string& name, bool bias = true) const {
    s.w = p(name + ".weight"); s.dw = g(name + ".weight");
    if (bias) { s.b = p(name + ".bias"); s.db = g(name + ".bias"); }
  }
};

struct Engine {
  Ctx c;
  const Params& P;
  WS& w;
  int B, T, cin;
  const float* maskT;
  float drop_p = 0.f;                   // transformer-block dropout (0 = eval): Plan::set_train_dropout
  unsigned long long drop_seed = 0;
  const char* rn[6] = {"down_blocks.0.0", "down_blocks.1.0", "mid_blocks.0.0", "mid_blocks.1.0", "up_blocks.0.0", "up_blocks.1.0"};
  const char* tn[6] = {"down_blocks.0.1.0", "down_blocks.1.1.0", "mid_blocks.0.1.0", "mid_blocks.1.1.0", "up_blocks.0.1.0", "up_blocks.1.1.0"};

  void zero(float* p, size_t n) { if (cudaMemsetAsync(p, 0, n * 4, c.st) != cudaSuccess) throw Error("cudaMemsetAsync failed"); }
  void mul_mask(float* x, const float* mask, long rows, int N) {
    trn::mul_mask_kernel<<<blocks256(rows * N), 256, 0, c.st>>>(x, mask, rows, N); c.check(); c.launches++;
  }
  // site = 2 * transformer block + {0: after attn1.to_out (transformer.py:200 -> to_out.1), 1: after the FF activation (:124)}
  void dropout(const float* in, float* out, long n, int site) {
    if (drop_p <= 0.f) { if (in != out) MB_CUDA(cudaMemcpyAsync(out, in, n * 4, cudaMemcpyDeviceToDevice, c.st)); return; }
    const unsigned int thresh = static_cast<unsigned int>(static_cast<double>(drop_p) * 4294967296.0);
    trn::dropout_kernel<<<blocks256(n), 256, 0, c.st>>>(in, out, n, drop_seed, static_cast<unsigned int>(site), thresh, 1.f / (1.f - drop_p));
    c.check(); c.launches++;
  }
  ConvSpec linear(const std::string& name, int in, int out, bool bias = true) {
    ConvSpec s = conv1d_spec(in, out, 1);
    P.bind(s, name, bias);
    return s;
  }

  // ---------------------------------------------------------------- GroupNorm + Mish block tail
  void gn_fwd(const float* raw, float* stats, const std::string& gn, const float* tadd, const float* mask, float* out, int L) {
    trn::gn_stats_kernel<<<dim3(8, B), 256, 0, c.st>>>(raw, L, stats); c.check();
    const long total = static_cast<long>(B) * L * 256;
    trn::gn_mish_fwd_kernel<<<blocks256(total), 256, 0, c.st>>>(raw, stats, P.p(gn + ".weight"), P.p(gn + ".bias"), tadd, 1536, mask, out, L, total);
    c.check(); c.launches += 2;
  }
  void gn_bwd(const float* dout, const float* raw, const float* stats, const std::string& gn, const float* mask, float* draw,
              float* dtadd, int L) {
    trn::gn_mish_bwd_reduce_kernel<<<dim3(8, B), 256, 0, c.st>>>(dout, raw, stats, P.p(gn + ".weight"), P.p(gn + ".bias"), mask, L, w.gsum,
                                                               P.g(gn + ".weight"), P.g(gn + ".bias"), dtadd, 1536);
    c.check();
    const long total = static_cast<long>(B) * L * 256;
    trn::gn_mish_bwd_apply_kernel<<<blocks256(total), 256, 0, c.st>>>(dout, raw, stats, w.gsum, P.p(gn + ".weight"), P.p(gn + ".bias"), mask, draw, L, total);
    c.check(); c.launches += 2;
  }
  void ln_fwd(const float* in, const std::string& ln, float* out, long rows) {
    trn::ln_fwd_kernel<<<static_cast<unsigned>((rows + 7) / 8), 256, 0, c.st>>>(in, P.p(ln + ".weight"), P.p(ln + ".bias"), out, rows);
    c.check(); c.launches++;
  }
  void ln_bwd(const float* dout, const float* in, const std::string& ln, float* din, long rows) {
    const unsigned grid = static_cast<unsigned>(std::min<long>((rows + 7) / 8, 592));
    trn::ln_bwd_kernel<<<grid, 256, 0, c.st>>>(dout, in, P.p(ln + ".weight"), din, P.g(ln + ".weight"), P.g(ln + ".bias"), rows);
    c.check(); c.launches++;
  }

  // ---------------------------------------------------------------- ResnetBlock1D (decoder.py:46-61)
  void resnet_fwd(int r, const Part* in, int n_in, int L, const float* mask) {
    const std::string p = std::string(rn[r]) + ".";
    int cin_total = 0;
    for (int q = 0; q < n_in; ++q) cin_total += in[q].ch;
    ResSave& R = w.rs[r];
    ConvSpec c1 = conv1d_spec(cin_total, 256, 3); P.bind(c1, p + "block1.block.0");
    ConvSpec c2 = conv1d_spec(256, 256, 3); P.bind(c2, p + "block2.block.0");
    ConvSpec rc = conv1d_spec(cin_total, 256, 1); P.bind(rc, p + "res_conv");
    conv_forward(c, c1, in, n_in, View{R.raw1, 256, L}, B, L);
    gn_fwd(R.raw1, R.st1, p + "block1.block.1", w.tproj + r * 256, mask, R.h1, L);
    Part h1{View{R.h1, 256, L}, 0, 256};
    conv_forward(c, c2, &h1, 1, View{R.raw2, 256, L}, B, L);
    gn_fwd(R.raw2, R.st2, p + "block2.block.1", nullptr, mask, w.TA, L);          // TA = block2 output
    conv_forward(c, rc, in, n_in, View{R.res, 256, L}, B, L);
    add(R.res, w.TA, static_cast<long>(B) * L * 256);
  }
  void add(float* dst, const float* src, long n) {
    // dst += src through the GEMM-free path: a 1-row "mask" trick is not worth a kernel; reuse sgemm with K = 1 would be
    // wasteful, so use cudaMemcpy-free axpy kernel below
    axpy_launch(dst, src, n);
  }
  void axpy_launch(float* dst, const float* src, long n);

  // upstream G = d(res) [L][256]; din[q] (zero-initialised or accumulating views) receive the input gradients
  void resnet_bwd(int r, const Part* in, const Part* din, int n_in, int L, const float* mask) {
    const std::string p = std::string(rn[r]) + ".";
    int cin_total = 0;
    for (int q = 0; q < n_in; ++q) cin_total += in[q].ch;
    ResSave& R = w.rs[r];
    ConvSpec c1 = conv1d_spec(cin_total, 256, 3); P.bind(c1, p + "block1.block.0");
    ConvSpec c2 = conv1d_spec(256, 256, 3); P.bind(c2, p + "block2.block.0");
    ConvSpec rc = conv1d_spec(cin_total, 256, 1); P.bind(rc, p + "res_conv");
    const size_t n = static_cast<size_t>(B) * L * 256;
    View g{w.G, 256, L};
    conv_wgrad(c, rc, in, n_in, g, B, L);
    conv_dgrad(c, rc, din, n_in, g, B, L);
    gn_bwd(w.G, R.raw2, R.st2, p + "block2.block.1", mask, w.TA, nullptr, L);       // TA = d raw2
    Part h1{View{R.h1, 256, L}, 0, 256};
    conv_wgrad(c, c2, &h1, 1, View{w.TA, 256, L}, B, L);
    zero(w.TB, n);
    Part dh1{View{w.TB, 256, L}, 0, 256};
    conv_dgrad(c, c2, &dh1, 1, View{w.TA, 256, L}, B, L);                              // TB = d h1
    gn_bwd(w.TB, R.raw1, R.st1, p + "block1.block.1", mask, w.TA, w.dtproj + r * 256, L);   // TA = d raw1
    conv_wgrad(c, c1, in, n_in, View{w.TA, 256, L}, B, L);
    conv_dgrad(c, c1, din, n_in, View{w.TA, 256, L}, B, L);
  }

  // ---------------------------------------------------------------- BasicTransformerBlock (transformer.py:243-316)
  void attention_fwd(TrSave& X, int L, const float* mask) {
    const long sP = static_cast<long>(L) * L;
    Gemm g{};
    g.M = L; g.N = L; g.K = 64; g.nz2 = 2; g.alpha = 0.125f; g.mode = 0;
    g.A = X.qkv; g.sAm = 384; g.sAk = 1; g.bA1 = static_cast<long>(L) * 384; g.bA2 = 64;
    g.B = X.qkv + 128; g.sBk = 1; g.sBn = 384; g.bB1 = static_cast<long>(L) * 384; g.bB2 = 64;
    g.C = X.P; g.sCm = L; g.sCn = 1; g.bC1 = 2 * sP; g.bC2 = sP;
    gemm(c, g, 2 * B);
    trn::softmax_fwd_kernel<<<static_cast<unsigned>(static_cast<long>(B) * 2 * L), 256, 0, c.st>>>(X.P, mask, L, 2 * L);
    c.check(); c.launches++;
    Gemm o{};
    o.M = L; o.N = 64; o.K = L; o.nz2 = 2; o.alpha = 1.f; o.mode = 0;
    o.A = X.P; o.sAm = L; o.sAk = 1; o.bA1 = 2 * sP; o.bA2 = sP;
    o.B = X.qkv + 256; o.sBk = 384; o.sBn = 1; o.bB1 = static_cast<long>(L) * 384; o.bB2 = 64;
    o.C = X.att; o.sCm = 128; o.sCn = 1; o.bC1 = static_cast<long>(L) * 128; o.bC2 = 64;
    gemm(c, o, 2 * B);
  }
  // GATT = d att -> GQKV = d qkv
  void attention_bwd(TrSave& X, int L) {
    const long sP = static_cast<long>(L) * L, sq = static_cast<long>(L) * 384, sa = static_cast<long>(L) * 128;
    Gemm g{};
    g.nz2 = 2; g.mode = 0;
    // dV = P^T dO
    g.M = L; g.N = 64; g.K = L; g.alpha = 1.f;
    g.A = X.P; g.sAm = 1; g.sAk = L; g.bA1 = 2 * sP; g.bA2 = sP;
    g.B = w.GATT; g.sBk = 128; g.sBn = 1; g.bB1 = sa; g.bB2 = 64;
    g.C = w.GQKV + 256; g.sCm = 384; g.sCn = 1; g.bC1 = sq; g.bC2 = 64;
    gemm(c, g, 2 * B);
    // dP = dO V^T
    g.M = L; g.N = L; g.K = 64;
    g.A = w.GATT; g.sAm = 128; g.sAk = 1; g.bA1 = sa; g.bA2 = 64;
    g.B = X.qkv + 256; g.sBk = 1; g.sBn = 384; g.bB1 = sq; g.bB2 = 64;
    g.C = w.GP; g.sCm = L; g.sCn = 1; g.bC1 = 2 * sP; g.bC2 = sP;
    gemm(c, g, 2 * B);
    trn::softmax_bwd_kernel<<<static_cast<unsigned>(static_cast<long>(B) * 2 * L), 256, 0, c.st>>>(X.P, w.GP, L);
    c.check(); c.launches++;
    // dQ = scale * dS K
    g.M = L; g.N = 64; g.K = L; g.alpha = 0.125f;
    g.A = w.GP; g.sAm = L; g.sAk = 1; g.bA1 = 2 * sP; g.bA2 = sP;
    g.B = X.qkv + 128; g.sBk = 384; g.sBn = 1; g.bB1 = sq; g.bB2 = 64;
    g.C = w.GQKV; g.sCm = 384; g.sCn = 1; g.bC1 = sq; g.bC2 = 64;
    gemm(c, g, 2 * B);
    // dK = scale * dS^T Q
    g.A = w.GP; g.sAm = 1; g.sAk = L;
    g.B = X.qkv; g.sBk = 384; g.sBn = 1;
    g.C = w.GQKV + 128;
    gemm(c, g, 2 * B);
  }
  void transformer_fwd(int x, const float* res, int L, const float* mask) {
    const std::string p = std::string(tn[x]) + ".";
    TrSave& X = w.ts[x];
    const long rows = static_cast<long>(B) * L;
    ln_fwd(res, p + "norm1", X.ln1, rows);
    Part ln1{View{X.ln1, 256, L}, 0, 256};
    const char* qn[3] = {"attn1.to_q", "attn1.to_k", "attn1.to_v"};
    for (int i = 0; i < 3; ++i) {
      ConvSpec s = linear(p + qn[i], 256, 128, false);
      conv_forward(c, s, &ln1, 1, View{X.qkv + 128 * i, 384, L}, B, L);
    }
    attention_fwd(X, L, mask);
    Part att{View{X.att, 128, L}, 0, 128};
    ConvSpec so = linear(p + "attn1.to_out.0", 128, 256);
    conv_forward(c, so, &att, 1, View{X.hid, 256, L}, B, L);
    if (drop_p > 0.f) dropout(X.hid, X.hid, rows * 256, 2 * x);
    axpy_launch(X.hid, res, rows * 256);
    ln_fwd(X.hid, p + "norm3", X.ln3, rows);
    Part ln3{View{X.ln3, 256, L}, 0, 256};
    ConvSpec f1 = linear(p + "ff.net.0.proj", 256, 1024);
    conv_forward(c, f1, &ln3, 1, View{X.pre, 1024, L}, B, L);
    trn::snake_fwd_kernel<<<blocks256(rows * 1024), 256, 0, c.st>>>(X.pre, P.p(p + "ff.net.0.alpha"), P.p(p + "ff.net.0.beta"), X.ff, rows * 1024);
    c.check(); c.launches++;
    if (drop_p > 0.f) dropout(X.ff, X.ff, rows * 1024, 2 * x + 1);   // X.ff = what ff.net.2 multiplies (and its wgrad reads)
    Part ff{View{X.ff, 1024, L}, 0, 1024};
    ConvSpec f2 = linear(p + "ff.net.2", 1024, 256);
    conv_forward(c, f2, &ff, 1, View{X.out, 256, L}, B, L);
    axpy_launch(X.out, X.hid, rows * 256);
    mul_mask(X.out, mask, rows, 256);
  }
  // upstream G = d out (consumer side, unmasked) -> G = d res
  void transformer_bwd(int x, const float* res, int L, const float* mask) {
    const std::string p = std::string(tn[x]) + ".";
    TrSave& X = w.ts[x];
    const long rows = static_cast<long>(B) * L;
    mul_mask(w.G, mask, rows, 256);
    View g{w.G, 256, L};
    Part ff{View{X.ff, 1024, L}, 0, 1024};
    ConvSpec f2 = linear(p + "ff.net.2", 1024, 256);
    conv_wgrad(c, f2, &ff, 1, g, B, L);
    zero(w.GFF, rows * 1024);
    Part dff{View{w.GFF, 1024, L}, 0, 1024};
    conv_dgrad(c, f2, &dff, 1, g, B, L);
    if (drop_p > 0.f) dropout(w.GFF, w.GFF, rows * 1024, 2 * x + 1);
    trn::snake_bwd_kernel<<<static_cast<unsigned>((rows + 63) / 64), 256, 0, c.st>>>(w.GFF, X.pre, P.p(p + "ff.net.0.alpha"), P.p(p + "ff.net.0.beta"),
                                                                                  P.g(p + "ff.net.0.alpha"), P.g(p + "ff.net.0.beta"), rows);
    c.check(); c.launches++;
    Part ln3{View{X.ln3, 256, L}, 0, 256};
    ConvSpec f1 = linear(p + "ff.net.0.proj", 256, 1024);
    conv_wgrad(c, f1, &ln3, 1, View{w.GFF, 1024, L}, B, L);
    zero(w.TA, rows * 256);
    Part dln3{View{w.TA, 256, L}, 0, 256};
    conv_dgrad(c, f1, &dln3, 1, View{w.GFF, 1024, L}, B, L);
    ln_bwd(w.TA, X.hid, p + "norm3", w.G, rows);                  // G = d hid (residual + LN path)
    Part att{View{X.att, 128, L}, 0, 128};
    ConvSpec so = linear(p + "attn1.to_out.0", 128, 256);
    View go = g;                                 // gradient at the out-projection's output: d hid through the dropout
    if (drop_p > 0.f) { dropout(w.G, w.TA, rows * 256, 2 * x); go = View{w.TA, 256, L}; }   // TA is free here (consumed by ln_bwd above)
    conv_wgrad(c, so, &att, 1, go, B, L);
    zero(w.GATT, rows * 128);
    Part datt{View{w.GATT, 128, L}, 0, 128};
    conv_dgrad(c, so, &datt, 1, go, B, L);
    attention_bwd(X, L);
    Part ln1{View{X.ln1, 256, L}, 0, 256};
    zero(w.TA, rows * 256);
    Part dln1{View{w.TA, 256, L}, 0, 256};
    const char* qn[3] = {"attn1.to_q", "attn1.to_k", "attn1.to_v"};
    for (int i = 0; i < 3; ++i) {
      ConvSpec s = linear(p + qn[i], 256, 128, false);
      View dq{w.GQKV + 128 * i, 384, L};
      conv_wgrad(c, s, &ln1, 1, dq, B, L);
      conv_dgrad(c, s, &dln1, 1, dq, B, L);
    }
    ln_bwd(w.TA, res, p + "norm1", w.G, rows);                    // G = d res
  }

  // masked conv between blocks: y = conv(x) * mask_out
  void mconv_fwd(ConvSpec s, const std::string& name, const float* x, int rows_in, float* y, int rows_out, int M, const float* mask_out) {
    P.bind(s, name);
    Part in{View{const_cast<float*>(x), 256, rows_in}, 0, 256};
    conv_forward(c, s, &in, 1, View{y, 256, rows_out}, B, M);
    mul_mask(y, mask_out, static_cast<long>(B) * rows_out, 256);
  }
  // dy (in place masked) -> dW, db, dx (accumulated into dx: caller zeroes or pre-loads it)
  void mconv_bwd(ConvSpec s, const std::string& name, const float* x, int rows_in, float* dy, int rows_out, int M, const float* mask_out, float* dx) {
    P.bind(s, name);
    mul_mask(dy, mask_out, static_cast<long>(B) * rows_out, 256);
    Part in{View{const_cast<float*>(x), 256, rows_in}, 0, 256};
    conv_wgrad(c, s, &in, 1, View{dy, 256, rows_out}, B, M);
    Part din{View{dx, 256, rows_in}, 0, 256};
    conv_dgrad(c, s, &din, 1, View{dy, 256, rows_out}, B, M);
  }
};

__global__ void axpy_kernel(float* __restrict__ dst, const float* __restrict__ src, long n) {
  const long i = blockIdx.x * 256L + threadIdx.x;
  if (i < n) dst[i] += src[i];
}
void Engine::axpy_launch(float* dst, const float* src, long n) {
  axpy_kernel<<<blocks256(n), 256, 0, c.st>>>(dst, src, n);
  c.check(); c.launches++;
}
__global__ void mask_half_train_kernel(const float* __restrict__ m, float* __restrict__ mh, int T, long total) {
  const long i = blockIdx.x * 256L + threadIdx.x;
  if (i >= total) return;
  const int Th = T / 2;
  mh[i] = m[(i / Th) * T + 2 * (i % Th)];   // mask[:, :, ::2] (decoder.py:407)
}

}  // namespace

void Plan::compute_loss_grad(const std::map<std::string, std::pair<const float*, float*>>& params, const float* x1,
                             const float* mask, const float* mu, const float* z, const float* t, const float* spks, int B,
                             int T, float sigma_min, float* loss_out, float* y_out, float* v_out, float* dmu, float* dspks,
                             bool backward, cudaStream_t st) {
  std::lock_guard<std::recursive_mutex> lk(mu_);
  if (T < 4 || T % 4) throw Error("compute_loss: T must be a multiple of 4 (utils/model.py:14-20 pads the batch to one)");
  const int cin = cfg.c_in;
  if ((cin > 160) != (spks != nullptr)) throw Error("spks must be given exactly when c_in > 160");
  WS w{};
  Arena probe;
  layout(probe, w, B, T, cin);
  if (probe.off > train_arena_bytes_) {
    if (train_arena_) { MB_CUDA(cudaDeviceSynchronize()); cudaFree(train_arena_); bytes_allocated -= static_cast<long long>(train_arena_bytes_); }
    train_arena_ = nullptr; train_arena_bytes_ = 0;
    MB_CUDA(cudaMalloc(&train_arena_, probe.off));
    train_arena_bytes_ = probe.off;
    bytes_allocated += static_cast<long long>(probe.off);
  }
  Arena arena;
  arena.base = static_cast<char*>(train_arena_);
  layout(arena, w, B, T, cin);

  Params P;
  P.m = params;
  Engine e{Ctx{st}, P, w, B, T, cin, mask};
  e.drop_p = train_dropout_p_;
  e.drop_seed = train_dropout_seed_;
  Ctx& c = e.c;
  const int Th = T / 2;
  const long rT = static_cast<long>(B) * T, rH = rT / 2;
  const float* maskT = mask;
  const float* maskH = w.maskH;

  // ---------------------------------------------------------------- forward
  {
    const long n = rT * 80;
    trn::make_y_kernel<<<blocks256(n), 256, 0, st>>>(x1, z, t, sigma_min, w.yb, 80L * T, n); c.check();
    if (y_out) MB_CUDA(cudaMemcpyAsync(y_out, w.yb, n * 4, cudaMemcpyDeviceToDevice, st));
    trn::pack_x0_kernel<<<blocks256(rT * cin), 256, 0, st>>>(w.yb, mu, spks, maskT, w.x0, cin, T, rT * cin); c.check();
    mask_half_train_kernel<<<blocks256(rH), 256, 0, st>>>(maskT, w.maskH, T, rH); c.check();
    c.launches += 3;
  }
  // time embedding (decoder.py:20-29, 105-117) and the six ResNet projections Linear(Mish(temb)) (decoder.py:49,58)
  ConvSpec tl1 = e.linear("time_mlp.linear_1", cin, 1024), tl2 = e.linear("time_mlp.linear_2", 1024, 1024);
  Part p_emb{View{w.emb, cin, B}, 0, cin}, p_act1{View{w.act1, 1024, B}, 0, 1024}, p_mt{View{w.mtemb, 1024, B}, 0, 1024};
  {
    MB_CUDA(launch_sinusoidal(t, B, cin, cin, w.emb, st)); c.launches++;
    conv_forward(c, tl1, &p_emb, 1, View{w.pre1, 1024, B}, 1, B);
    trn::unary_fwd_kernel<<<blocks256(B * 1024L), 256, 0, st>>>(w.pre1, w.act1, B * 1024L, 0); c.check();
    conv_forward(c, tl2, &p_act1, 1, View{w.temb, 1024, B}, 1, B);
    trn::unary_fwd_kernel<<<blocks256(B * 1024L), 256, 0, st>>>(w.temb, w.mtemb, B * 1024L, 1); c.check();
    c.launches += 2;
    for (int r = 0; r < 6; ++r) {
      ConvSpec s = e.linear(std::string(e.rn[r]) + ".mlp.1", 1024, 256);
      conv_forward(c, s, &p_mt, 1, View{w.tproj + r * 256, 1536, B}, 1, B);
    }
  }
  Part in0{View{w.x0, cin, T}, 0, cin};
  e.resnet_fwd(0, &in0, 1, T, maskT);
  e.transformer_fwd(0, w.rs[0].res, T, maskT);
  e.mconv_fwd(conv_s2_spec(256), "down_blocks.0.2.conv", w.ts[0].out, T, w.d0, Th, Th, maskH);
  Part in1{View{w.d0, 256, Th}, 0, 256};
  e.resnet_fwd(1, &in1, 1, Th, maskH);
  e.transformer_fwd(1, w.rs[1].res, Th, maskH);
  e.mconv_fwd(conv1d_spec(256, 256, 3), "down_blocks.1.2", w.ts[1].out, Th, w.d1, Th, Th, maskH);
  Part in2{View{w.d1, 256, Th}, 0, 256};
  e.resnet_fwd(2, &in2, 1, Th, maskH);
  e.transformer_fwd(2, w.rs[2].res, Th, maskH);
  Part in3{View{w.ts[2].out, 256, Th}, 0, 256};
  e.resnet_fwd(3, &in3, 1, Th, maskH);
  e.transformer_fwd(3, w.rs[3].res, Th, maskH);
  Part in4[2] = {{View{w.ts[3].out, 256, Th}, 0, 256}, {View{w.ts[1].out, 256, Th}, 256, 256}};   // cat[x, skip] (decoder.py:427)
  e.resnet_fwd(4, in4, 2, Th, maskH);
  e.transformer_fwd(4, w.rs[4].res, Th, maskH);
  e.mconv_fwd(convT_spec(256), "up_blocks.0.2.conv", w.ts[4].out, Th, w.u0, T, Th, maskT);
  Part in5[2] = {{View{w.u0, 256, T}, 0, 256}, {View{w.ts[0].out, 256, T}, 256, 256}};
  e.resnet_fwd(5, in5, 2, T, maskT);
  e.transformer_fwd(5, w.rs[5].res, T, maskT);
  e.mconv_fwd(conv1d_spec(256, 256, 3), "up_blocks.1.2", w.ts[5].out, T, w.u1, T, T, maskT);
  ConvSpec fc = conv1d_spec(256, 256, 3); P.bind(fc, "final_block.block.0");
  ConvSpec fp = conv1d_spec(256, 80, 1); P.bind(fp, "final_proj");
  Part p_u1{View{w.u1, 256, T}, 0, 256}, p_hf{View{w.hf, 256, T}, 0, 256};
  conv_forward(c, fc, &p_u1, 1, View{w.rawf, 256, T}, B, T);
  e.gn_fwd(w.rawf, w.stf, "final_block.block.1", nullptr, maskT, w.hf, T);
  conv_forward(c, fp, &p_hf, 1, View{w.v, 80, T}, B, T);
  e.mul_mask(w.v, maskT, rT, 80);
  if (v_out) { MB_CUDA(launch_btc_to_bct(w.v, v_out, B, 80, T, st)); c.launches++; }

  // ---------------------------------------------------------------- loss (flow_matching.py:115-117) and d v
  MB_CUDA(cudaMemsetAsync(w.den, 0, 8, st));
  trn::sum_kernel<<<64, 256, 0, st>>>(maskT, rT, w.den); c.check();
  trn::loss_kernel<<<256, 256, 0, st>>>(w.v, x1, z, maskT, w.den, sigma_min, T, rT * 80, w.GFF, w.den + 1); c.check();   // GFF = d v [B][T][80]
  c.launches += 2;
  MB_CUDA(cudaMemcpyAsync(loss_out, w.den + 1, 4, cudaMemcpyDeviceToDevice, st));
  if (!backward) { launches += c.launches; return; }

  // ---------------------------------------------------------------- backward
  View dv{w.GFF, 80, T};
  conv_wgrad(c, fp, &p_hf, 1, dv, B, T);
  e.zero(w.G, rT * 256);
  Part d_hf{View{w.G, 256, T}, 0, 256};
  conv_dgrad(c, fp, &d_hf, 1, dv, B, T);                                           // G = d hf
  e.gn_bwd(w.G, w.rawf, w.stf, "final_block.block.1", maskT, w.TA, nullptr, T);    // TA = d rawf
  conv_wgrad(c, fc, &p_u1, 1, View{w.TA, 256, T}, B, T);
  e.zero(w.IA, rT * 256);
  Part d_u1{View{w.IA, 256, T}, 0, 256};
  conv_dgrad(c, fc, &d_u1, 1, View{w.TA, 256, T}, B, T);                           // IA = d u1
  e.zero(w.G, rT * 256);
  e.mconv_bwd(conv1d_spec(256, 256, 3), "up_blocks.1.2", w.ts[5].out, T, w.IA, T, T, maskT, w.G);   // G = d ts5.out
  e.transformer_bwd(5, w.rs[5].res, T, maskT);                                     // G = d rs5.res
  e.zero(w.IA, rT * 256); e.zero(w.DSK0, rT * 256);
  Part din5[2] = {{View{w.IA, 256, T}, 0, 256}, {View{w.DSK0, 256, T}, 256, 256}};
  e.resnet_bwd(5, in5, din5, 2, T, maskT);                                         // IA = d u0, DSK0 = d skip0
  e.zero(w.G, rH * 256);
  e.mconv_bwd(convT_spec(256), "up_blocks.0.2.conv", w.ts[4].out, Th, w.IA, T, Th, maskT, w.G);     // G = d ts4.out
  e.transformer_bwd(4, w.rs[4].res, Th, maskH);
  e.zero(w.IA, rH * 256); e.zero(w.DSK1, rH * 256);
  Part din4[2] = {{View{w.IA, 256, Th}, 0, 256}, {View{w.DSK1, 256, Th}, 256, 256}};
  e.resnet_bwd(4, in4, din4, 2, Th, maskH);                                        // IA = d ts3.out, DSK1 = d skip1
  MB_CUDA(cudaMemcpyAsync(w.G, w.IA, rH * 256 * 4, cudaMemcpyDeviceToDevice, st));
  e.transformer_bwd(3, w.rs[3].res, Th, maskH);
  e.zero(w.IA, rH * 256);
  Part din3{View{w.IA, 256, Th}, 0, 256};
  e.resnet_bwd(3, &in3, &din3, 1, Th, maskH);                                      // IA = d ts2.out
  MB_CUDA(cudaMemcpyAsync(w.G, w.IA, rH * 256 * 4, cudaMemcpyDeviceToDevice, st));
  e.transformer_bwd(2, w.rs[2].res, Th, maskH);
  e.zero(w.IA, rH * 256);
  Part din2{View{w.IA, 256, Th}, 0, 256};
  e.resnet_bwd(2, &in2, &din2, 1, Th, maskH);                                      // IA = d d1
  e.mconv_bwd(conv1d_spec(256, 256, 3), "down_blocks.1.2", w.ts[1].out, Th, w.IA, Th, Th, maskH, w.DSK1);   // DSK1 += -> d ts1.out
  MB_CUDA(cudaMemcpyAsync(w.G, w.DSK1, rH * 256 * 4, cudaMemcpyDeviceToDevice, st));
  e.transformer_bwd(1, w.rs[1].res, Th, maskH);
  e.zero(w.IA, rH * 256);
  Part din1{View{w.IA, 256, Th}, 0, 256};
  e.resnet_bwd(1, &in1, &din1, 1, Th, maskH);                                      // IA = d d0
  e.mconv_bwd(conv_s2_spec(256), "down_blocks.0.2.conv", w.ts[0].out, T, w.IA, Th, Th, maskH, w.DSK0);      // DSK0 += -> d ts0.out
  MB_CUDA(cudaMemcpyAsync(w.G, w.DSK0, rT * 256 * 4, cudaMemcpyDeviceToDevice, st));
  e.transformer_bwd(0, w.rs[0].res, T, maskT);
  const bool want_dx0 = dmu || dspks;
  if (want_dx0) e.zero(w.dx0, rT * cin);
  Part din0{View{want_dx0 ? w.dx0 : nullptr, cin, T}, 0, cin};
  e.resnet_bwd(0, &in0, &din0, 1, T, maskT);
  if (want_dx0) {
    trn::unpack_dx0_kernel<<<B, 256, 0, st>>>(w.dx0, maskT, dmu, dspks, cin, T); c.check(); c.launches++;
  }
  // time MLP backward: dtproj [B][6][256] was written by the six ResNet blocks
  e.zero(w.dmt, B * 1024);
  Part d_mt{View{w.dmt, 1024, B}, 0, 1024};
  for (int r = 0; r < 6; ++r) {
    ConvSpec s = e.linear(std::string(e.rn[r]) + ".mlp.1", 1024, 256);
    View dtp{w.dtproj + r * 256, 1536, B};
    conv_wgrad(c, s, &p_mt, 1, dtp, 1, B);
    conv_dgrad(c, s, &d_mt, 1, dtp, 1, B);
  }
  trn::unary_bwd_kernel<<<blocks256(B * 1024L), 256, 0, st>>>(w.dmt, w.temb, B * 1024L, 1); c.check();   // dmt = d temb
  conv_wgrad(c, tl2, &p_act1, 1, View{w.dmt, 1024, B}, 1, B);
  e.zero(w.TA, B * 1024);
  Part d_act1{View{w.TA, 1024, B}, 0, 1024};
  conv_dgrad(c, tl2, &d_act1, 1, View{w.dmt, 1024, B}, 1, B);
  trn::unary_bwd_kernel<<<blocks256(B * 1024L), 256, 0, st>>>(w.TA, w.pre1, B * 1024L, 0); c.check();    // TA = d pre1
  conv_wgrad(c, tl1, &p_emb, 1, View{w.TA, 1024, B}, 1, B);
  c.launches += 2;
  launches += c.launches;
}

}  // namespace mb
