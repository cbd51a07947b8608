// Plan infrastructure: allocation, TMA descriptor encoding, weight repacking, conv op construction, program
// execution / CUDA-graph capture.
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <cudaTypedefs.h>

#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <cstring>

#include "plan.h"

namespace mb {

static thread_local std::string g_last_error;
void set_last_error(const std::string& s) { g_last_error = s; }
const char* last_error_cstr() { return g_last_error.c_str(); }

static PFN_cuTensorMapEncodeTiled get_encode() {
  static PFN_cuTensorMapEncodeTiled fn = nullptr;
  if (!fn) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult qres;
    cudaError_t e = cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres);
    if (e != cudaSuccess || qres != cudaDriverEntryPointSuccess || !p)
      throw Error("cuTensorMapEncodeTiled is not available from the CUDA driver");
    fn = reinterpret_cast<PFN_cuTensorMapEncodeTiled>(p);
  }
  return fn;
}

Plan::Plan(const mb200_cfg& c) : cfg(c) {
  if (cfg.dtype != MB200_F32 && cfg.dtype != MB200_BF16 && cfg.dtype != MB200_F16) throw Error("unknown dtype");
  if (cfg.c_in != 160 && cfg.c_in != 224) throw Error("c_in must be 160 (LJSpeech) or 224 (VCTK, 64-dim speaker embedding)");
  int dev = 0, major = 0;
  MB_CUDA(cudaGetDevice(&dev));
  MB_CUDA(cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, dev));
  MB_CUDA(cudaDeviceGetAttribute(&num_sms, cudaDevAttrMultiProcessorCount, dev));
  if (const char* v = std::getenv("MB200_PDL")) use_pdl = std::atoi(v);
  if (const char* v = std::getenv("MB200_TILE_POLICY")) tile_policy = std::atoi(v);
  if (const char* v = std::getenv("MB200_KERNEL")) kernel_version = std::atoi(v);
  if (const char* v = std::getenv("MB200_FUSE_PAIRS")) fuse_pairs = std::atoi(v);
  if (const char* v = std::getenv("MB200_FUSE_FF")) fuse_ff = std::atoi(v);
  if (const char* v = std::getenv("MB200_FUSE_RQ")) fuse_rq = std::atoi(v);
  if (const char* v = std::getenv("MB200_ATTN")) attn_umma = std::strcmp(v, "mma") != 0;
  if (const char* v = std::getenv("MB200_FP32")) fp32_split = std::strcmp(v, "simt") != 0;
  if (const char* v = std::getenv("MB200_FF_PAIR")) ff_pair = std::atoi(v) != 0;
  if (const char* v = std::getenv("MB200_RES_PAIR")) res_pair = std::atoi(v) != 0;
  if (const char* v = std::getenv("MB200_RQ_PAIR")) rq_pair = std::atoi(v) != 0;
  if (const char* v = std::getenv("MB200_PAIR_RESIDENT")) pair_resident = std::atoi(v) != 0;
  if (const char* v = std::getenv("MB200_PAIR")) use_pair = std::atoi(v);   // 0 off, 1 N = 256 layers only, 2 (default) also the N = 128 layers
  if (major != 10) throw Error("matcha_b200 kernels are built for sm_100a only; device compute capability major = " + std::to_string(major));
}

Plan::~Plan() {
  for (auto& kv : decode_cache_) if (kv.second->graph) cudaGraphExecDestroy(kv.second->graph);
  for (auto& kv : vocoder_cache_) if (kv.second->graph) cudaGraphExecDestroy(kv.second->graph);
  for (auto& kv : encoder_cache_) if (kv.second->graph) cudaGraphExecDestroy(kv.second->graph);
  for (void* p : owned_) cudaFree(p);
  if (train_arena_) cudaFree(train_arena_);
  for (auto* arr : {ws_est_, ws_voc_, ws_io_}) (void)arr;
  for (auto& b : ws_est_) if (b.p) cudaFree(b.p);
  for (auto& b : ws_voc_) if (b.p) cudaFree(b.p);
  for (auto& b : ws_io_) if (b.p) cudaFree(b.p);
  for (auto& b : ws_enc_) if (b.p) cudaFree(b.p);
  if (mel_ready_) cudaEventDestroy(mel_ready_);
  for (auto e : wav_ready_) if (e) cudaEventDestroy(e);
  if (copy_stream_) cudaStreamDestroy(copy_stream_);
  if (host_stream_) cudaStreamDestroy(host_stream_);
  if (upload_stream_) cudaStreamDestroy(upload_stream_);
  for (auto& sl : slots_) {
    for (DevBuf* b : {&sl.mu, &sl.z, &sl.mask, &sl.spk, &sl.mel, &sl.wav}) if (b->p) cudaFree(b->p);
    for (cudaEvent_t e : {sl.uploaded, sl.computed, sl.copied}) if (e) cudaEventDestroy(e);
  }
  if (busy_) cudaEventDestroy(busy_);
}

void Plan::mark_busy(cudaStream_t st) {
  if (!busy_) MB_CUDA(cudaEventCreateWithFlags(&busy_, cudaEventDisableTiming));
  MB_CUDA(cudaEventRecord(busy_, st));
}

void* Plan::dalloc(size_t bytes) {
  void* p = nullptr;
  bytes = (bytes + 255) & ~size_t(255);
  MB_CUDA(cudaMalloc(&p, bytes));
  MB_CUDA(cudaMemset(p, 0, bytes));
  MB_CUDA(cudaDeviceSynchronize());   // the memset runs on the legacy default stream, which non-blocking streams do not wait for
  owned_.push_back(p);
  bytes_allocated += static_cast<long long>(bytes);
  return p;
}

float* Plan::upload_f32(const float* host, size_t n) {
  float* d = static_cast<float*>(dalloc(n * sizeof(float)));
  MB_CUDA(cudaMemcpy(d, host, n * sizeof(float), cudaMemcpyHostToDevice));
  return d;
}

void* Plan::ensure(DevBuf& b, size_t bytes) {
  if (b.bytes >= bytes) return b.p;
  // workspaces are referenced by cached programs: growing one invalidates them
  for (auto& kv : decode_cache_) if (kv.second->graph) cudaGraphExecDestroy(kv.second->graph);
  for (auto& kv : vocoder_cache_) if (kv.second->graph) cudaGraphExecDestroy(kv.second->graph);
  for (auto& kv : encoder_cache_) if (kv.second->graph) cudaGraphExecDestroy(kv.second->graph);
  decode_cache_.clear();
  vocoder_cache_.clear();
  encoder_cache_.clear();
  if (b.p) { MB_CUDA(cudaDeviceSynchronize()); cudaFree(b.p); bytes_allocated -= static_cast<long long>(b.bytes); }
  bytes = (bytes + 1023) & ~size_t(1023);
  MB_CUDA(cudaMalloc(&b.p, bytes));
  MB_CUDA(cudaMemset(b.p, 0, bytes));
  MB_CUDA(cudaDeviceSynchronize());   // as in dalloc: callers' streams (torch side streams, synthesise_host's private stream)
                                      // are not ordered against the legacy stream's memset; growth is rare
  b.bytes = bytes;
  bytes_allocated += static_cast<long long>(bytes);
  return b.p;
}

void Plan::reserve_split(size_t elems) {
  if (!split_mode()) return;
  ensure(ws_split_[0], elems * 2);
  ensure(ws_split_[1], elems * 2);
}

void* Plan::upload_operand(const std::vector<float>& host) {
  const size_t n = host.size();
  if (cfg.dtype == MB200_F32) return upload_f32(host.data(), n);
  std::vector<uint16_t> tmp(n);
  if (cfg.dtype == MB200_BF16) {
    for (size_t i = 0; i < n; ++i) { __nv_bfloat16 h = __float2bfloat16_rn(host[i]); std::memcpy(&tmp[i], &h, 2); }
  } else {
    for (size_t i = 0; i < n; ++i) { __half h = __float2half_rn(host[i]); std::memcpy(&tmp[i], &h, 2); }
  }
  void* d = dalloc(n * 2);
  MB_CUDA(cudaMemcpy(d, tmp.data(), n * 2, cudaMemcpyHostToDevice));
  return d;
}

// Activation view [B][rows][ld] (channels contiguous): 3-D map, box = kc channels x 128 rows x 1 batch element.
// Rows outside [0, rows) are zero-filled by TMA: that IS the conv's zero padding at the utterance edges.
CUtensorMap Plan::tmap_act(const void* p, int ld, int rows, int B, int kc, int box_rows, int dtype) const {
  CUtensorMap m;
  if (dtype < 0) dtype = cfg.dtype;
  const CUtensorMapDataType dt = dtype == MB200_BF16 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT16;
  cuuint64_t dims[3] = {static_cast<cuuint64_t>(ld), static_cast<cuuint64_t>(rows), static_cast<cuuint64_t>(B)};
  cuuint64_t strides[2] = {static_cast<cuuint64_t>(ld) * 2, static_cast<cuuint64_t>(ld) * 2 * rows};
  cuuint32_t box[3] = {static_cast<cuuint32_t>(kc), static_cast<cuuint32_t>(box_rows), 1};
  cuuint32_t estr[3] = {1, 1, 1};
  CUresult r = get_encode()(&m, dt, 3, const_cast<void*>(p), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                            kc == 64 ? CU_TENSOR_MAP_SWIZZLE_128B : CU_TENSOR_MAP_SWIZZLE_64B,
                            CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) throw Error("cuTensorMapEncodeTiled(activation) failed: " + std::to_string(static_cast<int>(r)));
  return m;
}

CUtensorMap Plan::tmap_w(const void* p, int K_total, int rows, int kc, int ntile, int dtype) const {
  CUtensorMap m;
  if (dtype < 0) dtype = cfg.dtype;
  const CUtensorMapDataType dt = dtype == MB200_BF16 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT16;
  cuuint64_t dims[2] = {static_cast<cuuint64_t>(K_total), static_cast<cuuint64_t>(rows)};
  cuuint64_t strides[1] = {static_cast<cuuint64_t>(K_total) * 2};
  cuuint32_t box[2] = {static_cast<cuuint32_t>(kc), static_cast<cuuint32_t>(ntile)};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = get_encode()(&m, dt, 2, const_cast<void*>(p), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                            kc == 64 ? CU_TENSOR_MAP_SWIZZLE_128B : CU_TENSOR_MAP_SWIZZLE_64B,
                            CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) throw Error("cuTensorMapEncodeTiled(weights) failed: " + std::to_string(static_cast<int>(r)));
  return m;
}

PackedW Plan::pack_conv(const std::function<float(int, int, int)>& weight, const float* bias, int n_real, int n_pad,
                        const std::vector<std::pair<int, int>>& segs, const std::vector<std::vector<int>>& taps_j) {
  PackedW W;
  W.n_phases = static_cast<int>(taps_j.size());
  W.n_taps = static_cast<int>(taps_j[0].size());
  if (W.n_taps > kMaxTaps || W.n_phases > kMaxPhases) throw Error("pack_conv: too many taps/phases");
  int pad_total = 0;
  bool all64 = true;
  for (auto& s : segs) { pad_total += s.second; if (s.second % 64) all64 = false; if (s.second % 32) throw Error("pack_conv: channel padding must be a multiple of 32"); }
  W.kc = all64 ? 64 : 32;
  W.N = n_pad;
  W.n_valid = n_real;
  W.K_total = W.n_taps * pad_total;
  for (auto& sgm : segs) W.k_real += W.n_taps * sgm.first;
  W.ntile = (n_pad % 256 == 0) ? 256 : (n_pad % 128 == 0) ? 128 : (n_pad % 64 == 0) ? 64 : 32;
  if (n_pad % 32) throw Error("pack_conv: N must be a multiple of 32");
  if (W.kc == 32 && W.ntile != 32) throw Error("pack_conv: 32-channel K chunks are only built for N tile 32");
  std::vector<float> host(static_cast<size_t>(W.n_phases) * n_pad * W.K_total, 0.f);
  for (int ph = 0; ph < W.n_phases; ++ph)
    for (int n = 0; n < n_real; ++n) {
      float* row = host.data() + (static_cast<size_t>(ph) * n_pad + n) * W.K_total;
      for (int tp = 0; tp < W.n_taps; ++tp) {
        int koff = tp * pad_total, cglob = 0;
        for (auto& s : segs) {
          for (int c = 0; c < s.first; ++c) row[koff + c] = weight(n, cglob + c, taps_j[ph][tp]);
          koff += s.second;
          cglob += s.first;
        }
      }
    }
  W.w = upload_operand(host);
  if (bias) {
    std::vector<float> bp(n_pad, 0.f);
    for (int n = 0; n < n_real; ++n) bp[n] = bias[n];
    W.bias = upload_f32(bp.data(), n_pad);
  }
  if (tensor_core()) W.tm = tmap_w(W.w, W.K_total, W.n_phases * n_pad, W.kc, W.ntile);
  if (split_mode()) {   // (hi | lo) bf16 halves of every weight row
    const size_t rows = static_cast<size_t>(W.n_phases) * n_pad, K = W.K_total;
    std::vector<uint16_t> hl(rows * 2 * K);
    for (size_t r = 0; r < rows; ++r)
      for (size_t k = 0; k < K; ++k) {
        const float x = host[r * K + k];
        const __nv_bfloat16 h = __float2bfloat16_rn(x), l = __float2bfloat16_rn(x - __bfloat162float(h));
        std::memcpy(&hl[r * 2 * K + k], &h, 2);
        std::memcpy(&hl[r * 2 * K + K + k], &l, 2);
      }
    W.w3 = dalloc(hl.size() * 2);
    MB_CUDA(cudaMemcpy(W.w3, hl.data(), hl.size() * 2, cudaMemcpyHostToDevice));
    W.tm3 = tmap_w(W.w3, 2 * W.K_total, W.n_phases * n_pad, W.kc, W.ntile, MB200_BF16);
  }
  return W;
}

// Conv1d weight (Cout, Cin, K), "same" padding, dilation dil.  Optionally scales the first `scaled_rows` output rows
// (used to fold the 1/sqrt(head_dim) attention scale into W_q).
PackedW Plan::pack_conv1d(const HostTensor& w, const float* bias, int dil, const std::vector<std::pair<int, int>>& segs,
                          int n_pad, float row_scale_first, int scaled_rows) {
  const int Cout = static_cast<int>(w.shape[0]), Cin = static_cast<int>(w.shape[1]);
  const int K = w.shape.size() > 2 ? static_cast<int>(w.shape[2]) : 1;
  int real = 0;
  for (auto& s : segs) real += s.first;
  if (real != Cin) throw Error("pack_conv1d: channel segments do not add up to Cin");
  if (!n_pad) n_pad = (Cout + 31) / 32 * 32;
  std::vector<std::vector<int>> taps(1);
  for (int j = 0; j < K; ++j) taps[0].push_back(j);
  const float* d = w.data;
  PackedW W = pack_conv(
      [&](int n, int c, int j) { return d[(static_cast<size_t>(n) * Cin + c) * K + j] * (n < scaled_rows ? row_scale_first : 1.f); },
      bias, Cout, n_pad, segs, taps);
  for (int j = 0; j < K; ++j) { W.tap_row[0][j] = (j - (K - 1) / 2) * dil; W.tap_ch[0][j] = 0; }
  return W;
}

// Conv1d(C, Cout, 3, stride 2, padding 1) (decoder.py:67) on the [T/2][2C] view of the input:
// out[t] = w0 x[2t-1] + w1 x[2t] + w2 x[2t+1]  ->  taps (row -1, ch C), (row 0, ch 0), (row 0, ch C).
PackedW Plan::pack_conv_stride2(const HostTensor& w, const float* bias, int C) {
  PackedW W = pack_conv1d(w, bias, 1, {{C, C}});
  W.tap_row[0][0] = -1; W.tap_ch[0][0] = C;
  W.tap_row[0][1] = 0;  W.tap_ch[0][1] = 0;
  W.tap_row[0][2] = 0;  W.tap_ch[0][2] = C;
  return W;
}

// ConvTranspose1d weight (Cin, Cout, K) with K == 2*stride, padding (K-stride)/2 (decoder.py:144; hifigan/models.py:158-169).
// Output phase ph (t = stride*m + ph) is a 2-tap conv: j = ph + p + stride*delta reads input row m - delta.
PackedW Plan::pack_convT(const HostTensor& w, const float* bias, int stride) {
  const int Cin = static_cast<int>(w.shape[0]), Cout = static_cast<int>(w.shape[1]), K = static_cast<int>(w.shape[2]);
  if (K != 2 * stride) throw Error("pack_convT: only kernel == 2*stride is supported");
  const int p = (K - stride) / 2;
  std::vector<std::vector<int>> taps(stride);
  int rows[kMaxPhases][2];
  for (int ph = 0; ph < stride; ++ph) {
    const int d0 = (ph + p < stride) ? 0 : -1;
    for (int q = 0; q < 2; ++q) {
      const int delta = d0 + q;
      taps[ph].push_back(ph + p + stride * delta);
      rows[ph][q] = -delta;
    }
  }
  const float* d = w.data;
  const int cpad = (Cin + 63) / 64 * 64;
  PackedW W = pack_conv([&](int n, int c, int j) { return d[(static_cast<size_t>(c) * Cout + n) * K + j]; }, bias, Cout,
                        (Cout + 31) / 32 * 32, {{Cin, cpad}}, taps);
  for (int ph = 0; ph < stride; ++ph)
    for (int q = 0; q < 2; ++q) { W.tap_row[ph][q] = rows[ph][q]; W.tap_ch[ph][q] = 0; }
  W.o_mul = stride;
  return W;
}

Launch Plan::conv_op(const PackedW& W, Src s0, const Src* s1, int B, int T_out, EpiParams e) {
  ConvLaunch L{};
  ConvGeom& g = L.g;
  // Split-precision fp32 mode: the fp32 source(s) are first copied into bf16 hi / lo buffers (one row = the concat of the
  // sources), and the GEMM runs on the bf16 tcgen05 kernel as a_hi w_hi + a_hi w_lo + a_lo w_hi with the fp32 epilogue.
  const bool split = split_mode() && W.w3 != nullptr;
  const Src f0 = s0;                       // the fp32 views
  const Src f1 = s1 ? *s1 : Src{nullptr, 0, 0, 0};
  const bool concat = s1 != nullptr;
  Src lo_src{};
  int split_c0 = 0, split_c1 = 0;
  const bool presplit = split && s0.lo != nullptr;
  if (presplit) {
    if (concat) throw Error("conv_op(split): a pre-split source cannot be concatenated");
    lo_src = Src{f0.lo, f0.ld, f0.ch, f0.rows};
    s1 = &lo_src;
    e.all_f32 = 1;
  } else if (split) {
    if (concat) {
      for (int tp = 0; tp < W.n_taps; ++tp)
        if (W.tap_ch[0][tp] != 0) throw Error("conv_op(split): concat sources with channel-offset taps");
      split_c0 = f0.ch; split_c1 = f1.ch;
    } else {
      split_c0 = f0.ld;                    // whole rows of the view: strided convs address channel bases inside them
    }
    const int chm = split_c0 + split_c1;
    const size_t need = static_cast<size_t>(B) * f0.rows * chm * 2;
    if (ws_split_[0].bytes < need || ws_split_[1].bytes < need) throw Error("conv_op(split): reserve_split() was not called for this shape");
    const int per_tap_ch = concat ? chm : f0.ch;
    s0 = Src{ws_split_[0].p, chm, per_tap_ch, f0.rows};
    lo_src = Src{ws_split_[1].p, chm, per_tap_ch, f0.rows};
    s1 = &lo_src;
    e.all_f32 = 1;
  }
  g.B = B; g.T_in = s0.rows; g.T_out = T_out;
  g.n_src = s1 ? 2 : 1;
  g.src_ch[0] = s0.ch; g.src_ld[0] = s0.ld;
  g.src_ch[1] = s1 ? s1->ch : 0; g.src_ld[1] = s1 ? s1->ld : 0;
  if (split) g.src_ch[1] = 0;              // hi and lo carry the SAME channels: one tap spans src_ch[0] of them
  g.n_taps = W.n_taps; g.n_phases = W.n_phases;
  std::memcpy(g.tap_row, W.tap_row, sizeof(g.tap_row));
  std::memcpy(g.tap_ch, W.tap_ch, sizeof(g.tap_ch));
  g.N = W.N; g.K_total = W.K_total;
  if (W.n_taps * (g.src_ch[0] + g.src_ch[1]) != W.K_total) throw Error("conv_op: sources do not match the packed weight K");
  if (s1 && s1->rows != s0.rows) throw Error("conv_op: concat sources differ in length");
  e.o_mul = W.o_mul; e.o_phase_step = 1; e.T_o = T_out * W.o_mul; e.n_valid = W.n_valid; e.bias = W.bias;
  L.e = e;
  L.a[0] = s0.p; L.a[1] = s1 ? s1->p : s0.p;
  // Tile shape.  Weights are streamed from L2 once per tile, so a 128-row x 256-channel tile re-reads them twice as
  // often per output element as a 256-row x 128-channel one; at batch 64 that stream is what saturates L2 in the
  // estimator and in HiFi-GAN stage 1.  Only the fused-LayerNorm layers need the whole 256-channel row in one tile.
  int ntile = W.ntile;
  if (ntile == 256 && !e.ln_out && T_out > 128 && tile_policy == 1) ntile = 128;
  L.w = W.w; L.ntile = ntile; L.kc = W.kc;
  if (e.ln_out && (W.N != 256 || !tensor_core())) throw Error("conv_op: fused LayerNorm needs the tensor-core kernel and N == 256");
  if (tensor_core() || split) {
    const int adt = split ? MB200_BF16 : -1;
    const int Kw = split ? 2 * W.K_total : W.K_total;   // K extent of a packed weight row
    L.tmA[0] = tmap_act(s0.p, s0.ld, s0.rows, B, W.kc, 128, adt);
    L.tmA[1] = s1 ? tmap_act(s1->p, s1->ld, s1->rows, B, W.kc, 128, adt) : L.tmA[0];
    // CTA pairs (cta_group::2) for the N = 256 layers with streamed weights: each CTA of a pair fetches half of the weight
    // rows (conv_gemm2.cu).  Needs one N tile and one phase (both CTAs of a pair then always want the same weights).
    const long w_total_bytes = static_cast<long>(W.n_phases) * W.N * Kw * 2;
    const bool pair = use_pair && kernel_version >= 2 && (ntile == 256 || (ntile == 128 && T_out >= 2048 && use_pair > 1)) && W.N == ntile && W.n_phases == 1 && W.kc == 64 &&
                      (split || w_total_bytes > 100 * 1024 || (pair_resident && T_out >= 2048)) && static_cast<long>(B) * ((T_out + 127) / 128) >= 2;
    const int wbox = pair ? ntile / 2 : ntile;
    if (split) L.tmW = (ntile == W.ntile && !pair) ? W.tm3 : tmap_w(W.w3, Kw, W.n_phases * W.N, W.kc, wbox, MB200_BF16);
    else L.tmW = (ntile == W.ntile && !pair) ? W.tm : tmap_w(W.w, W.K_total, W.n_phases * W.N, W.kc, wbox);
    // ---- persistent kernel geometry: A-load groups = (source, channel base); taps of a group share one halo tile
    ConvGeom2& g2 = L.g2;
    g2.B = B; g2.T_in = s0.rows; g2.T_out = T_out; g2.N = W.N; g2.n_phases = W.n_phases; g2.K_total = Kw;
    const int per_tap = g.src_ch[0] + g.src_ch[1];
    int ng = 0;
    // normal: one pass per source.  split: three passes = the products (hi, w_hi), (hi, w_lo), (lo, w_hi); the w_lo half
    // of a weight row starts at K_total.
    const int n_pass = split ? 3 : g.n_src;
    for (int pass = 0; pass < n_pass; ++pass) {
      const int s = split ? (pass == 2 ? 1 : 0) : pass;
      const int wbase = split ? (pass == 1 ? W.K_total : 0) : (s ? g.src_ch[0] : 0);
      const int chs = split ? g.src_ch[0] : g.src_ch[s];
      std::vector<int> bases;
      for (int tp = 0; tp < W.n_taps; ++tp)
        if (std::find(bases.begin(), bases.end(), W.tap_ch[0][tp]) == bases.end()) bases.push_back(W.tap_ch[0][tp]);
      for (int cb : bases) {
        if (ng >= kMaxGroups) throw Error("conv_op: too many A-load groups");
        g2.g_src[ng] = s; g2.g_ch[ng] = cb; g2.g_chunks[ng] = chs / W.kc;
        int nt = 0;
        for (int tp = 0; tp < W.n_taps; ++tp)
          if (W.tap_ch[0][tp] == cb) g2.g_wk[ng][nt++] = tp * per_tap + wbase;
        g2.g_ntaps[ng] = nt;
        for (int ph = 0; ph < W.n_phases; ++ph) {
          int rmin = 1 << 30, rmax = -(1 << 30);
          for (int tp = 0; tp < W.n_taps; ++tp)
            if (W.tap_ch[0][tp] == cb) { rmin = std::min(rmin, W.tap_row[ph][tp]); rmax = std::max(rmax, W.tap_row[ph][tp]); }
          g2.g_rowmin[ph][ng] = rmin;
          g2.g_extra[ph][ng] = (rmax - rmin + 7) / 8;
          if (g2.g_extra[ph][ng] > 8) throw Error("conv_op: tap span exceeds the halo capacity");
          int k = 0;
          for (int tp = 0; tp < W.n_taps; ++tp)
            if (W.tap_ch[0][tp] == cb) g2.g_shift[ph][ng][k++] = W.tap_row[ph][tp] - rmin;
        }
        ++ng;
      }
    }
    g2.n_groups = ng;
    // narrow-N layers (HiFi-GAN stages 2-4) process 2-4 blocks of 128 rows per tile to amortise per-tile latencies
    g2.mb = 1;
    if (T_out >= 2048) g2.mb = ntile == 32 ? 4 : (ntile <= 128 ? 2 : 1);
    else if (ntile == 128 && T_out > 128 && tile_policy == 1) g2.mb = 2;
    g2.m_tiles = (T_out + 128 * g2.mb - 1) / (128 * g2.mb);
    g2.n_tiles = W.N / ntile;
    g2.total_tiles = B * g2.m_tiles * g2.n_tiles * W.n_phases;
    // shared memory: A ring | weights | 1 KB alignment slack, 512 B barriers, 4.5 KB epilogue staging per epilogue warp,
    // LayerNorm exchange.  Sixteen epilogue warps (eg = 4) when the rings still get their depth: the epilogue is
    // latency-bound, so more warps in flight is what speeds it up; layers with large rings keep eight.
    const int a_bytes = (128 * g2.mb + 64) * W.kc * 2, w_tile = wbox * W.kc * 2;   // weight rows of one K chunk in ONE CTA
    g2.pair = pair ? 1 : 0;
    const long w_total = static_cast<long>(W.n_phases) * W.N * Kw * 2;
    g2.a_bytes = a_bytes;
    int max_taps = 1;
    for (int q = 0; q < ng; ++q) max_taps = std::max(max_taps, g2.g_ntaps[q]);
    auto layout = [&](int eg) -> bool {   // fills g2 ring fields + L.smem_bytes; false = rings too shallow for this eg
      const int tail_bytes = 1024 + 512 + 4 * eg * 4608 + 8192, budget = 227 * 1024 - tail_bytes;
      g2.eg = eg;
      // (split mode streams: the resident loader walks a weight row front to back, the three products revisit its hi half)
      // pair: each CTA keeps half of the resident weights, which buys a third activation stage for the HBM-bound K = 3 convs
      const long w_res = pair ? w_total / 2 : w_total;
      if (!split && w_total <= 100 * 1024 && (budget - w_res) / a_bytes >= 2) {
        g2.w_resident = 1;
        g2.wt = 1;
        g2.a_stages = std::min<long>(8, (budget - w_res) / a_bytes);
        g2.w_stages = 0;
        L.smem_bytes = static_cast<int>(g2.a_stages * a_bytes + w_res) + tail_bytes;
        return true;
      }
      g2.w_resident = 0;
      g2.a_stages = (ntile == 256 || g2.mb > 1) ? 2 : 3;
      // pair: half-size weight stages leave room; 1x1 convs (one 512-cycle MMA group per A stage) want the deeper A ring
      if (pair) { g2.a_stages = max_taps == 1 ? 4 : 2; if (const char* v = std::getenv("MB200_PAIR_A")) g2.a_stages = std::atoi(v); }
      // narrow tiles pair two taps per ring stage: half as many producer/issuer handshakes per tensor-pipe cycle
      g2.wt = (w_tile <= 16 * 1024 && max_taps >= 2) ? 2 : 1;
      g2.w_stages = std::min(8, (budget - g2.a_stages * a_bytes) / (g2.wt * w_tile));
      L.smem_bytes = g2.a_stages * a_bytes + g2.w_stages * g2.wt * w_tile + tail_bytes;
      return g2.w_stages >= 3;
    };
    // Epilogue warps: 8 by default.  The estimator layers whose epilogue carries the GroupNorm+Mish / LayerNorm / SnakeBeta /
    // Euler work are bound by that epilogue's latency, not by the tensor pipe: they get 12 (measured -7 % on the
    // estimator's GEMM time); 16 capped the kernel at 96 registers and spilled (measured slower, removed).  MB200_EG=2|3 overrides.
    const bool heavy_epi = e.gn_raw || e.ln_out || e.snake_a || e.stats_out || e.x_state;
    int want_eg = heavy_epi ? 3 : 2;
    if (const char* v = std::getenv("MB200_EG")) want_eg = std::atoi(v) == 3 ? 3 : 2;
    if (want_eg == 3 && !(g2.mb == 1 && ntile >= 128)) want_eg = 2;
    if (want_eg == 2 || !layout(want_eg)) {
      if (!layout(2) && g2.w_stages < 2) throw Error("conv_op: weight ring does not fit");
    }
    if (e.ln_out && g2.mb != 1) throw Error("conv_op: fused LayerNorm needs single-block tiles");
    if (e.ln_out && (!e.out0 || e.out0_f32)) throw Error("conv_op: fused LayerNorm re-reads out0 in the operand dtype");
    g2.dbg = 0;
    g2.dbg_out = nullptr;
    if (const char* v = std::getenv("MB200_DBG")) {
      static int op_index = 0;              // MB200_DBG_OP=n instruments only the n-th conv op built (developer probe)
      const char* only = std::getenv("MB200_DBG_OP");
      const bool hit = !only || std::atoi(only) == op_index;
      if (hit) {
        g2.dbg = std::atoi(v);
        g2.dbg_out = static_cast<long long*>(dalloc(148 * 16 * sizeof(long long)));
        probe_out = g2.dbg_out;
      }
      if (only && (hit || std::atoi(only) < 0)) {   // MB200_DBG_OP=-1 lists every op
        {
          std::fprintf(stderr, "[dbg op %d] B=%d T_out=%d K=%d N=%d taps=%d ntile=%d mb=%d eg=%d tiles=%d flags: gn=%d resid=%d snake=%d mask=%d stats=%d x=%d ln=%d out1=%d\n",
                       op_index, B, T_out, W.K_total, W.N, W.n_taps, ntile, g2.mb, want_eg, g2.total_tiles, e.gn_raw != nullptr, e.resid != nullptr,
                       e.snake_a != nullptr, e.mask != nullptr, e.stats_out != nullptr, e.x_state != nullptr, e.ln_out != nullptr, e.out1 != nullptr);
        }
      }
      ++op_index;
    }
    // flat MMA issue schedule (see ConvGeom2::sched)
    {
      int steps = 0;
      for (int q = 0; q < ng; ++q) steps += g2.g_chunks[q] * ((g2.g_ntaps[q] + 1) / 2 * 2);
      if (steps * W.n_phases > kMaxSched) throw Error("conv_op: MMA schedule exceeds kMaxSched");
      g2.n_steps = steps;
      const int row16 = W.kc * 2 / 16;
      for (int ph = 0; ph < W.n_phases; ++ph) {
        int s = 0;
        for (int q = 0; q < ng; ++q)
          for (int c = 0; c < g2.g_chunks[q]; ++c) {
            const int nt = g2.g_ntaps[q];
            for (int tp = 0; tp < nt; ++tp) {
              uint32_t en = static_cast<uint32_t>(g2.g_shift[ph][q][tp] * row16);
              if (g2.w_resident) {
                en |= static_cast<uint32_t>(g2.g_wk[q][tp] / W.kc + c) << 16;
              } else {
                en |= static_cast<uint32_t>(tp % g2.wt) << 16;
                if (tp % g2.wt == 0) en |= kSchedWFirst;
                if (tp % g2.wt == g2.wt - 1 || tp == nt - 1) en |= kSchedWLast;
              }
              if (tp == 0) en |= kSchedFirst;
              if (tp == nt - 1) en |= kSchedLast;
              g2.sched[ph * steps + s++] = en;
            }
            if (nt & 1) g2.sched[ph * steps + s++] = kSchedNull;
          }
      }
    }
    L.tmA8[0] = tmap_act(s0.p, s0.ld, s0.rows, B, W.kc, 8, adt);
    L.tmA8[1] = s1 ? tmap_act(s1->p, s1->ld, s1->rows, B, W.kc, 8, adt) : L.tmA8[0];
    L.grid = pair ? 2 * std::min((g2.total_tiles + 1) / 2, num_sms / 2) : std::min(g2.total_tiles, num_sms);
    L.pdl = use_pdl;
  }
  const int dtype = cfg.dtype;
  const bool tc = tensor_core();
  const bool v2 = kernel_version >= 2;
  const long split_rows = static_cast<long>(B) * f0.rows;
  void* const split_hi = ws_split_[0].p;
  void* const split_lo = ws_split_[1].p;
  Launch op([=](cudaStream_t st) -> cudaError_t {
    if (split && presplit) return launch_conv_umma2<__nv_bfloat16>(L, st);
    if (split) {
      cudaError_t err = launch_split_bf16x2(static_cast<const float*>(f0.p), f0.ld, split_c0, static_cast<const float*>(f1.p), f1.ld,
                                            split_c1, split_rows, split_hi, split_lo, st);
      if (err != cudaSuccess) return err;
      return launch_conv_umma2<__nv_bfloat16>(L, st);
    }
    if (dtype == MB200_F32) return launch_conv_simt<float>(L, st);
    if (dtype == MB200_BF16)
      return tc ? (v2 ? launch_conv_umma2<__nv_bfloat16>(L, st) : launch_conv_umma<__nv_bfloat16>(L, st))
                : launch_conv_simt<__nv_bfloat16>(L, st);
    return tc ? (v2 ? launch_conv_umma2<__half>(L, st) : launch_conv_umma<__half>(L, st)) : launch_conv_simt<__half>(L, st);
  });
  op.kind = 0;
  op.flops = 2.0 * B * T_out * W.n_phases * static_cast<double>(W.n_valid) * W.k_real;
  return op;
}

Launch Plan::pair_op(const PackedW& W1, const PackedW& W2, int dil, const void* in, int C, int B, int T, EpiParams e) {
  if (!pair_supported(C)) throw Error("pair_op: unsupported configuration");
  const int K = W1.n_taps;
  if (W2.n_taps != K || !(K & 1) || K > kMaxTaps) throw Error("pair_op: both convs need the same odd kernel size");
  if (W1.N != C || W2.N != C || W1.K_total != K * C || W2.K_total != K * C || W1.kc != C || W1.ntile != C || W2.kc != C ||
      W2.ntile != C || W1.n_phases != 1 || W2.n_phases != 1)
    throw Error("pair_op: packed weights do not match a C -> C conv pair");
  PairLaunch L{};
  L.C = C;
  L.mb = C == 64 ? 2 : 4;
  const int R = 128 * L.mb, rowb = C * 2, w_tile_full = C * rowb;
  PairGeom& g = L.g;
  g.B = B; g.T = T; g.K = K; g.dil = dil;
  g.S = R - (K - 1);
  g.tiles_per_b = (T + g.S - 1) / g.S;
  g.total_tiles = B * g.tiles_per_b;
  g.extra = ((K - 1) * dil + 7) / 8;
  if (g.extra > 8) throw Error("pair_op: tap span exceeds the halo capacity");
  g.a_bytes = ((R + 8 * g.extra) * rowb + 1023) / 1024 * 1024;
  g.mid_bytes = (R + 16) * rowb;
  g.bias1 = W1.bias;
  g.slope = 0.1f;
  g.dbg_out = nullptr;
  if (const char* v = std::getenv("MB200_PAIR_DBG")) {   // developer probe: instrument the n-th pair op built
    static int built = 0;
    if (built++ == std::atoi(v)) {
      g.dbg_out = static_cast<long long*>(dalloc(148 * 16 * sizeof(long long)));
      probe_out = g.dbg_out;
      std::fprintf(stderr, "[pair probe] C=%d K=%d dil=%d T=%d\n", C, K, dil, T);
    }
  }
  // shared memory: x ring | mid tile(s) | weights (resident, or a ring of two-tap stages) | barriers + E2 staging.
  // The mid tile is double-buffered when that still leaves two x stages and a usable weight ring: E1 of tile n+1 then
  // overlaps c2 of tile n (the short-kernel steps are bound by that chain).
  const int tail_bytes = 1024 + 512 + 8 * 4608;
  // CTA pairs (cta_group::2): each CTA of a 2-CTA cluster holds half of every weight tile's output channels
  L.cta_pair = (res_pair && g.total_tiles >= 2) ? 1 : 0;
  const int w_tile = L.cta_pair ? w_tile_full / 2 : w_tile_full;
  const int w_all = 2 * K * w_tile;
  int want_mid = 2;
  if (const char* v = std::getenv("MB200_PAIR_MID")) want_mid = std::atoi(v) == 1 ? 1 : 2;
  bool placed = false;
  for (int ms = want_mid; ms >= 1 && !placed; --ms) {
    const int budget = 227 * 1024 - tail_bytes - ms * g.mid_bytes;
    g.mid_stages = ms;
    if ((budget - w_all) / g.a_bytes >= 2) {
      g.w_resident = 1; g.w_stages = 0;
      g.a_stages = std::min(4, (budget - w_all) / g.a_bytes);
      L.smem_bytes = g.a_stages * g.a_bytes + ms * g.mid_bytes + w_all + tail_bytes;
      placed = true;
    } else {
      g.w_resident = 0; g.a_stages = 2;
      g.w_stages = std::min(8, (budget - 2 * g.a_bytes) / (2 * w_tile));
      L.smem_bytes = 2 * g.a_bytes + ms * g.mid_bytes + g.w_stages * 2 * w_tile + tail_bytes;
      placed = g.w_stages >= 3;
    }
  }
  if (!placed) throw Error("pair_op: weight ring does not fit");
  e.o_mul = 1; e.o_phase_step = 1; e.T_o = T; e.n_valid = C; e.bias = W2.bias;
  L.e = e;
  L.tmA = tmap_act(in, C, T, B, C);
  L.tmA8 = tmap_act(in, C, T, B, C, 8);
  L.tmW1 = L.cta_pair ? tmap_w(W1.w, W1.K_total, C, C, C / 2) : W1.tm;
  L.tmW2 = L.cta_pair ? tmap_w(W2.w, W2.K_total, C, C, C / 2) : W2.tm;
  L.grid = L.cta_pair ? 2 * std::min((g.total_tiles + 1) / 2, num_sms / 2) : std::min(g.total_tiles, num_sms);
  const int dtype = cfg.dtype;
  Launch op([L, dtype](cudaStream_t st) -> cudaError_t {
    return dtype == MB200_BF16 ? launch_resblock_pair<__nv_bfloat16>(L, st) : launch_resblock_pair<__half>(L, st);
  });
  op.kind = 0;
  op.flops = 2.0 * B * T * static_cast<double>(C) * (W1.k_real + W2.k_real);
  return op;
}

Launch Plan::ff_op(const PackedW& W1, const PackedW& W2, const float* snake_a, const float* snake_invb, const void* x,
                   void* resid, const float* mask, void* out, long rows, const PackedW* Wo, const void* res_in,
                   const float* ln_g, const float* ln_b) {
  if (!ff_supported()) throw Error("ff_op: unsupported configuration");
  if (W1.N != 1024 || W1.K_total != 256 || W2.N != 256 || W2.K_total != 1024 || W1.kc != 64 || W2.kc != 64 ||
      W1.n_phases != 1 || W2.n_phases != 1 || !W1.bias || !W2.bias)
    throw Error("ff_op: packed weights do not match a 256 -> 1024 -> 256 feed-forward");
  FFLaunch L{};
  L.g.rows = rows;
  L.g.total_tiles = static_cast<int>((rows + 127) / 128);
  L.g.bias1 = W1.bias; L.g.snake_a = snake_a; L.g.snake_invb = snake_invb; L.g.bias2 = W2.bias;
  L.g.mask = mask; L.g.resid = resid; L.g.out = out;
  L.g.dbg_out = nullptr;
  if (const char* v = std::getenv("MB200_FF_DBG")) {   // developer probe: instrument the n-th feed-forward op built
    static int built = 0;
    if (built++ == std::atoi(v)) {
      L.g.dbg_out = static_cast<long long*>(dalloc(148 * 16 * sizeof(long long)));
      probe_out = L.g.dbg_out;
      std::fprintf(stderr, "[ff probe] rows=%ld tiles=%d\n", rows, L.g.total_tiles);
    }
  }
  L.pair = (ff_pair && L.g.total_tiles >= 2) ? 1 : 0;
  L.tmW1 = tmap_w(W1.w, 256, 1024, 64, L.pair ? 64 : 128);
  L.tmW2 = tmap_w(W2.w, 1024, 256, 64, L.pair ? 128 : 256);
  if (Wo) {
    if (Wo->N != 256 || Wo->K_total != 128 || Wo->kc != 64 || Wo->n_phases != 1 || !Wo->bias || !res_in || !ln_g || !ln_b)
      throw Error("ff_op: the pre-stage needs a packed 128 -> 256 projection with bias, its residual and LayerNorm parameters");
    L.g.pre = 1; L.g.bias_o = Wo->bias; L.g.res_in = res_in; L.g.hid_out = resid; L.g.ln_g = ln_g; L.g.ln_b = ln_b;
    L.tmX = tmap_act(x, 128, static_cast<int>(rows), 1, 64);
    L.tmWo = tmap_w(Wo->w, 128, 256, 64, L.pair ? 128 : 256);
  } else {
    L.g.pre = 0;
    L.tmX = tmap_act(x, 256, static_cast<int>(rows), 1, 64);
    L.tmWo = L.tmW2;
  }
  L.smem_bytes = ff_fused_smem_bytes();
  L.grid = L.pair ? 2 * std::min((L.g.total_tiles + 1) / 2, num_sms / 2) : std::min(L.g.total_tiles, num_sms);
  const int dtype = cfg.dtype;
  Launch op([L, dtype](cudaStream_t st) -> cudaError_t {
    return dtype == MB200_BF16 ? launch_ff_fused<__nv_bfloat16>(L, st) : launch_ff_fused<__half>(L, st);
  });
  op.kind = 0;
  op.flops = 2.0 * static_cast<double>(rows) * (2.0 * 256.0 * 1024.0 + (Wo ? 256.0 * 128.0 : 0.0));
  return op;
}

Launch Plan::resqkv_op(const PackedW& Wres, const PackedW& Wqkv, Src s0, const Src* s1, const void* gn_raw,
                       const EpiParams& gn_producer, const float* gn_gamma, const float* gn_beta, const float* mask, void* res_out,
                       const float* ln_g, const float* ln_b, void* qkv_out, int B, int L) {
  if (!rq_supported()) throw Error("resqkv_op: unsupported configuration");
  const int kres = s0.ch + (s1 ? s1->ch : 0);
  if (Wres.N != 256 || Wres.K_total != kres || Wres.n_taps != 1 || Wres.kc != 64 || Wres.ntile != 256 || !Wres.bias ||
      Wqkv.N != 384 || Wqkv.K_total != 256 || Wqkv.kc != 64 || Wqkv.bias || s0.ch % 64 || (s1 && s1->ch % 64) || kres > 512)
    throw Error("resqkv_op: packed weights do not match a 1x1 res_conv (K <= 512) followed by a 256 -> 384 projection");
  if (s0.rows != L || (s1 && s1->rows != L)) throw Error("resqkv_op: sources must be dense [B][L][ld] buffers");
  RQLaunch Lq{};
  RQGeom& g = Lq.g;
  g.rows = static_cast<long>(B) * L; g.L = L;
  g.total_tiles = static_cast<int>((g.rows + 127) / 128);
  g.nk0 = s0.ch / 64; g.nk1 = s1 ? s1->ch / 64 : 0;
  g.bias = Wres.bias; g.gn_raw = gn_raw; g.gn_stats = gn_producer.stats_out; g.gn_gamma = gn_gamma; g.gn_beta = gn_beta;
  g.gn_part = gn_producer.stats_nofinish ? gn_producer.stats_part : nullptr;   // the conv left its partial sums un-added
  g.gn_slots = gn_producer.stats_slots;
  g.gn_inv_count = 1.f / (32.f * L);
  g.mask = mask; g.res_out = res_out; g.ln_g = ln_g; g.ln_b = ln_b; g.qkv_out = qkv_out;
  g.dbg_out = nullptr;
  if (const char* v = std::getenv("MB200_RQ_DBG")) {   // developer probe: instrument the n-th op built
    static int built = 0;
    if (built++ == std::atoi(v)) {
      g.dbg_out = static_cast<long long*>(dalloc(148 * 16 * sizeof(long long)));
      probe_out = g.dbg_out;
      std::fprintf(stderr, "[rq probe] rows=%ld tiles=%d nk=%d\n", g.rows, g.total_tiles, g.nk0 + g.nk1);
    }
  }
  Lq.tmA0 = tmap_act(s0.p, s0.ld, static_cast<int>(g.rows), 1, 64);
  Lq.tmA1 = s1 ? tmap_act(s1->p, s1->ld, static_cast<int>(g.rows), 1, 64) : Lq.tmA0;
  Lq.pair = (rq_pair && g.total_tiles >= 2) ? 1 : 0;   // CTA pairs: every weight box split between the two CTAs of a cluster
  Lq.tmWres = Lq.pair ? tmap_w(Wres.w, Wres.K_total, 256, 64, 128) : Wres.tm;
  Lq.tmWqkv = tmap_w(Wqkv.w, 256, 384, 64, Lq.pair ? 64 : 128);
  Lq.smem_bytes = resqkv_fused_smem_bytes();
  Lq.grid = Lq.pair ? 2 * std::min((g.total_tiles + 1) / 2, num_sms / 2) : std::min(g.total_tiles, num_sms);
  const int dtype = cfg.dtype;
  Launch op([Lq, dtype](cudaStream_t st) -> cudaError_t {
    return dtype == MB200_BF16 ? launch_resqkv_fused<__nv_bfloat16>(Lq, st) : launch_resqkv_fused<__half>(Lq, st);
  });
  op.kind = 0;
  op.flops = 2.0 * static_cast<double>(g.rows) * (256.0 * Wres.k_real + 384.0 * 256.0);
  return op;
}

Launch Plan::attention_op(const void* qkv, const float* mask, void* out, int B, int L) {
  const int dtype = cfg.dtype;
  const bool tc = tensor_core();
  Launch at;
  if (tc && attn_umma) {
    AttnLaunch A{};
    A.tmQKV = tmap_act(qkv, 384, L, B, 64);
    A.mask = mask; A.out = out; A.B = B; A.L = L;
    A.grid = std::min(B * 2 * ((L + 255) / 256), num_sms);
    A.dbg = nullptr;
    if (std::getenv("MB200_ATTN_DBG")) {
      A.dbg = static_cast<long long*>(dalloc(148 * 16 * sizeof(long long)));
      probe_out = A.dbg;
    }
    at = Launch([A, dtype](cudaStream_t s) -> cudaError_t {
      return dtype == MB200_BF16 ? launch_attention_umma<__nv_bfloat16>(A, s) : launch_attention_umma<__half>(A, s);
    });
  } else {
    at = Launch([=](cudaStream_t s) -> cudaError_t {
      if (dtype == MB200_F32) return launch_attention_simt<float>((const float*)qkv, mask, (float*)out, B, L, s);
      if (dtype == MB200_BF16) return tc ? launch_attention_mma<__nv_bfloat16>((const __nv_bfloat16*)qkv, mask, (__nv_bfloat16*)out, B, L, s)
                                         : launch_attention_simt<__nv_bfloat16>((const __nv_bfloat16*)qkv, mask, (__nv_bfloat16*)out, B, L, s);
      return tc ? launch_attention_mma<__half>((const __half*)qkv, mask, (__half*)out, B, L, s)
                : launch_attention_simt<__half>((const __half*)qkv, mask, (__half*)out, B, L, s);
    });
  }
  at.kind = 1;
  at.flops = 4.0 * B * 2 * static_cast<double>(L) * L * 64;   // QK^T + PV, dense over all L keys as the reference computes it
  return at;
}

void Plan::run(Program& prog, cudaStream_t st) {
  prog.last_use = ++use_clock_;
  if (profiling) {
    std::vector<cudaEvent_t> ev(prog.ops.size() + 1);
    for (auto& e : ev) MB_CUDA(cudaEventCreate(&e));
    MB_CUDA(cudaEventRecord(ev[0], st));
    for (size_t i = 0; i < prog.ops.size(); ++i) {
      MB_CUDA(prog.ops[i](st));
      MB_CUDA(cudaEventRecord(ev[i + 1], st));
    }
    MB_CUDA(cudaStreamSynchronize(st));
    for (size_t i = 0; i < prog.ops.size(); ++i) {
      float ms = 0.f;
      MB_CUDA(cudaEventElapsedTime(&ms, ev[i], ev[i + 1]));
      const Launch& l = prog.ops[i];
      double* slot = prof[l.region < kProfRegions ? l.region : kProfRegions - 1][l.kind < kProfKinds ? l.kind : kProfKinds - 1];
      slot[0] += 1; slot[1] += ms; slot[2] += l.flops;
    }
    for (auto& e : ev) cudaEventDestroy(e);
    launches += prog.kernels;
    return;
  }
  if (prog.graph) {
    MB_CUDA(cudaGraphLaunch(prog.graph, st));
    launches += prog.kernels;
    return;
  }
  for (auto& op : prog.ops) MB_CUDA(op(st));
  launches += prog.kernels;
  if (cfg.use_graph) {
    // The eager pass above already produced this call's result (and configured every kernel); record the same
    // launches into a graph for the following calls.  Capture happens on a private stream: the caller's may be
    // the legacy default stream, which cannot be captured.
    cudaStream_t cs;
    MB_CUDA(cudaStreamCreateWithFlags(&cs, cudaStreamNonBlocking));
    cudaGraph_t graph = nullptr;
    cudaError_t err = cudaStreamBeginCapture(cs, cudaStreamCaptureModeThreadLocal);
    if (err == cudaSuccess) {
      for (auto& op : prog.ops) { err = op(cs); if (err != cudaSuccess) break; }
      cudaError_t e2 = cudaStreamEndCapture(cs, &graph);
      if (err == cudaSuccess) err = e2;
    }
    if (err == cudaSuccess) err = cudaGraphInstantiate(&prog.graph, graph, 0);
    if (graph) cudaGraphDestroy(graph);
    cudaStreamDestroy(cs);
    if (err != cudaSuccess) { prog.graph = nullptr; throw Error(std::string("CUDA graph capture failed: ") + cudaGetErrorString(err)); }
  }
}

}  // namespace mb
