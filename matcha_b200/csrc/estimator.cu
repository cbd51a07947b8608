// The U-Net velocity estimator (matcha/models/components/decoder.py:363-443) and the Euler loop around it
// (matcha/models/components/flow_matching.py:55-85) as a launch program over [B][T][C] workspaces.
#include <cmath>

#include "plan.h"

namespace mb {

namespace {
enum EstBuf { XIN, RAW, H1, RAW2, RES, LN, QKV, ATT, HID, FF, YA, YB, SKIP0, SKIP1, XSTATE, MU_NTC, MASK_T, MASK_H, SPK,
              STATS, TVALS, TEMB0, TEMB1, TPROJ, STATS_PART, STATS_CNT };

const HostTensor& need(const StateDict& sd, const std::string& k) {
  auto it = sd.find(k);
  if (it == sd.end()) throw Error("state_dict is missing '" + k + "'");
  return it->second;
}
void expect_shape(const StateDict& sd, const std::string& k, std::vector<long long> shape) {
  if (need(sd, k).shape != shape) {
    std::string s = "state_dict tensor '" + k + "' has shape (";
    for (auto v : need(sd, k).shape) s += std::to_string(v) + ",";
    s += "), expected (";
    for (auto v : shape) s += std::to_string(v) + ",";
    throw Error(s + ")");
  }
}
EpiParams epi() {
  EpiParams e{};
  e.scale = 1.f; e.out0_slope = 1.f; e.out1_slope = 1.f; e.resid_inv = 1.f; e.o_mul = 1;
  return e;
}
}  // namespace

void Plan::load_estimator(const StateDict& sd) {
  std::lock_guard<std::recursive_mutex> lk(mu_);
  const int cin = cfg.c_in, cpad = (cin + 63) / 64 * 64;
  expect_shape(sd, "time_mlp.linear_1.weight", {1024, cin});
  expect_shape(sd, "final_proj.weight", {80, 256, 1});
  EstimatorW& E = est_;
  auto up = [&](const std::string& k) { const HostTensor& t = need(sd, k); return upload_f32(t.data, t.numel()); };
  E.t_w1 = up("time_mlp.linear_1.weight"); E.t_b1 = up("time_mlp.linear_1.bias");
  E.t_w2 = up("time_mlp.linear_2.weight"); E.t_b2 = up("time_mlp.linear_2.bias");
  const char* rn[6] = {"down_blocks.0.0", "down_blocks.1.0", "mid_blocks.0.0", "mid_blocks.1.0", "up_blocks.0.0", "up_blocks.1.0"};
  const char* tn[6] = {"down_blocks.0.1.0", "down_blocks.1.1.0", "mid_blocks.0.1.0", "mid_blocks.1.1.0", "up_blocks.0.1.0", "up_blocks.1.1.0"};
  std::vector<float> wp(6 * 256 * 1024), bp(6 * 256);
  for (int r = 0; r < 6; ++r) {
    const std::string p = std::string(rn[r]) + ".";
    expect_shape(sd, p + "mlp.1.weight", {256, 1024});
    std::copy(need(sd, p + "mlp.1.weight").data, need(sd, p + "mlp.1.weight").data + 256 * 1024, wp.begin() + r * 256 * 1024);
    std::copy(need(sd, p + "mlp.1.bias").data, need(sd, p + "mlp.1.bias").data + 256, bp.begin() + r * 256);
    std::vector<std::pair<int, int>> segs;
    if (r == 0) segs = {{cin, cpad}};
    else if (r >= 4) segs = {{256, 256}, {256, 256}};   // cat[x, skip] (decoder.py:427)
    else segs = {{256, 256}};
    ResnetW& R = E.res[r];
    R.conv1 = pack_conv1d(need(sd, p + "block1.block.0.weight"), need(sd, p + "block1.block.0.bias").data, 1, segs);
    R.conv2 = pack_conv1d(need(sd, p + "block2.block.0.weight"), need(sd, p + "block2.block.0.bias").data, 1, {{256, 256}});
    R.res = pack_conv1d(need(sd, p + "res_conv.weight"), need(sd, p + "res_conv.bias").data, 1, segs);
    R.gn1_g = up(p + "block1.block.1.weight"); R.gn1_b = up(p + "block1.block.1.bias");
    R.gn2_g = up(p + "block2.block.1.weight"); R.gn2_b = up(p + "block2.block.1.bias");
  }
  E.t_wp = upload_f32(wp.data(), wp.size());
  E.t_bp = upload_f32(bp.data(), bp.size());
  for (int x = 0; x < 6; ++x) {
    const std::string p = std::string(tn[x]) + ".";
    TransformerW& X = E.tr[x];
    // fused QKV weight, rows [q*0.125 | k | v]; the softmax scale 64^-0.5 is exact in every dtype
    std::vector<float> qkv(384 * 256);
    const char* names[3] = {"attn1.to_q.weight", "attn1.to_k.weight", "attn1.to_v.weight"};
    for (int i = 0; i < 3; ++i) {
      expect_shape(sd, p + names[i], {128, 256});
      const float* src = need(sd, p + names[i]).data;
      for (int j = 0; j < 128 * 256; ++j) qkv[i * 128 * 256 + j] = src[j] * (i == 0 ? 0.125f : 1.f);
    }
    HostTensor tq{qkv.data(), {384, 256}};
    X.qkv = pack_conv1d(tq, nullptr, 1, {{256, 256}});
    X.out = pack_conv1d(need(sd, p + "attn1.to_out.0.weight"), need(sd, p + "attn1.to_out.0.bias").data, 1, {{128, 128}});
    X.ff1 = pack_conv1d(need(sd, p + "ff.net.0.proj.weight"), need(sd, p + "ff.net.0.proj.bias").data, 1, {{256, 256}});
    X.ff2 = pack_conv1d(need(sd, p + "ff.net.2.weight"), need(sd, p + "ff.net.2.bias").data, 1, {{1024, 1024}});
    X.ln1_g = up(p + "norm1.weight"); X.ln1_b = up(p + "norm1.bias");
    X.ln3_g = up(p + "norm3.weight"); X.ln3_b = up(p + "norm3.bias");
    std::vector<float> sa(1024), sb(1024);
    const float* al = need(sd, p + "ff.net.0.alpha").data;
    const float* be = need(sd, p + "ff.net.0.beta").data;
    for (int i = 0; i < 1024; ++i) { sa[i] = std::exp(al[i]); sb[i] = 1.f / (std::exp(be[i]) + 1e-9f); }
    X.snake_a = upload_f32(sa.data(), 1024);
    X.snake_invb = upload_f32(sb.data(), 1024);
  }
  E.down0 = pack_conv_stride2(need(sd, "down_blocks.0.2.conv.weight"), need(sd, "down_blocks.0.2.conv.bias").data, 256);
  E.down1 = pack_conv1d(need(sd, "down_blocks.1.2.weight"), need(sd, "down_blocks.1.2.bias").data, 1, {{256, 256}});
  E.up0 = pack_convT(need(sd, "up_blocks.0.2.conv.weight"), need(sd, "up_blocks.0.2.conv.bias").data, 2);
  E.up1 = pack_conv1d(need(sd, "up_blocks.1.2.weight"), need(sd, "up_blocks.1.2.bias").data, 1, {{256, 256}});
  E.final_conv = pack_conv1d(need(sd, "final_block.block.0.weight"), need(sd, "final_block.block.0.bias").data, 1, {{256, 256}});
  E.fgn_g = up("final_block.block.1.weight"); E.fgn_b = up("final_block.block.1.bias");
  E.final_proj = pack_conv1d(need(sd, "final_proj.weight"), need(sd, "final_proj.bias").data, 1, {{256, 256}}, 128);
  E.loaded = true;
  decode_cache_.clear();
}

void Plan::estimator_buffers(int B, int T) {
  const size_t es = esize(cfg.dtype);
  const size_t rows = static_cast<size_t>(B) * T;
  const int cpad = (cfg.c_in + 63) / 64 * 64;
  ensure(ws_est_[XIN], rows * cpad * es);
  reserve_split(rows * 1024);          // widest conv input: the feed-forward down projection (1024 channels)
  for (int b : {RAW, H1, RAW2, RES, LN, HID, YA, YB, SKIP0}) ensure(ws_est_[b], rows * 256 * es);
  ensure(ws_est_[SKIP1], rows / 2 * 256 * es);
  ensure(ws_est_[QKV], rows * 384 * es);
  ensure(ws_est_[ATT], rows * 128 * es);
  ensure(ws_est_[FF], rows * 1024 * es);
  ensure(ws_est_[XSTATE], rows * 80 * 4);
  ensure(ws_est_[MU_NTC], rows * 80 * 4);
  ensure(ws_est_[MASK_T], rows * 4);
  ensure(ws_est_[MASK_H], rows / 2 * 4);
  ensure(ws_est_[SPK], static_cast<size_t>(B) * 64 * 4);
  // GroupNorm statistics of the 13 Block1D convs: final sums, per-32-row partial sums and one arrival counter per conv
  // (conv.cuh, gn_stats_store / gn_stats_finish).  The counters start at zero (ensure() clears new allocations) and every
  // launch re-arms its own.
  ensure(ws_est_[STATS], static_cast<size_t>(13) * B * 16 * 4);
  ensure(ws_est_[STATS_PART], static_cast<size_t>(13) * B * 8 * ((T + 31) / 32) * 8);
  ensure(ws_est_[STATS_CNT], static_cast<size_t>(13) * B * 8 * 4);
}

// Appends one Decoder.forward to `prog`.  tproj: [.. ][6][256] time projections for this call (bstride 0 when t is a
// scalar).  euler: fuse x += dt*v into final_proj (flow_matching.py:79), else write v to v_out_ntc ([B][T][80] fp32).
void Plan::build_estimator_step(Program& prog, int B, int T, const float* tproj, int tproj_bstride, bool euler, float dt,
                                float* v_out_ntc) {
  const int dtype = cfg.dtype;
  const bool tc = tensor_core();
  const int Th = T / 2;
  const int cpad = (cfg.c_in + 63) / 64 * 64;
  auto buf = [&](int i) { return ws_est_[i].p; };
  const float* maskT = static_cast<const float*>(buf(MASK_T));
  const float* maskH = static_cast<const float*>(buf(MASK_H));
  float* stats_base = static_cast<float*>(buf(STATS));
  float2* part_base = static_cast<float2*>(buf(STATS_PART));
  int* cnt_base = static_cast<int*>(buf(STATS_CNT));
  int stat_slot = 0;
  auto& ops = prog.ops;
  auto add = [&](Launch l) { ops.push_back(std::move(l)); prog.kernels++; };
  // statistics of the next Block1D conv over Tn rows: fills the epilogue's fields, returns the [B][8][2] sums the
  // consumer reads.  Sums are written (not accumulated) in a fixed order, so nothing has to be cleared between calls.
  // Statistics whose consumer is gn_mish_kernel are left as per-slot partial sums (the consumer adds them): the producing conv
  // then has no ordered-sum tail.  Only with the tensor-core kernels: the FFMA kernel keeps its own finish (MB200_GN_FINISH=1 too).
  const int stats_by_consumer = (tc || split_mode()) && !std::getenv("MB200_GN_FINISH") ? 1 : 0;
  // ... for sequences of at most 16 slots (512 frames): every consumer CTA / thread re-adds the slots it needs, which at cfg4's
  // 94 slots made the GroupNorm + Mish pass 3.6 x slower (measured: 56.0 against 52.6 ms per step); longer sequences keep the finish
  // in the conv.
  constexpr int kConsumerSumSlots = 16;
  auto next_stats = [&](EpiParams& e, int Tn) {
    const int k = stat_slot++;
    e.stats_out = stats_base + static_cast<size_t>(k) * B * 16;
    e.stats_slots = (Tn + 31) / 32;
    e.stats_part = part_base + static_cast<size_t>(k) * B * 8 * ((T + 31) / 32);
    e.stats_cnt = cnt_base + static_cast<size_t>(k) * B * 8;
    return e.stats_out;
  };

  // e: the EpiParams of the conv that produced `raw` (its statistics are consumed here as per-slot partial sums when the
  // conv ran without its ordered-sum tail, see use_partials below)
  auto gn_mish = [&](const void* raw, const EpiParams& pe, const float* g, const float* bta, const float* tadd, int tbs,
                     const float* mask, void* out, int Tn) {
    const float* st = pe.stats_out;
    const float2* part = pe.stats_nofinish ? pe.stats_part : nullptr;
    const int slots = pe.stats_slots;
    add([=](cudaStream_t s) -> cudaError_t {
      if (dtype == MB200_F32) return launch_gn_mish<float>((const float*)raw, st, part, slots, g, bta, tadd, tbs, mask, (float*)out, B, Tn, s);
      if (dtype == MB200_BF16) return launch_gn_mish<__nv_bfloat16>((const __nv_bfloat16*)raw, st, part, slots, g, bta, tadd, tbs, mask, (__nv_bfloat16*)out, B, Tn, s);
      return launch_gn_mish<__half>((const __half*)raw, st, part, slots, g, bta, tadd, tbs, mask, (__half*)out, B, Tn, s);
    });
  };
  auto layernorm = [&](const void* in, const float* g, const float* bta, void* out, long rows) {
    add([=](cudaStream_t s) -> cudaError_t {
      if (dtype == MB200_F32) return launch_layernorm<float>((const float*)in, g, bta, (float*)out, rows, s);
      if (dtype == MB200_BF16) return launch_layernorm<__nv_bfloat16>((const __nv_bfloat16*)in, g, bta, (__nv_bfloat16*)out, rows, s);
      return launch_layernorm<__half>((const __half*)in, g, bta, (__half*)out, rows, s);
    });
  };
  auto attention = [&](const void* qkv, const float* mask, void* out, int L) { add(attention_op(qkv, mask, out, B, L)); };

  // ResnetBlock1D (decoder.py:46-61): in (1 or 2 sources, already masked) -> RES (unmasked), LN1(RES) -> LN
  auto resnet = [&](int r, Src s0, const Src* s1, int Tn, const float* mask, const TransformerW& X) {
    const ResnetW& R = est_.res[r];
    EpiParams e = epi();
    e.out0 = buf(RAW); e.out0_ld = 256; e.out0_f32 = dtype == MB200_F32;
    next_stats(e, Tn);
    e.stats_nofinish = stats_by_consumer && e.stats_slots <= kConsumerSumSlots;
    add(conv_op(R.conv1, s0, s1, B, Tn, e));
    gn_mish(buf(RAW), e, R.gn1_g, R.gn1_b, tproj + r * 256, tproj_bstride, mask, buf(H1), Tn);
    e = epi();
    e.out0 = buf(RAW2); e.out0_ld = 256; e.out0_f32 = dtype == MB200_F32;
    const float* st2 = next_stats(e, Tn);
    // the fused consumer adds the partial sums itself when that is cheap (one load per slot and group: up to 16 slots = 512 frames)
    e.stats_nofinish = (stats_by_consumer && rq_supported() && e.stats_slots <= kConsumerSumSlots) ? 1 : 0;
    add(conv_op(R.conv2, Src{buf(H1), 256, 256, Tn}, nullptr, B, Tn, e));
    if (rq_supported()) {   // res_conv + GroupNorm/Mish + LayerNorm1 + QKV projection: one kernel (RES and QKV written)
      add(resqkv_op(R.res, X.qkv, s0, s1, buf(RAW2), e, R.gn2_g, R.gn2_b, mask, buf(RES), X.ln1_g, X.ln1_b, buf(QKV), B, Tn));
      return;
    }
    e = epi();
    e.gn_raw = buf(RAW2); e.gn_ld = 256; e.gn_stats = st2; e.gn_gamma = R.gn2_g; e.gn_beta = R.gn2_b;
    e.gn_inv_count = 1.f / (32.f * Tn); e.mask = mask;
    e.out0 = buf(RES); e.out0_ld = 256; e.out0_f32 = dtype == MB200_F32;
    if (tc) { e.ln_out = buf(LN); e.ln_ld = 256; e.ln_g = X.ln1_g; e.ln_b = X.ln1_b; }
    add(conv_op(R.res, s0, s1, B, Tn, e));
    if (!tc) layernorm(buf(RES), X.ln1_g, X.ln1_b, buf(LN), static_cast<long>(B) * Tn);
  };
  // BasicTransformerBlock (transformer.py:262-316): RES/LN -> out (masked with this level's mask)
  auto transformer = [&](int x, int Tn, const float* mask, void* out) {
    const TransformerW& X = est_.tr[x];
    EpiParams e = epi();
    if (!rq_supported()) {   // otherwise the ResNet block's fused kernel has already written QKV
      e.out0 = buf(QKV); e.out0_ld = 384; e.out0_f32 = dtype == MB200_F32;
      add(conv_op(X.qkv, Src{buf(LN), 256, 256, Tn}, nullptr, B, Tn, e));
    }
    attention(buf(QKV), mask, buf(ATT), Tn);
    if (ff_supported() && fuse_ff >= 2) {   // out-projection + residual + LayerNorm + feed-forward + residual + mask: one kernel
      add(ff_op(X.ff1, X.ff2, X.snake_a, X.snake_invb, buf(ATT), buf(HID), mask, out, static_cast<long>(B) * Tn, &X.out, buf(RES),
                X.ln3_g, X.ln3_b));
      return;
    }
    e = epi();
    e.resid = buf(RES); e.resid_ld = 256; e.resid_f32 = dtype == MB200_F32;
    e.out0 = buf(HID); e.out0_ld = 256; e.out0_f32 = dtype == MB200_F32;
    if (tc) { e.ln_out = buf(LN); e.ln_ld = 256; e.ln_g = X.ln3_g; e.ln_b = X.ln3_b; }
    add(conv_op(X.out, Src{buf(ATT), 128, 128, Tn}, nullptr, B, Tn, e));
    if (!tc) layernorm(buf(HID), X.ln3_g, X.ln3_b, buf(LN), static_cast<long>(B) * Tn);
    if (ff_supported()) {   // feed-forward as one kernel: the 1024-wide hidden activation stays in shared memory
      add(ff_op(X.ff1, X.ff2, X.snake_a, X.snake_invb, buf(LN), buf(HID), mask, out, static_cast<long>(B) * Tn));
      return;
    }
    e = epi();
    e.snake_a = X.snake_a; e.snake_invb = X.snake_invb;
    e.out0 = buf(FF); e.out0_ld = 1024; e.out0_f32 = dtype == MB200_F32;
    add(conv_op(X.ff1, Src{buf(LN), 256, 256, Tn}, nullptr, B, Tn, e));
    e = epi();
    e.resid = buf(HID); e.resid_ld = 256; e.resid_f32 = dtype == MB200_F32;
    e.mask = mask; e.mask_out = 1;   // every consumer of a transformer output multiplies by this level's mask
    e.out0 = out; e.out0_ld = 256; e.out0_f32 = dtype == MB200_F32;
    add(conv_op(X.ff2, Src{buf(FF), 1024, 1024, Tn}, nullptr, B, Tn, e));
  };
  auto masked_conv = [&](const PackedW& W, Src s, int T_out, const float* mask_out, void* out) {
    EpiParams e = epi();
    e.mask = mask_out; e.mask_out = 1;
    e.out0 = out; e.out0_ld = 256; e.out0_f32 = dtype == MB200_F32;
    add(conv_op(W, s, nullptr, B, T_out, e));
  };

  // down 0 (T) -------------------------------------------------------------------------------------
  resnet(0, Src{buf(XIN), cpad, cpad, T}, nullptr, T, maskT, est_.tr[0]);
  transformer(0, T, maskT, buf(SKIP0));
  masked_conv(est_.down0, Src{buf(SKIP0), 512, 256, Th}, Th, maskH, buf(YA));   // [T/2][2C] view, stride-2 conv
  // down 1 (T/2)
  resnet(1, Src{buf(YA), 256, 256, Th}, nullptr, Th, maskH, est_.tr[1]);
  transformer(1, Th, maskH, buf(SKIP1));
  masked_conv(est_.down1, Src{buf(SKIP1), 256, 256, Th}, Th, maskH, buf(YB));
  // mid
  resnet(2, Src{buf(YB), 256, 256, Th}, nullptr, Th, maskH, est_.tr[2]);
  transformer(2, Th, maskH, buf(YA));
  resnet(3, Src{buf(YA), 256, 256, Th}, nullptr, Th, maskH, est_.tr[3]);
  transformer(3, Th, maskH, buf(YB));
  // up 0 (T/2 -> T)
  {
    Src skip{buf(SKIP1), 256, 256, Th};
    resnet(4, Src{buf(YB), 256, 256, Th}, &skip, Th, maskH, est_.tr[4]);
  }
  transformer(4, Th, maskH, buf(YA));
  masked_conv(est_.up0, Src{buf(YA), 256, 256, Th}, Th, maskT, buf(YB));        // ConvTranspose, output rows 2m+phase
  // up 1 (T)
  {
    Src skip{buf(SKIP0), 256, 256, T};
    resnet(5, Src{buf(YB), 256, 256, T}, &skip, T, maskT, est_.tr[5]);
  }
  transformer(5, T, maskT, buf(YA));
  masked_conv(est_.up1, Src{buf(YA), 256, 256, T}, T, maskT, buf(YB));
  // final block + projection (decoder.py:440-443)
  {
    EpiParams e = epi();
    e.out0 = buf(RAW); e.out0_ld = 256; e.out0_f32 = dtype == MB200_F32;
    next_stats(e, T);
    e.stats_nofinish = stats_by_consumer && e.stats_slots <= kConsumerSumSlots;
    add(conv_op(est_.final_conv, Src{buf(YB), 256, 256, T}, nullptr, B, T, e));
    gn_mish(buf(RAW), e, est_.fgn_g, est_.fgn_b, nullptr, 0, maskT, buf(H1), T);
    e = epi();
    e.mask = maskT; e.mask_out = 1;
    if (euler) {
      e.x_state = static_cast<float*>(buf(XSTATE)); e.x_ld = 80; e.dt = dt;
      e.out1 = buf(XIN); e.out1_ld = cpad;   // next step's masked x operand
    } else {
      e.out0 = v_out_ntc; e.out0_ld = 80; e.out0_f32 = 1;
    }
    add(conv_op(est_.final_proj, Src{buf(H1), 256, 256, T}, nullptr, B, T, e));
  }
}

}  // namespace mb

namespace mb {


// torch.linspace(0, 1, n+1) in fp32 (symmetric evaluation, as ATen does) and the (t, dt) recurrence of
// BASECFM.solve_euler (flow_matching.py:70-83): t accumulates by += dt, dt = t_span[step+1] - t.
void euler_schedule_host(int n_steps, float* ts, float* dts) {
  const int steps = n_steps + 1;
  std::vector<float> span(steps);
  const float step = (1.0f - 0.0f) / static_cast<float>(steps - 1);
  const int halfway = steps / 2;
  // ATen evaluates both halves with a fused multiply-add: one rounding of the exact start + step*i / end - step*k
  for (int i = 0; i < steps; ++i)
    span[i] = i < halfway ? static_cast<float>(static_cast<double>(step) * i)
                          : static_cast<float>(1.0 - static_cast<double>(step) * (steps - i - 1));
  float t = span[0], dt = span[1] - span[0];
  for (int s = 1; s <= n_steps; ++s) {
    ts[s - 1] = t;
    dts[s - 1] = dt;
    t = t + dt;
    if (s < n_steps) dt = span[s + 1] - t;
  }
}

// sinusoidal embedding -> Linear -> SiLU -> Linear -> (Mish -> six Linear(1024,256)) for n_t timesteps
void Plan::build_time_embed(Program& prog, const float* t_dev, int n_t) {
  const int cin = cfg.c_in;
  float* emb = static_cast<float*>(ws_est_[TEMB0].p);
  float* hid = static_cast<float*>(ws_est_[TEMB1].p);
  float* tproj = static_cast<float*>(ws_est_[TPROJ].p);
  const EstimatorW& E = est_;
  auto add = [&](Launch l) { prog.ops.push_back(std::move(l)); prog.kernels++; };
  add([=](cudaStream_t s) { return launch_sinusoidal(t_dev, n_t, cin, cin, emb, s); });   // rows of pitch c_in
  add([=](cudaStream_t s) { return launch_small_linear(emb, E.t_w1, E.t_b1, hid, n_t, cin, 1024, 1, s); });
  add([=](cudaStream_t s) { return launch_small_linear(hid, E.t_w2, E.t_b2, emb, n_t, 1024, 1024, 2, s); });  // stores Mish(temb)
  add([=](cudaStream_t s) { return launch_small_linear(emb, E.t_wp, E.t_bp, tproj, n_t, 1024, 6 * 256, 0, s); });
}

void Plan::estimator_forward(const float* x, const float* mask, const float* mu, const float* t, bool batched,
                             const float* spks, int B, int T, float* out, cudaStream_t st) {
  std::lock_guard<std::recursive_mutex> lk(mu_);
  if (!est_.loaded) throw Error("estimator weights are not loaded");
  if (T < 2 || T % 2) throw Error("T must be even (the reference pads to a multiple of 4: utils/model.py:14-20)");
  if ((cfg.c_in > 160) != (spks != nullptr)) throw Error("spks must be given exactly when c_in > 160");
  estimator_buffers(B, T);
  const int n_t = batched ? B : 1;
  ensure(ws_est_[TVALS], n_t * 4);
  ensure(ws_est_[TEMB0], static_cast<size_t>(n_t) * 1024 * 4);
  ensure(ws_est_[TEMB1], static_cast<size_t>(n_t) * 1024 * 4);
  ensure(ws_est_[TPROJ], static_cast<size_t>(n_t) * 6 * 256 * 4);
  DevBuf& vout = ws_io_[0];
  ensure(vout, static_cast<size_t>(B) * T * 80 * 4);
  float* xs = static_cast<float*>(ws_est_[XSTATE].p);
  float* mun = static_cast<float*>(ws_est_[MU_NTC].p);
  float* mT = static_cast<float*>(ws_est_[MASK_T].p);
  float* mH = static_cast<float*>(ws_est_[MASK_H].p);
  float* spk = static_cast<float*>(ws_est_[SPK].p);
  const int spk_dim = cfg.c_in - 160;
  const int cpad = (cfg.c_in + 63) / 64 * 64;
  Program prog;
  auto add = [&](Launch l) { prog.ops.push_back(std::move(l)); prog.kernels++; };
  prog.ops.push_back([=](cudaStream_t s) { return cudaMemcpyAsync(mT, mask, static_cast<size_t>(B) * T * 4, cudaMemcpyDeviceToDevice, s); });
  if (spks) prog.ops.push_back([=](cudaStream_t s) { return cudaMemcpyAsync(spk, spks, static_cast<size_t>(B) * spk_dim * 4, cudaMemcpyDeviceToDevice, s); });
  add([=](cudaStream_t s) { return launch_mask_half(mT, mH, B, T, s); });
  add([=](cudaStream_t s) { return launch_bct_to_btc(x, xs, B, 80, T, s); });
  add([=](cudaStream_t s) { return launch_bct_to_btc(mu, mun, B, 80, T, s); });
  {
    const int dtype = cfg.dtype;
    void* xin = ws_est_[XIN].p;
    add([=](cudaStream_t s) -> cudaError_t {
      if (dtype == MB200_F32) return launch_pack_xin<float>(xs, mun, spk, spk_dim, mT, (float*)xin, cpad, B, T, s);
      if (dtype == MB200_BF16) return launch_pack_xin<__nv_bfloat16>(xs, mun, spk, spk_dim, mT, (__nv_bfloat16*)xin, cpad, B, T, s);
      return launch_pack_xin<__half>(xs, mun, spk, spk_dim, mT, (__half*)xin, cpad, B, T, s);
    });
  }
  build_time_embed(prog, t, n_t);
  float* v = static_cast<float*>(vout.p);
  build_estimator_step(prog, B, T, static_cast<float*>(ws_est_[TPROJ].p), batched ? 6 * 256 : 0, false, 0.f, v);
  add([=](cudaStream_t s) { return launch_btc_to_bct(v, out, B, 80, T, s); });
  for (auto& op : prog.ops) MB_CUDA(op(st));
  launches += prog.kernels;
  mark_busy(st);
}

void Plan::cfm_decode(const float* mu, const float* mask, const float* z, const float* spks, int B, int T, int n_steps,
                      float* out, cudaStream_t st) {
  std::lock_guard<std::recursive_mutex> lk(mu_);
  if (!est_.loaded) throw Error("estimator weights are not loaded");
  if (T < 2 || T % 2) throw Error("T must be even (the reference pads to a multiple of 4: utils/model.py:14-20)");
  if (n_steps < 1 || n_steps > 1024) throw Error("n_steps out of range");
  if ((cfg.c_in > 160) != (spks != nullptr)) throw Error("spks must be given exactly when c_in > 160");
  estimator_buffers(B, T);
  ensure(ws_est_[TEMB0], static_cast<size_t>(n_steps) * 1024 * 4);
  ensure(ws_est_[TEMB1], static_cast<size_t>(n_steps) * 1024 * 4);
  ensure(ws_est_[TPROJ], static_cast<size_t>(n_steps) * 6 * 256 * 4);
  float* xs = static_cast<float*>(ws_est_[XSTATE].p);
  float* mun = static_cast<float*>(ws_est_[MU_NTC].p);
  float* mT = static_cast<float*>(ws_est_[MASK_T].p);
  float* mH = static_cast<float*>(ws_est_[MASK_H].p);
  float* spk = static_cast<float*>(ws_est_[SPK].p);
  const int spk_dim = cfg.c_in - 160;
  const int cpad = (cfg.c_in + 63) / 64 * 64;

  // boundary: user pointers -> plan workspaces (outside the cached program, so the program never sees them)
  MB_CUDA(cudaMemcpyAsync(mT, mask, static_cast<size_t>(B) * T * 4, cudaMemcpyDeviceToDevice, st));
  if (spks) MB_CUDA(cudaMemcpyAsync(spk, spks, static_cast<size_t>(B) * spk_dim * 4, cudaMemcpyDeviceToDevice, st));
  MB_CUDA(launch_bct_to_btc(z, xs, B, 80, T, st));
  MB_CUDA(launch_bct_to_btc(mu, mun, B, 80, T, st));
  launches += 2;

  auto key = std::make_tuple(B, T, n_steps, spks ? 1 : 0, 0LL);
  auto it = decode_cache_.find(key);
  if (it == decode_cache_.end()) {
    auto prog = std::make_unique<Program>();
    std::vector<float> ts(n_steps), dts(n_steps);
    euler_schedule_host(n_steps, ts.data(), dts.data());
    // The schedule is a constant of the program: its own device copy, uploaded once here.  (It used to be refreshed every
    // call from pageable host memory; the runtime drains the stream before such a copy, which serialised back-to-back
    // host-buffer calls against each other.)
    float* tv = upload_f32(ts.data(), ts.size());
    auto add = [&](Launch l) { prog->ops.push_back(std::move(l)); prog->kernels++; };
    add([=](cudaStream_t s) { return launch_mask_half(mT, mH, B, T, s); });
    const int dtype = cfg.dtype;
    void* xin = ws_est_[XIN].p;
    add([=](cudaStream_t s) -> cudaError_t {
      if (dtype == MB200_F32) return launch_pack_xin<float>(xs, mun, spk, spk_dim, mT, (float*)xin, cpad, B, T, s);
      if (dtype == MB200_BF16) return launch_pack_xin<__nv_bfloat16>(xs, mun, spk, spk_dim, mT, (__nv_bfloat16*)xin, cpad, B, T, s);
      return launch_pack_xin<__half>(xs, mun, spk, spk_dim, mT, (__half*)xin, cpad, B, T, s);
    });
    build_time_embed(*prog, tv, n_steps);
    const float* tproj = static_cast<const float*>(ws_est_[TPROJ].p);
    for (int s = 0; s < n_steps; ++s) build_estimator_step(*prog, B, T, tproj + static_cast<size_t>(s) * 6 * 256, 0, true, dts[s], nullptr);
    make_room(decode_cache_);
    it = decode_cache_.emplace(key, std::move(prog)).first;
  }
  run(*it->second, st);
  if (out) { MB_CUDA(launch_btc_to_bct(xs, out, B, 80, T, st)); launches += 1; }
  mark_busy(st);
}

}  // namespace mb
