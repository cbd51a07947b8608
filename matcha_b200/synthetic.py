"""Deterministic synthetic (random-init) state_dicts with the reference key layout, for benchmarks and tests.

Key names / shapes follow SURVEY.md Appendix A (decoder.py:221-313, hifigan/models.py:149-177).
Every parameter that the reference initialises to a constant (GroupNorm/LayerNorm affine, biases,
SnakeBeta alpha/beta: decoder.py:345-361, transformer.py:53-54) is randomised here so that parity
tests exercise it (SURVEY.md section 8(c) protocol).  torch's CPU generator makes this reproducible
on the GPU box without shipping 100 MB of weights.
"""
import math

import torch


def _g(seed):
    g = torch.Generator(device="cpu")
    g.manual_seed(seed)
    return g


def estimator_state_dict(c_in=160, n_feats=80, seed=1234, prefix="estimator."):
    g = _g(seed)
    sd = {}

    def lin(name, out_f, in_f, bias=True, gain=1.0):
        sd[name + ".weight"] = torch.randn(out_f, in_f, generator=g) * (gain * math.sqrt(2.0 / in_f))
        if bias:
            sd[name + ".bias"] = torch.randn(out_f, generator=g) * 0.1

    def conv(name, out_c, in_c, k, gain=1.0):
        sd[name + ".weight"] = torch.randn(out_c, in_c, k, generator=g) * (gain * math.sqrt(2.0 / (in_c * k)))
        sd[name + ".bias"] = torch.randn(out_c, generator=g) * 0.1

    def norm(name, c):
        sd[name + ".weight"] = 1.0 + 0.2 * torch.randn(c, generator=g)
        sd[name + ".bias"] = 0.1 * torch.randn(c, generator=g)

    def resnet(name, cin):
        lin(name + ".mlp.1", 256, 1024, gain=0.7)
        conv(name + ".block1.block.0", 256, cin, 3)
        norm(name + ".block1.block.1", 256)
        conv(name + ".block2.block.0", 256, 256, 3)
        norm(name + ".block2.block.1", 256)
        conv(name + ".res_conv", 256, cin, 1, gain=0.7)

    def transformer(name):
        norm(name + ".norm1", 256)
        for p in ("to_q", "to_k", "to_v"):
            lin(name + ".attn1." + p, 128, 256, bias=False, gain=1.2)
        lin(name + ".attn1.to_out.0", 256, 128, gain=0.7)
        norm(name + ".norm3", 256)
        sd[name + ".ff.net.0.alpha"] = 0.3 * torch.randn(1024, generator=g)
        sd[name + ".ff.net.0.beta"] = 0.3 * torch.randn(1024, generator=g)
        lin(name + ".ff.net.0.proj", 1024, 256)
        lin(name + ".ff.net.2", 256, 1024, gain=0.7)

    lin("time_mlp.linear_1", 1024, c_in)
    lin("time_mlp.linear_2", 1024, 1024)
    for i, cin in enumerate((c_in, 256)):
        resnet(f"down_blocks.{i}.0", cin)
        transformer(f"down_blocks.{i}.1.0")
    conv("down_blocks.0.2.conv", 256, 256, 3)
    conv("down_blocks.1.2", 256, 256, 3)
    for i in range(2):
        resnet(f"mid_blocks.{i}.0", 256)
        transformer(f"mid_blocks.{i}.1.0")
    for i in range(2):
        resnet(f"up_blocks.{i}.0", 512)
        transformer(f"up_blocks.{i}.1.0")
    # ConvTranspose1d weight layout is (Cin, Cout, K) (decoder.py:144)
    sd["up_blocks.0.2.conv.weight"] = torch.randn(256, 256, 4, generator=g) * math.sqrt(2.0 / (256 * 2))
    sd["up_blocks.0.2.conv.bias"] = torch.randn(256, generator=g) * 0.1
    conv("up_blocks.1.2", 256, 256, 3)
    conv("final_block.block.0", 256, 256, 3)
    norm("final_block.block.1", 256)
    conv("final_proj", n_feats, 256, 1, gain=0.7)
    return {prefix + k: v for k, v in sd.items()}


def generator_state_dict(seed=1234, weight_norm=False, gain=1.0):
    """HiFi-GAN v1 generator (hifigan/config.py:2-15).  weight_norm=True emits weight_g/weight_v
    (the on-disk checkpoint layout, cli.py:87); False emits the post-remove_weight_norm layout."""
    g = _g(seed)
    sd = {}

    def put(name, w, nbias):
        if weight_norm:
            sd[name + ".weight_g"] = w.flatten(1).norm(dim=1).view(-1, 1, 1) * (1.0 + 0.1 * torch.randn(
                w.shape[0], 1, 1, generator=g))
            sd[name + ".weight_v"] = w * (0.5 + torch.rand(w.shape[0], 1, 1, generator=g))
        else:
            sd[name + ".weight"] = w
        sd[name + ".bias"] = torch.randn(nbias, generator=g) * 0.05

    put("conv_pre", torch.randn(512, 80, 7, generator=g) * (gain * 0.18 * math.sqrt(1.0 / (80 * 7))), 512)
    ch = 512
    for i, (u, k) in enumerate(zip((8, 8, 2, 2), (16, 16, 4, 4))):
        put(f"ups.{i}", torch.randn(ch, ch // 2, k, generator=g) * (gain * math.sqrt(2.0 * u / (ch * k))), ch // 2)
        ch //= 2
        for j, rk in enumerate((3, 7, 11)):
            for m in range(3):
                for cs in ("convs1", "convs2"):
                    put(f"resblocks.{3 * i + j}.{cs}.{m}",
                        torch.randn(ch, ch, rk, generator=g) * (gain * 0.6 * math.sqrt(2.0 / (ch * rk))), ch)
    put("conv_post", torch.randn(1, 32, 7, generator=g) * (gain * 0.7 * math.sqrt(1.0 / (32 * 7))), 1)
    return sd


def prefix_mask(lengths, t_pad):
    """sequence_mask (matcha/utils/model.py:7-11) as float (B,1,T)."""
    lengths = torch.as_tensor(lengths)
    return (torch.arange(t_pad).unsqueeze(0) < lengths.unsqueeze(1)).float().unsqueeze(1)


def text_encoder_state_dict(n_spks=1, spk_emb_dim=64, seed=1234, n_vocab=178, n_channels=192, filter_channels=768,
                            filter_channels_dp=256, n_layers=6, kernel_size=3, n_feats=80, prefix=""):
    """`TextEncoder` parameters with the reference key layout (text_encoder.py:313-375); every constant-initialised
    tensor (prenet.proj zeros, LayerNorm gamma/beta, biases) is randomised.  Durations stay moderate: the duration
    head's output scale gives exp(logw) of a few frames per token."""
    g = _g(seed)
    sd = {}
    ch = n_channels + (spk_emb_dim if n_spks > 1 else 0)

    def conv(name, out_c, in_c, k, gain=1.0):
        sd[prefix + name + ".weight"] = torch.randn(out_c, in_c, k, generator=g) * (gain * math.sqrt(1.0 / (in_c * k)))
        sd[prefix + name + ".bias"] = torch.randn(out_c, generator=g) * 0.05

    def norm(name, c):
        sd[prefix + name + ".gamma"] = 1.0 + 0.2 * torch.randn(c, generator=g)
        sd[prefix + name + ".beta"] = 0.1 * torch.randn(c, generator=g)

    sd[prefix + "emb.weight"] = torch.randn(n_vocab, n_channels, generator=g) * n_channels ** -0.5
    for i in range(3):
        conv(f"prenet.conv_layers.{i}", n_channels, n_channels, 5)
        norm(f"prenet.norm_layers.{i}", n_channels)
    conv("prenet.proj", n_channels, n_channels, 1, gain=0.5)
    for i in range(n_layers):
        for p in ("conv_q", "conv_k", "conv_v", "conv_o"):
            conv(f"encoder.attn_layers.{i}.{p}", ch, ch, 1)
        norm(f"encoder.norm_layers_1.{i}", ch)
        conv(f"encoder.ffn_layers.{i}.conv_1", filter_channels, ch, kernel_size)
        conv(f"encoder.ffn_layers.{i}.conv_2", ch, filter_channels, kernel_size)
        norm(f"encoder.norm_layers_2.{i}", ch)
    conv("proj_m", n_feats, ch, 1)
    conv("proj_w.conv_1", filter_channels_dp, ch, kernel_size)
    norm("proj_w.norm_1", filter_channels_dp)
    conv("proj_w.conv_2", filter_channels_dp, filter_channels_dp, kernel_size)
    norm("proj_w.norm_2", filter_channels_dp)
    conv("proj_w.proj", 1, filter_channels_dp, 1, gain=0.6)
    sd[prefix + "proj_w.proj.bias"] = torch.full((1,), 0.6)
    return sd
