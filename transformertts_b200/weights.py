"""state_dict (reference key names) -> packed device buffers + the C `TtsWeights` struct.

Host-side preparation done once per checkpoint:
  * eval-mode BatchNorm folded into the preceding Conv1d (model/module.py:18-19,28-33):
        W' = W * g / sqrt(var + eps),  b' = (b - mean) * g / sqrt(var + eps) + beta
  * conv weights re-laid out [Cout][tap][Cin_pad] (K-major rows for the implicit GEMM), Cin padded to 64
  * 1/sqrt(head_dim) = 1/8 folded into the q rows of every in_proj (exact in bf16: power of two)
    (torch/nn/functional.py:6632 scales q before q k^T)
  * mel (80x256) and stop (1x256) heads stacked into one [81][256] matrix
  * matrices cast to bf16, biases / LayerNorm / PE table kept fp32
  * K-blocked transposes [K/8][256][8] of the decoder out-projections, FFN second matrices and prenet second matrix
    for the AR decode kernel's K-split contractions
"""
from __future__ import annotations

import ctypes as C
from typing import Dict

import torch

from . import _ext

D, H, DH, FFN, N_MELS, KS = 256, 4, 64, 1024, 80, 5
BN_EPS = 1e-5
MAX_PE = 5000

EXPECTED_CONFIG = dict(
    encoder_prenet_n_layers=3, encoder_prenet_in_channel=256, encoder_prenet_out_channel=256,
    encoder_prenet_kernel_size=5, encoder_n_layers=3, encoder_n_head=4, encoder_d_ffn=1024,
    decoder_n_layers=3, decoder_n_head=4, decoder_d_ffn=1024, postnet_n_layers=5, postnet_kernel_size=5,
    d_model=256, n_mels=80,
)


def fold_conv_bn(sd: Dict[str, torch.Tensor], prefix: str, cin_pad: int):
    """Returns (W [Cout][KS*cin_pad] fp32, b [Cout] fp32) with BN folded (eval mode)."""
    w = sd[f"{prefix}.conv.weight"].double()          # [Cout][Cin][KS]
    b = sd[f"{prefix}.conv.bias"].double()
    g, beta = sd[f"{prefix}.bn.weight"].double(), sd[f"{prefix}.bn.bias"].double()
    mean, var = sd[f"{prefix}.bn.running_mean"].double(), sd[f"{prefix}.bn.running_var"].double()
    scale = g / torch.sqrt(var + BN_EPS)
    w = w * scale[:, None, None]
    b = (b - mean) * scale + beta
    cout, cin, ks = w.shape
    wp = torch.zeros(cout, ks, cin_pad, dtype=torch.float64)
    wp[:, :, :cin] = w.permute(0, 2, 1)
    return wp.reshape(cout, ks * cin_pad).float(), b.float()


class PackedWeights:
    """Owns the device copies; `.struct` is the ctypes TtsWeights to pass by reference."""

    def __init__(self, sd: Dict[str, torch.Tensor], device):
        self.device = torch.device(device)
        sd = {k: v.detach().to("cpu") for k, v in sd.items()}
        self._check_shapes(sd)
        self._keep = []
        s = _ext.Weights()

        def mat(t):      # bf16 matrix on device
            d = t.to(torch.bfloat16).contiguous().to(self.device)
            self._keep.append(d)
            return C.c_void_p(d.data_ptr())

        def vec(t):      # fp32 vector on device
            d = t.to(torch.float32).contiguous().to(self.device)
            self._keep.append(d)
            return C.c_void_p(d.data_ptr())

        def attn(dst, prefix):
            w_in = sd[f"{prefix}.in_proj_weight"].clone().float()
            b_in = sd[f"{prefix}.in_proj_bias"].clone().float()
            w_in[:D] *= 1.0 / 8.0
            b_in[:D] *= 1.0 / 8.0
            dst.w_in, dst.b_in = mat(w_in), vec(b_in)
            dst.w_out, dst.b_out = mat(sd[f"{prefix}.out_proj.weight"]), vec(sd[f"{prefix}.out_proj.bias"])

        def ffn(dst, prefix):
            dst.w1, dst.b1 = mat(sd[f"{prefix}.linear1.weight"]), vec(sd[f"{prefix}.linear1.bias"])
            dst.w2, dst.b2 = mat(sd[f"{prefix}.linear2.weight"]), vec(sd[f"{prefix}.linear2.bias"])

        def ln(dst, prefix):
            dst.g, dst.b = vec(sd[f"{prefix}.weight"]), vec(sd[f"{prefix}.bias"])

        s.emb = mat(sd["emb.weight"])
        s.n_phon = sd["emb.weight"].shape[0]
        for i in range(3):
            w, b = fold_conv_bn(sd, f"enc_prenet.layers.{2 * i}", D)
            s.enc_conv_w[i], s.enc_conv_b[i] = mat(w), vec(b)
        s.enc_lin_w, s.enc_lin_b = mat(sd["enc_prenet.linear.linear.weight"]), vec(sd["enc_prenet.linear.linear.bias"])
        for l in range(3):
            p = f"encoder.layers.{l}"
            attn(s.enc[l].sa, f"{p}.self_attn")
            ffn(s.enc[l].ffn, p)
            ln(s.enc[l].ln1, f"{p}.norm1")
            ln(s.enc[l].ln2, f"{p}.norm2")
        w1 = torch.zeros(D, 128)
        w1[:, :N_MELS] = sd["dec_prenet.linear1.linear.weight"].float()
        s.pre1_w, s.pre1_b = mat(w1), vec(sd["dec_prenet.linear1.linear.bias"])
        s.pre2_w, s.pre2_b = mat(sd["dec_prenet.linear2.linear.weight"]), vec(sd["dec_prenet.linear2.linear.bias"])
        for l in range(3):
            p = f"decoder.layers.{l}"
            attn(s.dec[l].sa, f"{p}.self_attn")
            attn(s.dec[l].ca, f"{p}.multihead_attn")
            ffn(s.dec[l].ffn, p)
            ln(s.dec[l].ln1, f"{p}.norm1")
            ln(s.dec[l].ln2, f"{p}.norm2")
            ln(s.dec[l].ln3, f"{p}.norm3")
        s.head_w = mat(torch.cat([sd["linear1.linear.weight"].float(), sd["linear2.linear.weight"].float()], dim=0))
        s.head_b = vec(torch.cat([sd["linear1.linear.bias"].float(), sd["linear2.linear.bias"].float()], dim=0))
        for i in range(5):
            cin = N_MELS if i == 0 else D
            w, b = fold_conv_bn(sd, f"postnet.layers.{3 * i}", 128 if i == 0 else D)
            assert w.shape[1] == KS * (128 if cin == N_MELS else D)
            s.post_w[i], s.post_b[i] = mat(w), vec(b)
        # cross-attention K/V projections of all decoder layers stacked (rows 256-767 of each in_proj): one GEMM
        s.ckv_w = mat(torch.cat([sd[f"decoder.layers.{l}.multihead_attn.in_proj_weight"][D:].float() for l in range(3)]))
        s.ckv_b = vec(torch.cat([sd[f"decoder.layers.{l}.multihead_attn.in_proj_bias"][D:].float() for l in range(3)]))
        # K-blocked transposes [K/8][256][8] of the matrices the AR decode kernel contracts K-split (decode.cu)
        def kt(w):
            n, k = w.shape
            return mat(w.float().view(n, k // 8, 8).permute(1, 0, 2))
        for l in range(3):
            p = f"decoder.layers.{l}"
            s.dec_kt[l][0] = kt(sd[f"{p}.self_attn.out_proj.weight"])
            s.dec_kt[l][1] = kt(sd[f"{p}.multihead_attn.out_proj.weight"])
            s.dec_kt[l][2] = kt(sd[f"{p}.linear2.weight"])
        s.pre2_kt = kt(sd["dec_prenet.linear2.linear.weight"])
        s.pe = vec(sd["pe.pe"])
        s.alpha = float(sd["pe.alpha"].reshape(-1)[0])
        self.struct = s

    @staticmethod
    def _check_shapes(sd):
        need = {
            "emb.weight": (None, D), "enc_prenet.layers.0.conv.weight": (D, D, KS),
            "enc_prenet.linear.linear.weight": (D, D), "dec_prenet.linear1.linear.weight": (D, N_MELS),
            "dec_prenet.linear2.linear.weight": (D, D), "encoder.layers.2.linear1.weight": (FFN, D),
            "decoder.layers.2.multihead_attn.in_proj_weight": (3 * D, D),
            "postnet.layers.0.conv.weight": (D, N_MELS, KS), "postnet.layers.12.conv.weight": (N_MELS, D, KS),
            "linear1.linear.weight": (N_MELS, D), "linear2.linear.weight": (1, D), "pe.pe": (MAX_PE, D),
        }
        for k, shp in need.items():
            if k not in sd:
                raise ValueError(f"state_dict is missing '{k}' (expected the reference TransformerTTS layout)")
            got = tuple(sd[k].shape)
            if len(got) != len(shp) or any(e is not None and e != g for e, g in zip(shp, got)):
                raise ValueError(f"'{k}' has shape {got}; the sm_100a kernels are specialised to {shp}")
        for stack, n in (("encoder", 3), ("decoder", 3)):
            if f"{stack}.layers.{n}.linear1.weight" in sd:
                raise ValueError(f"more than {n} {stack} layers: unsupported by the specialised kernels")
