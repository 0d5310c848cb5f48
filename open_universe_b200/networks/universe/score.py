"""State (parameters only) of the reference ScoreNetwork (networks/universe/score.py:27-276).
The forward pass lives in the native kernels; see open_universe_b200/program.py."""
import torch
from torch import nn

from .layers import ConvBlockP, PReLUConv, PReLUw, SigmaBlockP, SimpleTimeEmbeddingP, WConv


class ScoreEncoder(nn.Module):
    def __init__(self, ds_factors, input_channels, noise_cond_dim, with_extra_conv_block, use_weight_norm, use_antialiasing):
        super().__init__()
        c = input_channels
        self.ds_modules = nn.ModuleList(
            [ConvBlockP(c * 2**i, r, "down", weight_norm=use_weight_norm, antialias=use_antialiasing) for i, r in enumerate(ds_factors)]
        )
        self.cond_proj = nn.ModuleList(
            [WConv((c * 2 ** (i + 1), noise_cond_dim), weight_norm=use_weight_norm) for i in range(len(ds_factors))]
        )
        oc = c * 2 ** len(ds_factors)
        if with_extra_conv_block:
            self.ds_modules.append(ConvBlockP(oc, weight_norm=use_weight_norm))
            self.cond_proj.append(WConv((2 * oc, noise_cond_dim), weight_norm=use_weight_norm))
        self.gru = nn.GRU(oc, oc // 2, num_layers=1, bidirectional=True, batch_first=True)


class ScoreDecoder(nn.Module):
    def __init__(self, up_factors, input_channels, noise_cond_dim, with_extra_conv_block, use_weight_norm, use_antialiasing):
        super().__init__()
        n_channels = [input_channels * 2 ** (len(up_factors) - i - 1) for i in range(len(up_factors))]
        self.up_modules = nn.ModuleList()
        self.noise_cond_proj = nn.ModuleList()
        self.signal_cond_proj = nn.ModuleList()
        if with_extra_conv_block:
            oc = input_channels * 2 ** len(up_factors)
            self.up_modules.append(ConvBlockP(oc, weight_norm=use_weight_norm))
            self.noise_cond_proj.append(WConv((2 * oc, noise_cond_dim), weight_norm=use_weight_norm))
            self.signal_cond_proj.append(WConv((oc, oc, 1), weight_norm=use_weight_norm))
        for c, r in zip(n_channels, up_factors):
            self.up_modules.append(ConvBlockP(c, r, "up", weight_norm=use_weight_norm, antialias=use_antialiasing))
            self.noise_cond_proj.append(WConv((2 * c, noise_cond_dim), weight_norm=use_weight_norm))
            self.signal_cond_proj.append(WConv((c, c, 1), weight_norm=use_weight_norm))


class ScoreNetwork(nn.Module):
    def __init__(self, fb_kernel_size=3, rate_factors=(2, 4, 4, 5), n_channels=32, n_rff=32, noise_cond_dim=512,
                 encoder_gru_conv_sandwich=False, extra_conv_block=False, encoder_act_type="prelu",
                 decoder_act_type="prelu", precoding=None, input_channels=1, output_channels=1,
                 use_weight_norm=False, seq_model="gru", use_antialiasing=False, time_embedding=None):
        super().__init__()
        if encoder_gru_conv_sandwich or precoding or seq_model != "gru" or input_channels != 1 or output_channels != 1:
            raise NotImplementedError("ScoreNetwork option outside the supported UNIVERSE / UNIVERSE++ configurations")
        if encoder_act_type != "prelu" or decoder_act_type != "prelu":
            raise NotImplementedError("only PReLU activations are supported in the score network")
        rate_factors = [int(r) for r in rate_factors]
        if time_embedding == "simple":
            self.sigma_block = SimpleTimeEmbeddingP(n_dim=noise_cond_dim)
        else:
            self.sigma_block = SigmaBlockP(n_rff, noise_cond_dim)
        self.input_conv = WConv((n_channels, input_channels, fb_kernel_size))
        self.encoder = ScoreEncoder(rate_factors, n_channels, noise_cond_dim, extra_conv_block, use_weight_norm, use_antialiasing)
        self.decoder = ScoreDecoder(rate_factors[::-1], n_channels, noise_cond_dim, extra_conv_block, use_weight_norm, use_antialiasing)
        self.prelu = PReLUw()
        self.output_conv = PReLUConv(n_channels, output_channels, fb_kernel_size, weight_norm=use_weight_norm)

    def forward(self, *args, **kwargs):
        raise RuntimeError("ScoreNetwork has no PyTorch forward in open_universe_b200: use Universe.enhance (native sm_100a path)")
