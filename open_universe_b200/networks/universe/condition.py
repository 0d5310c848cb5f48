"""State (parameters only) of the reference ConditionerNetwork
(networks/universe/condition.py:33-344).  The forward pass lives in the native kernels."""
import math

from torch import nn

from .layers import ConvBlockP, MelAdapterP, PReLUConv, WConv


class ConditionerEncoder(nn.Module):
    def __init__(self, ds_factors, input_channels, with_extra_conv_block, use_weight_norm):
        super().__init__()
        c = input_channels
        # the reference forces use_antialiasing=False in the conditioner encoder (condition.py:333)
        self.ds_modules = nn.ModuleList([ConvBlockP(c * 2**i, r, "down", weight_norm=use_weight_norm) for i, r in enumerate(ds_factors)])
        rates = [ds_factors[-1]]
        for r in ds_factors[-2::-1]:
            rates.append(rates[-1] * r)
        rates = rates[::-1]
        oc = c * 2 ** len(ds_factors)
        self.st_convs = nn.ModuleList()
        for i in range(len(ds_factors)):
            if i >= len(ds_factors) - 1:
                self.st_convs.append(None)
            else:
                self.st_convs.append(PReLUConv(c * 2**i, oc, rates[i], weight_norm=use_weight_norm))
        if with_extra_conv_block:
            self.ds_modules.append(ConvBlockP(oc, weight_norm=use_weight_norm))
            self.st_convs.append(None)
        self.gru = nn.GRU(oc, oc // 2, num_layers=2, bidirectional=True, batch_first=True)
        self.conv_block1 = ConvBlockP(oc, weight_norm=use_weight_norm)
        self.conv_block2 = ConvBlockP(oc, weight_norm=use_weight_norm)


class ConditionerDecoder(nn.Module):
    def __init__(self, up_factors, input_channels, with_extra_conv_block, use_weight_norm, use_antialiasing):
        super().__init__()
        n_channels = [input_channels * 2 ** (len(up_factors) - i - 1) for i in range(len(up_factors))]
        self.input_conv_block = ConvBlockP(n_channels[0] * 2, weight_norm=use_weight_norm)
        ups = [ConvBlockP(c, r, "up", weight_norm=use_weight_norm, antialias=use_antialiasing) for c, r in zip(n_channels, up_factors)]
        if with_extra_conv_block:
            ups = [ConvBlockP(2 * n_channels[0], weight_norm=use_weight_norm)] + ups
        self.up_modules = nn.ModuleList(ups)


class ConditionerNetwork(nn.Module):
    def __init__(self, fb_kernel_size=3, rate_factors=(2, 4, 4, 5), n_channels=32, n_mels=80, n_mel_oversample=4,
                 encoder_gru_residual=False, extra_conv_block=False, encoder_act_type="prelu",
                 decoder_act_type="prelu", precoding=None, input_channels=1, output_channels=None,
                 use_weight_norm=False, seq_model="gru", use_antialiasing=False):
        super().__init__()
        if precoding or seq_model != "gru" or input_channels != 1 or output_channels is not None:
            raise NotImplementedError("ConditionerNetwork option outside the supported UNIVERSE / UNIVERSE++ configurations")
        if encoder_act_type != "prelu" or decoder_act_type != "prelu":
            raise NotImplementedError("only PReLU activations are supported in the conditioner")
        rate_factors = [int(r) for r in rate_factors]
        self.input_conv = WConv((n_channels, input_channels, fb_kernel_size), weight_norm=use_weight_norm)
        total_ds = math.prod(rate_factors)
        total_channels = 2 ** len(rate_factors) * n_channels
        self.input_mel = MelAdapterP(n_mels, total_channels, total_ds * input_channels, n_mel_oversample, weight_norm=use_weight_norm)
        self.encoder = ConditionerEncoder(rate_factors, n_channels, extra_conv_block, use_weight_norm)
        self.decoder = ConditionerDecoder(rate_factors[::-1], n_channels, extra_conv_block, use_weight_norm, use_antialiasing)

    def forward(self, *args, **kwargs):
        raise RuntimeError("ConditionerNetwork has no PyTorch forward in open_universe_b200: use Universe.enhance (native sm_100a path)")
