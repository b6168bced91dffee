"""Decoder modules with the reference's parameter names and shapes (model/decoder.py:7-180, nn.Linear branch) so that
state_dict keys (`decoder.sdf_net.model.0.weight`, ...), optimiser parameter groups (coslam.py:615-622) and checkpoints
(coslam.py:112-129) stay interchangeable.  The modules are containers for the weights: on the hot path the four
matrices are handed to the fused kernels by JointEncoding; calling a module directly runs the same kernels' math through
plain torch ops on whatever device the inputs live on (used for shape checks and ONNX-style export only).
"""
import torch
import torch.nn as nn

from . import tcnn


def _stack(dims):
    layers = []
    for i in range(len(dims) - 1):
        layers.append(nn.Linear(dims[i], dims[i + 1], bias=False))
        if i != len(dims) - 2:
            layers.append(nn.ReLU(inplace=True))
    return nn.Sequential(*layers)


class SDFNet(nn.Module):
    """model/decoder.py:57-108."""

    def __init__(self, config, input_ch=3, geo_feat_dim=15, hidden_dim=64, num_layers=2):
        super().__init__()
        self.config, self.input_ch, self.geo_feat_dim = config, input_ch, geo_feat_dim
        self.hidden_dim, self.num_layers = hidden_dim, num_layers
        if config['decoder']['tcnn_network']:
            self.model = tcnn.Network(input_ch, 1 + geo_feat_dim, {
                'otype': 'FullyFusedMLP', 'activation': 'ReLU', 'output_activation': 'None',
                'n_neurons': hidden_dim, 'n_hidden_layers': num_layers - 1})
        else:
            self.model = _stack([input_ch] + [hidden_dim] * (num_layers - 1) + [1 + geo_feat_dim])

    def forward(self, x, return_geo=True):
        out = self.model(x)
        return out if return_geo else out[..., :1]


class ColorNet(nn.Module):
    """model/decoder.py:7-55."""

    def __init__(self, config, input_ch=4, geo_feat_dim=15, hidden_dim_color=64, num_layers_color=3):
        super().__init__()
        self.config, self.input_ch, self.geo_feat_dim = config, input_ch, geo_feat_dim
        self.hidden_dim_color, self.num_layers_color = hidden_dim_color, num_layers_color
        if config['decoder']['tcnn_network']:
            self.model = tcnn.Network(input_ch + geo_feat_dim, 3, {
                'otype': 'FullyFusedMLP', 'activation': 'ReLU', 'output_activation': 'None',
                'n_neurons': hidden_dim_color, 'n_hidden_layers': num_layers_color - 1})
        else:
            self.model = _stack([input_ch + geo_feat_dim] + [hidden_dim_color] * (num_layers_color - 1) + [3])

    def forward(self, input_feat):
        return self.model(input_feat)


class ColorSDFNet_v2(nn.Module):
    """One-grid decoder (model/decoder.py:145-180): sdf_net on [embed | embed_pos], color_net on [embed_pos | geo]."""

    def __init__(self, config, input_ch=3, input_ch_pos=12):
        super().__init__()
        d = config['decoder']
        self.config = config
        self.color_net = ColorNet(config, input_ch=input_ch_pos, geo_feat_dim=d['geo_feat_dim'],
                                  hidden_dim_color=d['hidden_dim_color'], num_layers_color=d['num_layers_color'])
        self.sdf_net = SDFNet(config, input_ch=input_ch + input_ch_pos, geo_feat_dim=d['geo_feat_dim'],
                              hidden_dim=d['hidden_dim'], num_layers=d['num_layers'])

    def forward(self, embed, embed_pos):
        h = self.sdf_net(torch.cat([embed, embed_pos], dim=-1), return_geo=True)
        sdf, geo_feat = h[..., :1], h[..., 1:]
        rgb = self.color_net(torch.cat([embed_pos, geo_feat], dim=-1))
        return torch.cat([rgb, sdf], -1)

    def fused_weights(self):
        """The four matrices in the order the kernels take them, or None if the layer stack is not the 2+2 one."""
        try:
            s, c = self.sdf_net.model, self.color_net.model
            if len(s) != 3 or len(c) != 3:
                return None
            return s[0].weight, s[2].weight, c[0].weight, c[2].weight
        except (TypeError, AttributeError):
            return None


class ColorSDFNet(nn.Module):
    """Two-grid variant (model/decoder.py:110-143, `oneGrid: False`): colour net additionally sees a colour grid
    feature.  Config-reachable but unused by the shipped configs; served through the stand-alone encodings."""

    def __init__(self, config, input_ch=3, input_ch_pos=12):
        super().__init__()
        d = config['decoder']
        self.config = config
        self.color_net = ColorNet(config, input_ch=input_ch + input_ch_pos, geo_feat_dim=d['geo_feat_dim'],
                                  hidden_dim_color=d['hidden_dim_color'], num_layers_color=d['num_layers_color'])
        self.sdf_net = SDFNet(config, input_ch=input_ch + input_ch_pos, geo_feat_dim=d['geo_feat_dim'],
                              hidden_dim=d['hidden_dim'], num_layers=d['num_layers'])

    def forward(self, embed, embed_pos, embed_color):
        """[sdf-grid feature (| OneBlob)] -> SDF head; [(OneBlob |) colour-grid feature | geometric feature] -> colour head."""
        with_pos = embed_pos is not None
        sdf_in = (embed, embed_pos) if with_pos else (embed,)
        out = self.sdf_net(torch.cat(sdf_in, dim=-1), return_geo=True)
        col_in = ((embed_pos,) if with_pos else ()) + (embed_color, out[..., 1:])
        return torch.cat([self.color_net(torch.cat(col_in, dim=-1)), out[..., :1]], -1)
