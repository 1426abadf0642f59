"""Decoder parameter containers with the reference's module tree, so ``state_dict()`` keys
(``sdf_net.model.0.weight`` ...) round-trip with checkpoints written by the reference
(``mneslam_mp.py:294-315``).  Mirrors reference ``model/decoder.py:7-175`` (``ColorNet``, ``SDFNet``,
``ColorSDFNet``, ``ColorSDFNet_v2``) for ``decoder.tcnn_network: False``, the only mode its configs use.

The arithmetic is NOT done here: the fused CUDA kernels read these weights through raw pointers.  No
reference call site invokes a sub-network on its own (they go through ``JointEncoding.query_*`` /
``run_network*``), so the modules' ``forward`` raises instead of silently running a torch path.
"""
import torch.nn as nn


def _mlp(in_dim, hidden, out_dim, num_layers):
    layers = []
    for l in range(num_layers):
        i = in_dim if l == 0 else hidden
        o = out_dim if l == num_layers - 1 else hidden
        layers.append(nn.Linear(i, o, bias=False))
        if l != num_layers - 1:
            layers.append(nn.ReLU(inplace=True))
    return nn.Sequential(*layers)


class _Net(nn.Module):
    def forward(self, *a, **k):
        raise RuntimeError('mneslam_b200: the decoder runs inside the fused CUDA kernels; call '
                           'JointEncoding.query_color_sdf / run_network_flat instead of the sub-network')


class ColorNet(_Net):
    def __init__(self, config, input_ch=4, geo_feat_dim=15, hidden_dim_color=64, num_layers_color=3):
        super().__init__()
        if config['decoder'].get('tcnn_network', False):
            raise NotImplementedError('decoder.tcnn_network: True is dormant in the reference and not built')
        self.model = _mlp(input_ch + geo_feat_dim, hidden_dim_color, 3, num_layers_color)


class SDFNet(_Net):
    def __init__(self, config, input_ch=3, geo_feat_dim=15, hidden_dim=64, num_layers=2):
        super().__init__()
        if config['decoder'].get('tcnn_network', False):
            raise NotImplementedError('decoder.tcnn_network: True is dormant in the reference and not built')
        self.model = _mlp(input_ch, hidden_dim, 1 + geo_feat_dim, num_layers)


class _ColorSDF(_Net):
    def __init__(self, config, input_ch, input_ch_pos, colour_grid):
        super().__init__()
        d = config['decoder']
        if (d['hidden_dim'], d['hidden_dim_color'], d['geo_feat_dim'], d['num_layers'], d['num_layers_color']) != \
                (32, 32, 15, 2, 2):
            raise NotImplementedError('the fused kernels are built for the decoder every reference config uses: '
                                      'hidden 32/32, geo_feat_dim 15, 2+2 layers; got %r' % (d,))
        col_in = input_ch_pos + (input_ch if colour_grid else 0)
        self.color_net = ColorNet(config, input_ch=col_in, geo_feat_dim=d['geo_feat_dim'],
                                  hidden_dim_color=d['hidden_dim_color'], num_layers_color=d['num_layers_color'])
        self.sdf_net = SDFNet(config, input_ch=input_ch + input_ch_pos, geo_feat_dim=d['geo_feat_dim'],
                              hidden_dim=d['hidden_dim'], num_layers=d['num_layers'])

    def weight_list(self):
        """[sdf0, sdf1, col0, col1] in the order the C ABI takes them."""
        return [self.sdf_net.model[0].weight, self.sdf_net.model[2].weight,
                self.color_net.model[0].weight, self.color_net.model[2].weight]


class ColorSDFNet(_ColorSDF):
    """Colour grid + SDF grid (``grid.oneGrid: False``)."""
    def __init__(self, config, input_ch=3, input_ch_pos=12):
        super().__init__(config, input_ch, input_ch_pos, colour_grid=True)


class ColorSDFNet_v2(_ColorSDF):
    """No colour grid (``grid.oneGrid: True``)."""
    def __init__(self, config, input_ch=3, input_ch_pos=12):
        super().__init__(config, input_ch, input_ch_pos, colour_grid=False)
