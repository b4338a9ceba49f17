"""TEST INFRASTRUCTURE -- CPU restatement of the pre-processing SplatLoc's render() does in PyTorch
before it calls the rasterizer, for the configuration it trains with (SH degree 0,
`convert_SHs_python`, covariance from scale / rotation):

    gaussian_splatting/scene/gaussian_model.py:75-110   getters: exp, normalize, sigmoid
    gaussian_splatting/gaussian_renderer/__init__.py:73-76, 85-102   isotropic repeat, SH -> colour, cat(kp_score)
    gaussian_splatting/utils/sh_utils.py:73   eval_sh at degree 0 = C0 * sh[..., 0]

Pinned by tests/golden/render_*.npz, which tests/golden/make_render_golden.py produced by running the
reference's own render() and GaussianModel getters on the CPU (tests/test_render_cpu.py: the activated
tensors bit for bit, the gradients of the raw parameters through autograd).  Only tests/, smoke() and
bench.py may import this module; the product (splatloc_b200/csrc: sl_forward_render) does the same
arithmetic inside its kernels.
"""
import torch

C0 = 0.28209479177387814


def preprocess(xyz, features_dc, kp_score, raw_opacity, raw_scaling, raw_rotation):
    """Raw optimiser parameters -> the rasterizer inputs render() builds:
    (means3D, colors_precomp (P, 4), opacities (P, 1), scales (P, 3), rotations (P, 4)).  Differentiable."""
    opacities = torch.sigmoid(raw_opacity)
    scales = torch.exp(raw_scaling)
    if scales.shape[-1] == 1:
        scales = scales.repeat(1, 3)
    rotations = torch.nn.functional.normalize(raw_rotation)
    shs_view = features_dc.transpose(1, 2).view(-1, 3, 1)           # (P, 1, 3) -> (P, 3, 1): channel-major
    colors = torch.clamp_min(C0 * shs_view[..., 0] + 0.5, 0.0)
    return xyz, torch.cat((colors, kp_score), dim=1), opacities, scales, rotations
