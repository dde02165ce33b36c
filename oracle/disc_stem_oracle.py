"""TEST INFRASTRUCTURE ONLY -- CPU restatement of the discriminator's first layer.

    models/gan.py:39   nn.Conv2d(input_nc, ndf, kernel_size=4, stride=2, padding=1), nn.LeakyReLU(0.2, True)

torch-CPU fp32 with the same two ATen ops the reference's modules call.  Pinned against the reference's own
NLayerDiscriminator (input_nc=2: the renderer's channel count; the shipped input_nc=3 cannot take the renderer's output,
SURVEY section 2) executed by oracle/make_golden.py --only-disc -> tests/golden/disc_stem.npz (forward and autograd
gradients).  Only tests/ and __graft_entry__.smoke() may import this module.
"""
import torch
import torch.nn.functional as F


def disc_stem(x: torch.Tensor, weight: torch.Tensor, bias: torch.Tensor) -> torch.Tensor:
    return F.leaky_relu(F.conv2d(x, weight, bias, stride=2, padding=1), 0.2)
