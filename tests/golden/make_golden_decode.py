"""Generates tests/golden/decode_reference.npz: the reference's own decode prologue on a seeded compressed representation.
quantize.py is imported from /root/reference; its third-party imports that are absent here and NOT on this path
(vector_quantize_pytorch, constriction, pytorch_msssim) are stubbed so that the module loads; UniformQuantizer.decompress
and VectorQuantizer.decompress are then the reference's unmodified code.  Run:  python tests/golden/make_golden_decode.py
"""
import sys
import types

import numpy as np
import torch

for name, attrs in (("vector_quantize_pytorch", ("VectorQuantize", "ResidualVQ")), ("constriction", ()), ("pytorch_msssim", ("ms_ssim", "ssim"))):
    m = types.ModuleType(name)
    for a in attrs:
        setattr(m, a, type(a, (), {}))
    sys.modules[name] = m
sys.path.insert(0, "/root/reference")
from quantize import UniformQuantizer, VectorQuantizer  # noqa: E402


class _Layer:
    def __init__(self, embed):
        self._codebook = types.SimpleNamespace(embed=embed)      # layer._codebook.embed[0, index] (quantize.py:149)


def main(out="tests/golden/decode_reference.npz"):
    g = torch.Generator().manual_seed(5)
    N, K, C = 777, 2, 8
    xyz = torch.atanh((torch.rand(N, 2, generator=g) * 2 - 1).clamp(-0.9995, 0.9995)).half()
    uq = UniformQuantizer(signed=False, bits=6, learned=True, num_channels=3)
    uq.scale.data = torch.rand(3, generator=g) * 0.02 + 0.005
    uq.beta.data = torch.rand(3, generator=g) * 0.1 - 0.05
    codes = torch.randint(0, 64, (N, 3), generator=g).float()
    bound = torch.tensor([0.5, 0.0, 0.5]).view(1, 3)
    vq = VectorQuantizer.__new__(VectorQuantizer)
    torch.nn.Module.__init__(vq)
    embeds = [torch.rand(1, C, 3, generator=g) for _ in range(K)]
    vq.quantizer = types.SimpleNamespace(layers=[_Layer(e) for e in embeds])
    index = torch.randint(0, C, (N, K), generator=g)
    with torch.no_grad():
        means = torch.tanh(xyz.float())                                   # gaussianimage_cholesky.py:163
        chol = uq.decompress(codes) + bound                               # :164-165
        colors = vq.decompress(index)                                     # :166
    np.savez_compressed(out, xyz=xyz.numpy(), codes=codes.numpy(), scale=uq.scale.data.numpy(), beta=uq.beta.data.numpy(),
                        bound=bound.numpy(), index=index.numpy(), codebooks=torch.cat(embeds, 0).numpy(), means=means.numpy(),
                        cholesky=chol.numpy(), colors=colors.numpy())
    print("wrote", out)


if __name__ == "__main__":
    main(*sys.argv[1:])
