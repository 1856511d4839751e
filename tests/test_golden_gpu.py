"""The CUDA layer against the COMMITTED vectors in tests/golden/ (made by tools/make_golden.py
from the oracle; the GPU box has neither /root/reference nor Paddle). Tolerance 2e-2 relative
(north_star) per tensor."""
import os

import numpy as np
import pytest
import torch

from conftest import rel_err

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(__file__), "golden")


@pytest.mark.parametrize("name", ["layer_gpu_small_fp32", "layer_gpu_small_fp32_dropout"])
def test_layer_matches_committed_golden_vectors(name):
    from fleetx_b200.layers import TransformerDecoderLayer
    from fleetx_b200.rng import set_seed

    fx = np.load(os.path.join(GOLD, name + ".npz"))
    b, s, h, nh, f, n, p1000, seed = [int(v) for v in fx["cfg"]]
    assert n == 1
    set_seed(seed)
    layer = TransformerDecoderLayer(h, nh, f, dropout=p1000 / 1000.0)
    layer.load_full_state({k[6:]: torch.from_numpy(fx[k]) for k in fx.files if k.startswith("param.")})
    layer.train()
    x = torch.from_numpy(fx["x"]).to(torch.bfloat16).cuda().requires_grad_(True)
    y = layer(x)
    y.backward(torch.from_numpy(fx["dy"]).to(torch.bfloat16).cuda())
    torch.cuda.synchronize()
    assert rel_err(y, torch.from_numpy(fx["y"])) < 2e-2
    assert rel_err(x.grad, torch.from_numpy(fx["dx"])) < 2e-2
    for pname, p in layer.named_parameters():
        assert rel_err(p.grad, torch.from_numpy(fx["grad." + pname])) < 2e-2, pname
