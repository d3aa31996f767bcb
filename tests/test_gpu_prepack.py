"""One-launch weight packing for a whole model (tc_conv.prepacked) against the per-layer pack kernel."""
import pytest
import torch

pytestmark = pytest.mark.gpu

import pytorch_distributed_learning_b200 as pdl  # noqa: E402
from pytorch_distributed_learning_b200.ops import tc_conv  # noqa: E402


@pytest.fixture(autouse=True)
def _tf32_convs():
    old = torch.backends.cudnn.allow_tf32
    torch.backends.cudnn.allow_tf32 = True
    yield
    torch.backends.cudnn.allow_tf32 = old


def test_prepacked_matches_per_layer_pack():
    torch.manual_seed(0)
    nn = pdl.nn
    model = nn.Sequential(nn.Conv2d(3, 64, 7, 2, 3, bias=False),      # stem: not eligible (Cin = 3), stays per-layer
                          nn.Conv2d(64, 64, 3, 1, 1, bias=False), nn.Conv2d(64, 256, 1, bias=False), nn.Conv2d(256, 128, 3, 2, 1, bias=False),
                          nn.Conv2d(128, 32, 1, bias=True), nn.Conv2d(32, 96, 3, bias=False)).cuda()  # 96 outputs: not a supported column tile
    eligible = [m.weight for m in list(model)[1:5]]
    singles = [tc_conv.pack_weights(w, True, True) for w in eligible]
    with tc_conv.prepacked(model):
        for w, (f0, b0) in zip(eligible, singles):
            f1, b1 = tc_conv.pack_weights(w, True, True)
            assert f1.data_ptr() != f0.data_ptr()  # came from the model-wide buffers, not a fresh launch
            assert torch.equal(f1, f0) and torch.equal(b1, b0)
            f2, b2 = tc_conv.pack_weights(w, True, False)
            assert b2 is None and f2.data_ptr() == f1.data_ptr()
        # ineligible layers silently fall back to the per-layer kernel
        f, _ = tc_conv.pack_weights(model[5].weight, True, False)
        assert f.shape == (9, 96, 32)
    # outside the block nothing is cached
    f3, _ = tc_conv.pack_weights(eligible[0], True, False)
    assert torch.equal(f3, singles[0][0])
    # weights change -> the next block repacks
    with torch.no_grad():
        eligible[0].mul_(2.0)
    with tc_conv.prepacked(model):
        f4, _ = tc_conv.pack_weights(eligible[0], True, False)
    torch.testing.assert_close(f4, singles[0][0] * 2.0, rtol=1e-3, atol=0)
