"""CPU checks of the Q-network restatement helpers (the network is PyTorch and not the hot path)."""
import pytest
import torch

from toric_rl_decoder_b200 import networks


@pytest.mark.parametrize("arch,d", (("NN_11", 5), ("NN_11", 7), ("NN_8", 5), ("NN_17", 9)))
def test_pad_channels_is_the_same_function(arch, d):
    """Zero-padded channels change nothing: same Q-values as the reference channel schedule."""
    torch.manual_seed(1)
    m = getattr(networks, arch)(d, 3).eval()
    p = networks.pad_channels(m, 8)
    for k in range(1, p.depth):
        assert getattr(p, "conv%d" % k).out_channels % 8 == 0
    x = (torch.rand(8, 2, d, d) < 0.25).float()
    with torch.no_grad():
        a, b = m(x), p(x)
    assert torch.allclose(a, b, rtol=1e-6, atol=1e-7)
