"""Q-networks the env path feeds (PyTorch; NOT the hot path -- north_star keeps the forward in PyTorch).

Table-driven restatement of the reference's plain CNN decoders (src/nn/torch/NN.py:10-44 NN_11,
:47-77 NN_8, :80-143 NN_17): circular padding by one site, a valid 3x3 convolution, `same` 3x3
convolutions with the channel schedule below, a last valid 3x3 convolution and one linear layer to the 3
Pauli Q-values.  Parameter names (conv1..convN, linear1) match the reference so its state_dicts load
unchanged (network/converged/*.pt; the legacy full-module pickle network/Size_5_NN_11.pt has to be
converted to a state_dict by the reference's own loader first; tests/golden/ holds the converted files).
"""
import torch
import torch.nn as nn
import torch.nn.functional as F

# output channels of conv1..convN; the first and the last convolution are unpadded ("valid")
CHANNELS = {
    "NN_11": (128, 128, 120, 111, 104, 103, 90, 80, 73, 71, 64),
    "NN_8": (256, 256, 240, 224, 220, 215, 205, 200),
    "NN_17": (256, 256, 251, 250, 240, 240, 235, 233, 233, 229, 225, 223, 220, 220, 220, 215, 214, 205, 204, 200),
}


class ConvDecoder(nn.Module):
    """Input (B,2,d,d) float32 perspectives -> (B,3) Q-values for X, Y, Z on the centred qubit."""

    # rows are independent, so policy._forward may pad a small batch with zero rows (fewer distinct cuDNN shapes)
    pad_small_batches = True

    def __init__(self, arch, system_size, number_of_actions=3, device=None):
        super().__init__()
        chans = CHANNELS[arch]
        self.arch, self.system_size, self.depth = arch, int(system_size), len(chans)
        cin = 2
        for k, cout in enumerate(chans, start=1):
            valid = k == 1 or k == len(chans)
            setattr(self, "conv%d" % k, nn.Conv2d(cin, cout, kernel_size=3, stride=1, padding=0 if valid else 1))
            cin = cout
        side = self.system_size + 2 - 2 - 2          # circular pad +2, two valid 3x3 convolutions -2 each
        self.linear1 = nn.Linear(chans[-1] * side * side, 3)
        self.device = device
        if device is not None:
            self.to(device)

    def forward(self, x):
        x = F.pad(x, (1, 1, 1, 1), mode="circular")  # pad_circular(x, 1), src/nn/torch/util.py:22-27
        for k in range(1, self.depth + 1):
            x = F.relu(getattr(self, "conv%d" % k)(x))
        return self.linear1(x.flatten(1))


def NN_11(system_size, number_of_actions=3, device=None):
    return ConvDecoder("NN_11", system_size, number_of_actions, device)


def NN_8(system_size, number_of_actions=3, device=None):
    return ConvDecoder("NN_8", system_size, number_of_actions, device)


def NN_17(system_size, number_of_actions=3, device=None):
    return ConvDecoder("NN_17", system_size, number_of_actions, device)


def load_state(model, path_or_state, device=None):
    """Load a reference state_dict (weights_only) into a ConvDecoder."""
    sd = path_or_state
    if isinstance(sd, (str, bytes)):
        sd = torch.load(sd, map_location="cpu", weights_only=True)
    model.load_state_dict(sd)
    if device is not None:
        model.to(device)
    return model.eval()


def pad_channels(model, multiple=8):
    """A ConvDecoder with every hidden channel count rounded up to `multiple` and the extra weights zero:
    the same function (a zero-weight, zero-bias channel is relu(0) = 0 and feeds zero weights), but every
    convolution is tensor-core aligned -- the reference's schedules (251, 233, 229, 111, 103, 73, 71 ...) are
    not, which drops cuDNN to its slow paths under TF32 / bf16.  For evaluation throughput only; parity tests
    use the unpadded network."""
    chans = CHANNELS[model.arch]
    up = [((c + multiple - 1) // multiple) * multiple for c in chans]
    dev = next(model.parameters()).device
    out = ConvDecoder.__new__(ConvDecoder)
    nn.Module.__init__(out)
    out.arch, out.system_size, out.depth, out.device = model.arch, model.system_size, model.depth, model.device
    cin_old, cin_new = 2, 2
    with torch.no_grad():
        for k, (c_old, c_new) in enumerate(zip(chans, up), start=1):
            src = getattr(model, "conv%d" % k)
            conv = nn.Conv2d(cin_new, c_new, kernel_size=3, stride=1, padding=src.padding)
            conv.weight.zero_()
            conv.bias.zero_()
            conv.weight[:c_old, :cin_old] = src.weight
            conv.bias[:c_old] = src.bias
            setattr(out, "conv%d" % k, conv)
            cin_old, cin_new = c_old, c_new
        side = model.system_size - 2
        lin = nn.Linear(cin_new * side * side, model.linear1.out_features)
        lin.weight.zero_()
        lin.weight.view(-1, cin_new, side, side)[:, :cin_old] = model.linear1.weight.view(-1, cin_old, side, side)
        lin.bias.copy_(model.linear1.bias)
        out.linear1 = lin
    return out.to(dev).eval()


def strict_fp32():
    """Q-value parity (1e-5 relative, north_star) needs true fp32 convolutions: TF32 off."""
    torch.backends.cuda.matmul.allow_tf32 = False
    torch.backends.cudnn.allow_tf32 = False
