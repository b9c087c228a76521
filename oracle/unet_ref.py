"""torch-fp32 restatement of the reference UNet forward (oracle; test-only).

Follows, layer for layer:
  UNet.forward / get_enc_feature      src/models/unet/unet.py:293-297
  UNetEncoder.forward / get_feature   src/models/unet/unet.py:80-93
  UNetDecoder.forward                 src/models/unet/unet.py:208-224
  PlainBlock                          src/models/unet/blocks.py:66-105
      Conv2d(k3, pad1, bias) -> Dropout2d -> BatchNorm2d(eps 1e-5) -> LeakyReLU(0.01)
in the configuration al_train builds (src/training/al_trainer.py:504-515):
2D, channels [32,64,128,256,512], plain blocks, batch norm, transposed-conv
upsampling, no deep supervision.

It works on a plain state_dict (the reference's key layout), so it travels to
the GPU box where /root/reference is absent.  Pinned against the reference
module in tests/test_oracle_vs_reference.py and against tests/golden/.
"""
from __future__ import annotations

from collections import OrderedDict

import torch
import torch.nn as nn
import torch.nn.functional as F

CHANNELS = (32, 64, 128, 256, 512)
BN_EPS = 1e-5          # blocks.py:98
LEAKY_SLOPE = 0.01     # nn.LeakyReLU default, blocks.py:100

# (state_dict prefix, Cin, Cout, stride) for the 18 PlainBlocks, forward order.
def plain_blocks(in_channels: int = 1):
    out = []
    prev = in_channels
    for l, c in enumerate(CHANNELS):
        out.append((f"encoder.levels.{l}.0.all", prev, c, 1 if l == 0 else 2))
        out.append((f"encoder.levels.{l}.1.all", c, c, 1))
        prev = c
    rev = CHANNELS[::-1]
    for l in range(4):
        c = rev[l + 1]
        out.append((f"decoder.levels.{l}.0.all", 2 * c, c, 1))
        out.append((f"decoder.levels.{l}.1.all", c, c, 1))
    return out


# Dropout sites in forward order; MC-dropout mask channel offsets (sum = 2944).
def dropout_sites(in_channels: int = 1):
    sites, off = [], 0
    for prefix, _cin, cout, _s in plain_blocks(in_channels):
        sites.append((prefix, off, cout))
        off += cout
    return sites, off


def make_state_dict(seed: int, num_out: int, in_channels: int = 1) -> "OrderedDict[str, torch.Tensor]":
    """Seeded default-initialised weights with the reference's key layout.

    Parameters are created in the order the reference constructs its modules
    (encoder levels; then per decoder level: ConvTranspose2d, two blocks; then
    seg_output - unet.py:53-78,131-178), so `torch.manual_seed(seed)` followed
    by this function gives bit-identical tensors to `torch.manual_seed(seed);
    UNet(...)` (checked in tests/test_oracle_vs_reference.py).
    """
    torch.manual_seed(seed)
    sd: "OrderedDict[str, torch.Tensor]" = OrderedDict()

    def add_block(prefix, cin, cout, stride):
        conv = nn.Conv2d(cin, cout, 3, stride=stride, padding=1, bias=True)
        bn = nn.BatchNorm2d(cout, eps=BN_EPS, affine=True)
        sd[f"{prefix}.0.weight"] = conv.weight.detach()
        sd[f"{prefix}.0.bias"] = conv.bias.detach()
        sd[f"{prefix}.2.weight"] = bn.weight.detach()
        sd[f"{prefix}.2.bias"] = bn.bias.detach()
        sd[f"{prefix}.2.running_mean"] = bn.running_mean.clone()
        sd[f"{prefix}.2.running_var"] = bn.running_var.clone()
        sd[f"{prefix}.2.num_batches_tracked"] = bn.num_batches_tracked.clone()

    blocks = plain_blocks(in_channels)
    for b in blocks[:10]:
        add_block(*b)
    rev = CHANNELS[::-1]
    # decoder: reference builds upsample l, then blocks l, appended to separate
    # ModuleLists; state_dict order is levels.* first, then upsamples.*.
    ups = []
    for l in range(4):
        ups.append(nn.ConvTranspose2d(rev[l], rev[l + 1], kernel_size=2, stride=2))
        add_block(*blocks[10 + 2 * l])
        add_block(*blocks[11 + 2 * l])
    for l, up in enumerate(ups):
        sd[f"decoder.upsamples.{l}.weight"] = up.weight.detach()
        sd[f"decoder.upsamples.{l}.bias"] = up.bias.detach()
    seg = nn.Conv2d(CHANNELS[0], num_out, 1)
    sd["decoder.seg_output.weight"] = seg.weight.detach()
    sd["decoder.seg_output.bias"] = seg.bias.detach()
    return sd


def _plain_block(sd, prefix, x, stride, mask=None, training_bn=False, bn_momentum=0.1):
    """conv -> (dropout mask) -> batchnorm -> leaky relu (blocks.py:83-104)."""
    z = F.conv2d(x, sd[f"{prefix}.0.weight"], sd[f"{prefix}.0.bias"], stride=stride, padding=1)
    if mask is not None:  # Dropout2d sits between conv and norm (blocks.py:102)
        z = z * mask[:, :, None, None]
    z = F.batch_norm(
        z, sd[f"{prefix}.2.running_mean"], sd[f"{prefix}.2.running_var"],
        sd[f"{prefix}.2.weight"], sd[f"{prefix}.2.bias"],
        training=training_bn, momentum=bn_momentum, eps=BN_EPS)
    return F.leaky_relu(z, LEAKY_SLOPE)


def unet_encode(sd, x, masks=None, training_bn=False, bn_momentum=0.1):
    """Encoder with skips (unet.py:80-87).  masks: [B, 2944] multipliers or None."""
    blocks = plain_blocks(x.shape[1])
    sites, _ = dropout_sites(x.shape[1])
    skips = []
    for i in range(10):
        prefix, _cin, _cout, stride = blocks[i]
        m = None
        if masks is not None:
            _, off, c = sites[i]
            m = masks[:, off:off + c]
        x = _plain_block(sd, prefix, x, stride, m, training_bn, bn_momentum)
        if i % 2 == 1:
            skips.append(x)
    return skips


def unet_embed(sd, x, masks=None):
    """UNet.get_enc_feature (unet.py:89-93, 296-297): bottleneck -> global avg pool."""
    bott = unet_encode(sd, x, masks)[-1]
    return F.adaptive_avg_pool2d(bott, (1, 1)).view(x.shape[0], -1)


def unet_forward(sd, x, masks=None, training_bn=False, return_features=False):
    """UNet.forward (unet.py:293): logits [B, K, H, W]."""
    return _forward_impl(sd, x, masks, training_bn, 0.1, return_features)


def _forward_impl(sd, x, masks, training_bn, bn_momentum, return_features=False):
    in_ch = x.shape[1]
    blocks = plain_blocks(in_ch)
    sites, _ = dropout_sites(in_ch)
    skips = unet_encode(sd, x, masks, training_bn, bn_momentum)
    y = skips[-1]
    for l in range(4):
        y = F.conv_transpose2d(y, sd[f"decoder.upsamples.{l}.weight"], sd[f"decoder.upsamples.{l}.bias"], stride=2)
        y = torch.cat([skips[3 - l], y], dim=1)  # skip first, upsampled second (unet.py:215)
        for j in range(2):
            i = 10 + 2 * l + j
            prefix, _cin, _cout, stride = blocks[i]
            m = None
            if masks is not None:
                _, off, c = sites[i]
                m = masks[:, off:off + c]
            y = _plain_block(sd, prefix, y, stride, m, training_bn, bn_momentum)
    logits = F.conv2d(y, sd["decoder.seg_output.weight"], sd["decoder.seg_output.bias"])
    if return_features:
        return logits, y, skips
    return logits


class OracleUNet(nn.Module):
    """nn.Module shim over a state_dict so selectors that take `model` work on
    the GPU box without the reference package.  Same state_dict keys, same
    forward / get_enc_feature results as the reference UNet (eval mode)."""

    def __init__(self, sd):
        super().__init__()
        self._keys = list(sd.keys())
        for k, v in sd.items():
            name = k.replace(".", "__")
            if v.dtype.is_floating_point and "running" not in k:
                self.register_parameter(name, nn.Parameter(v.clone(), requires_grad=False))
            else:
                self.register_buffer(name, v.clone())

    def state_dict(self, *a, **k):  # reference key layout
        return OrderedDict((key, getattr(self, key.replace(".", "__")).detach()) for key in self._keys)

    def forward(self, x):
        return unet_forward(self.state_dict(), x)

    def get_enc_feature(self, x):
        return unet_embed(self.state_dict(), x)


def conv_weight_digest(sd) -> str:
    """sha256 over all conv / transposed-conv weights+biases (not BN, not head):
    lets the GPU box check that re-seeding reproduced the golden weights."""
    import hashlib
    h = hashlib.sha256()
    for k, v in sd.items():
        if (k.endswith(".0.weight") or k.endswith(".0.bias") or "upsamples" in k):
            h.update(v.detach().contiguous().numpy().tobytes())
    return h.hexdigest()


GOLDEN_BN_KEYS = (".all.2.weight", ".all.2.bias", ".all.2.running_mean", ".all.2.running_var")


def build_calibrated_state_dict(num_out: int, images_u8: torch.Tensor, seed: int = 1337,
                                affine_seed: int = 4242, logit_std: float = 2.5):
    """Golden-model recipe (run once here by oracle/make_golden.py):
    1. seeded default init (== reference `UNet(...)` under the same seed);
    2. BN gamma ~ U(0.5,1.5), beta ~ N(0,0.3^2) from `affine_seed`;
    3. BN running stats := batch statistics of one train-mode pass over
       `images_u8` (momentum 1.0), so activations are O(1) at every depth;
    4. seg_output.weight rescaled so logits have std `logit_std` on that batch
       (a default-initialised head gives near-uniform softmax everywhere and
       scores that differ only in the 4th digit).
    """
    sd = make_state_dict(seed, num_out)
    g = torch.Generator().manual_seed(affine_seed)
    for k in sd:
        if k.endswith(".all.2.weight"):
            sd[k].copy_(0.5 + torch.rand(sd[k].shape, generator=g))
        elif k.endswith(".all.2.bias"):
            sd[k].copy_(0.3 * torch.randn(sd[k].shape, generator=g))
    x = images_u8.to(torch.float32).div(255.0).unsqueeze(1)
    with torch.no_grad():
        _forward_impl(sd, x, None, True, 1.0)
        y = unet_forward(sd, x[:16])
        sd["decoder.seg_output.weight"].mul_(logit_std / y.std())
    return sd


def golden_state_dict(num_out: int, golden_npz, seed: int = 1337):
    """Rebuild the golden model on any machine: seeded conv weights + the BN
    tensors and heads frozen in tests/golden/model.npz.  Raises if the
    re-seeded conv weights do not hash to the recorded digest."""
    sd = make_state_dict(seed, num_out)
    digest = conv_weight_digest(sd)
    want = str(golden_npz["conv_digest"])
    if digest != want:
        raise RuntimeError(f"re-seeded conv weights differ from the golden ones ({digest[:12]} != {want[:12]})")
    with torch.no_grad():
        for k in sd:
            if k.endswith(GOLDEN_BN_KEYS):
                sd[k].copy_(torch.from_numpy(golden_npz["bn/" + k]))
        sd["decoder.seg_output.weight"].copy_(torch.from_numpy(golden_npz[f"head{num_out}/weight"]))
        sd["decoder.seg_output.bias"].copy_(torch.from_numpy(golden_npz[f"head{num_out}/bias"]))
    return sd
