"""CPU oracle for the Zero-DCE / Zero-DCE++ / SCI inference forward.

TEST INFRASTRUCTURE ONLY.  Nothing in the product path
(`etss_07_imageenhance_b200/`) may import this module; only `tests/`,
`__graft_entry__.smoke()` and `bench.py`'s `cpu_baseline` / `--impl reference`
legs use it, and only as the checker / the timed CPU baseline.

This is a plain-torch (ATen on CPU, fp32) restatement of the reference's own
arithmetic.  Citations are relative to /root/reference/:

* `zero_dce_forward`   follows src/mon/vision/enhance/llie/zero_dce/zero_dce.py:173-191
  (layers :133-139) == upstream src/mon_extra/vision/enhance/llie/zero_dce/model.py:27-50
* `zero_dcepp_forward` follows src/mon/vision/enhance/llie/zero_dce/zero_dcepp.py:182-208
  (layers :140-146, DSConv2d = src/mon/nn/modules/conv/dsconv.py:106-149); for
  scale_factor != 1 the reference class raises (zero_dcepp.py:139 passes the
  factor as `size`), so the *intended* semantics are taken from upstream
  src/mon_extra/vision/enhance/llie/zero_dce++/model.py:39,64-82.
* `sci_forward`        follows src/mon_extra/vision/enhance/llie/sci/model.py:36-45,155-159
  (BatchNorm2d in eval mode).
* `ruas_forward`       follows src/mon_extra/vision/enhance/llie/ruas/model.py:46-66 (SearchBlock),
  :77-96 (IEM), :117-125 (EnhanceNetwork), :142-149 (DenoiseNetwork), :202-206 (Network.forward) with the
  cells of genotypes.py:15-16 and the ops of operations.py:25-54.

Parity pin: the reference holds no tests / golden vectors for this path
("parity unpinned" by the reference's own tests).  The pin we
create is `oracle/pin_against_reference.py`: it executes the reference's own
source files (the `mon` classes under a stub `mon` package, and the upstream
`mon_extra` model files unmodified) in the build container, asserts bit
equality with this restatement on CPU, and writes tests/golden/*.npz.

`bilinear_down_ref` / `bilinear_up_ref` are independent explicit-index numpy
restatements of ATen's `upsample_bilinear2d` index math
(area_pixel_compute_source_index, fp32) used to pin the kernels' index logic.
"""
from __future__ import annotations

import math

import numpy as np
import torch
import torch.nn.functional as F

# --------------------------------------------------------------------------- #
# state_dict layouts
# --------------------------------------------------------------------------- #
DCE_CHANNELS = [(3, 32), (32, 32), (32, 32), (32, 32), (64, 32), (64, 32), (64, 24)]
DCEPP_CHANNELS = [(3, 32), (32, 32), (32, 32), (32, 32), (64, 32), (64, 32), (64, 3)]


def make_dce_state_dict(seed: int, regime: str = "init", num_channels: int = 32, num_iters: int = 8,
                        in_channels: int = 3) -> dict:
    """Seeded Zero-DCE weights.  regime="init": the reference's random init
    (conv weights N(0,0.02), zero_dce.py:152-158; biases torch-default-like
    U(-1/sqrt(fan_in), 1/sqrt(fan_in))).  regime="stress": weights scaled so the
    curve maps span roughly (-0.9, 0.9) and indexing/sign bugs are visible."""
    g = torch.Generator().manual_seed(seed)
    nc = num_channels
    chans = [(in_channels, nc), (nc, nc), (nc, nc), (nc, nc), (2 * nc, nc), (2 * nc, nc),
             (2 * nc, in_channels * num_iters)]
    sd = {}
    for i, (ci, co) in enumerate(chans, 1):
        fan_in = ci * 9
        if regime == "init":
            w = torch.randn(co, ci, 3, 3, generator=g) * 0.02
        else:
            w = torch.randn(co, ci, 3, 3, generator=g) * (1.6 / fan_in ** 0.5)
        bound = 1.0 / fan_in ** 0.5
        b = (torch.rand(co, generator=g) * 2 - 1) * bound
        sd[f"e_conv{i}.weight"] = w
        sd[f"e_conv{i}.bias"] = b
    return sd


def make_dcepp_state_dict(seed: int, regime: str = "init", num_channels: int = 32,
                          in_channels: int = 3) -> dict:
    """Seeded Zero-DCE++ weights in the `mon` key layout
    (e_conv{i}.dw_conv/pw_conv.{weight,bias}, dsconv.py:122-149)."""
    g = torch.Generator().manual_seed(seed)
    nc = num_channels
    chans = [(in_channels, nc), (nc, nc), (nc, nc), (nc, nc), (2 * nc, nc), (2 * nc, nc),
             (2 * nc, in_channels)]
    sd = {}
    for i, (ci, co) in enumerate(chans, 1):
        if regime == "init":
            dw = torch.randn(ci, 1, 3, 3, generator=g) * 0.02
            pw = torch.randn(co, ci, 1, 1, generator=g) * 0.02
        else:
            dw = torch.randn(ci, 1, 3, 3, generator=g) * (1.4 / 3.0)
            pw = torch.randn(co, ci, 1, 1, generator=g) * (1.4 / ci ** 0.5)
        dwb = (torch.rand(ci, generator=g) * 2 - 1) / 3.0
        pwb = (torch.rand(co, generator=g) * 2 - 1) / ci ** 0.5
        sd[f"e_conv{i}.dw_conv.weight"] = dw
        sd[f"e_conv{i}.dw_conv.bias"] = dwb
        sd[f"e_conv{i}.pw_conv.weight"] = pw
        sd[f"e_conv{i}.pw_conv.bias"] = pwb
    return sd


def make_sci_state_dict(seed: int, regime: str = "init") -> dict:
    """Seeded SCI EnhanceNetwork(layers=1, channels=3) weights, keys as in
    sci/model.py:16-34 (without the `enhance.` prefix of Finetunemodel)."""
    g = torch.Generator().manual_seed(seed)
    s = 0.02 if regime == "init" else 0.35
    sd = {}
    for name in ("in_conv.0", "conv.0", "out_conv.0"):
        sd[f"{name}.weight"] = torch.randn(3, 3, 3, 3, generator=g) * s
        sd[f"{name}.bias"] = (torch.rand(3, generator=g) * 2 - 1) * (0.0 if regime == "init" else 0.2)
    if regime == "init":
        sd["conv.1.weight"] = 1.0 + torch.randn(3, generator=g) * 0.02
        sd["conv.1.bias"] = torch.zeros(3)
        sd["conv.1.running_mean"] = torch.zeros(3)
        sd["conv.1.running_var"] = torch.ones(3)
    else:
        sd["conv.1.weight"] = 1.0 + torch.randn(3, generator=g) * 0.3
        sd["conv.1.bias"] = torch.randn(3, generator=g) * 0.2
        sd["conv.1.running_mean"] = torch.randn(3, generator=g) * 0.1
        sd["conv.1.running_var"] = 0.5 + torch.rand(3, generator=g)
    sd["conv.1.num_batches_tracked"] = torch.tensor(0, dtype=torch.long)
    return sd


# RUAS cells (ruas/genotypes.py:15-16): op of c1_d, c1_r, c2_d, c2_r, c3_d, c3_r, c4
RUAS_IEM = ("skip_connect", "resconv_1x1", "resdilconv_3x3", "conv_3x3", "conv_3x3", "skip_connect", "conv_3x3")
RUAS_NRM = ("resconv_1x1", "resconv_1x1", "resdilconv_3x3", "skip_connect", "resconv_1x1", "resconv_1x1",
            "skip_connect")
RUAS_SLOTS = ("c1_d", "c1_r", "c2_d", "c2_r", "c3_d", "c3_r", "c4")


def make_ruas_state_dict(seed: int, regime: str = "init") -> dict:
    """Seeded weights with the key names of upstream `Network` (ruas/model.py:152-206): `enhance_net.iems.{i}.cell.*`
    (3 cells, 3 channels) and `denoise_net.{stem,nrms.{i}.*,activate.0}` (3 cells, 6 channels).  skip_connect slots
    own no parameters.  regime="stress" scales the weights up so that t leaves the sigmoid's flat middle."""
    g = torch.Generator().manual_seed(seed)
    scale = [1.0 if regime == "init" else 3.0]
    sd = {}

    def conv(key, co, ci, k):
        bound = 1.0 / math.sqrt(ci * k * k)
        sd[key + ".weight"] = (torch.rand(co, ci, k, k, generator=g) * 2 - 1) * bound * scale[0]
        sd[key + ".bias"] = (torch.rand(co, generator=g) * 2 - 1) * bound

    def cell(prefix, ch, geno):
        for slot, op in zip(RUAS_SLOTS, geno):
            if op != "skip_connect":
                conv(f"{prefix}.{slot}.op", ch, ch, 1 if op.endswith("1x1") else 3)
        conv(f"{prefix}.c5", ch, 4 * ch, 1)

    for i in range(3):
        cell(f"enhance_net.iems.{i}.cell", 3, RUAS_IEM)
    scale[0] = 1.0 if regime == "init" else 1.3  # three stacked 24 -> 6 cells amplify: keep the noise map O(1)
    conv("denoise_net.stem", 6, 3, 3)
    for i in range(3):
        cell(f"denoise_net.nrms.{i}", 6, RUAS_NRM)
    conv("denoise_net.activate.0", 3, 6, 3)
    return sd


def synthetic_lowlight(B: int, H: int, W: int, seed: int, device="cpu") -> torch.Tensor:
    """Seeded synthetic low-light frame batch: dark gamma'd
    noise blended with a smooth gradient field; a few pixels forced to exactly
    0 and 1.  NCHW fp32 in [0,1]."""
    g = torch.Generator().manual_seed(seed)
    u = torch.rand(B, 3, H, W, generator=g)
    noise = 0.02 + 0.33 * u.pow(2.2)
    field = torch.rand(B, 3, 9, 16, generator=g)
    smooth = 0.45 * F.interpolate(field, size=(H, W), mode="bilinear", align_corners=True)
    x = (0.5 * noise + 0.5 * smooth).clamp_(0, 1)
    x[:, :, 0, 0] = 0.0
    x[:, :, -1, -1] = 1.0
    if H > 2 and W > 2:
        x[:, 1, H // 2, W // 2] = 1.0
        x[:, 2, H // 2, W // 2 - 1] = 0.0
    return x.to(device)


# --------------------------------------------------------------------------- #
# forwards
# --------------------------------------------------------------------------- #
def curve_shared(x: torch.Tensor, a: torch.Tensor, num_iters: int = 8) -> torch.Tensor:
    """zero_dcepp.py:201-203 — y = y + a*(y**2 - y), same map every iteration."""
    y = x
    for _ in range(num_iters):
        y = y + a * (torch.pow(y, 2) - y)
    return y


def curve_per_iter(x: torch.Tensor, a: torch.Tensor, num_iters: int = 8) -> torch.Tensor:
    """zero_dce.py:183-188 — one C-channel map per iteration (split along dim 1)."""
    c = x.shape[1]
    a_s = torch.split(a, c, dim=1)
    y = x
    for i in range(num_iters):
        y = y + a_s[i] * (torch.pow(y, 2) - y)
    return y


@torch.no_grad()
def gce_curve_loop(image: torch.Tensor, adjust: torch.Tensor, num_iters: int, bam_gamma: float) -> dict:
    """GCENet.forward's enhancement loop, src/mon/vision/enhance/llie/gcenet/gcenet.py:469-484, with the brightness
    attention map of src/mon/nn/modules/prior/image.py:269-272 (tensor branch, denoise_ksize=None): V is kornia's HSV
    value = the channel maximum.  `enhanced` is the value BEFORE the guided filter (gcenet.py:485)."""
    if bam_gamma in [None, 0.0]:
        enhanced, bam, bright, dark = image, None, None, None
        for _ in range(num_iters):
            enhanced = enhanced + adjust * (torch.pow(enhanced, 2) - enhanced)
    else:
        v = image.max(dim=-3, keepdim=True)[0]
        bam = torch.pow((1 - v), bam_gamma)
        enhanced, bright, dark = image, None, None
        for _ in range(num_iters):
            bright = enhanced * (1 - bam)
            dark = enhanced * bam
            enhanced = bright + dark + adjust * (torch.pow(dark, 2) - dark)
    return {"enhanced": enhanced, "bam": bam, "bright": bright, "dark": dark}


@torch.no_grad()
def zero_dce_forward(x: torch.Tensor, sd: dict, num_iters: int = 8) -> dict:
    """ZeroDCE_RE.forward (zero_dce.py:173-191).  Returns the reference's dict
    (adjust_0.., adjust, enhanced — in that insertion order)."""
    def conv(t, i):
        return F.conv2d(t, sd[f"e_conv{i}.weight"], sd[f"e_conv{i}.bias"], stride=1, padding=1)
    x1 = F.relu(conv(x, 1))
    x2 = F.relu(conv(x1, 2))
    x3 = F.relu(conv(x2, 3))
    x4 = F.relu(conv(x3, 4))
    x5 = F.relu(conv(torch.cat([x3, x4], 1), 5))
    x6 = F.relu(conv(torch.cat([x2, x5], 1), 6))
    x_r = torch.tanh(conv(torch.cat([x1, x6], 1), 7))
    x_rs = torch.split(x_r, x.shape[1], dim=1)
    y = x
    out = {}
    for i in range(num_iters):
        y = y + x_rs[i] * (torch.pow(y, 2) - y)
        out[f"adjust_{i}"] = x_rs[i]
    out["adjust"] = x_r
    out["enhanced"] = y
    return out


@torch.no_grad()
def zero_dce_v_core(x: torch.Tensor, sd: dict, down_size: int = 256, num_iters: int = 15) -> dict:
    """ZeroDCEV.forward (src/mon/vision/enhance/llie/zero_dce/zero_dce_v.py:172-193) up to `image_v_fixed_lr`:
    `core.rgb_to_v` = V of kornia's rgb_to_hsv = the maximum over the channels (src/mon/core/image/color.py:198-221),
    `interpolate_image` = F.interpolate(size=(down_size, down_size)) (nearest, :204-206), the 7-layer DCE-Net with
    in_channels = 1 (:179-186) and `num_iters` curve steps on the low-resolution V (:189-192)."""
    image_v = x.max(dim=1, keepdim=True)[0]
    image_v_lr = F.interpolate(image_v, size=(down_size, down_size))

    def conv(t, i):
        return F.conv2d(t, sd[f"e_conv{i}.weight"], sd[f"e_conv{i}.bias"], stride=1, padding=1)
    x1 = F.relu(conv(image_v_lr, 1))
    x2 = F.relu(conv(x1, 2))
    x3 = F.relu(conv(x2, 3))
    x4 = F.relu(conv(x3, 4))
    x5 = F.relu(conv(torch.cat([x3, x4], 1), 5))
    x6 = F.relu(conv(torch.cat([x2, x5], 1), 6))
    x_r = torch.tanh(conv(torch.cat([x1, x6], 1), 7))
    x_rs = torch.split(x_r, 1, dim=1)
    y = image_v_lr
    for i in range(num_iters):
        y = y + x_rs[i] * (torch.pow(y, 2) - y)
    return {"image_v": image_v, "image_v_lr": image_v_lr, "adjust": x_r, "image_v_fixed_lr": y}


# ----------------------------------------------------------------------------- guided filter / HSV (ZeroDCE-V tail)
def _gf_diff(t: torch.Tensor, r: int, dim: int) -> torch.Tensor:
    """diff_x (dim=2) / diff_y (dim=3) of src/mon/vision/filtering/box_filter.py:26-66: window sums from a cumulative sum."""
    n = t.shape[dim]
    sl = lambda a, b: t.narrow(dim, a, b - a)  # noqa: E731
    left = sl(r, 2 * r + 1)
    middle = sl(2 * r + 1, n) - sl(0, n - 2 * r - 1)
    right = sl(n - 1, n) - sl(n - 2 * r - 1, n - r - 1)
    return torch.cat([left, middle, right], dim=dim)


def box_filter_cumsum(t: torch.Tensor, r: int) -> torch.Tensor:
    """box_filter, tensor branch (box_filter.py:69-117): diff_y(diff_x(t.cumsum(2), r).cumsum(3), r).  On the CPU ATen's
    cumsum accumulates fp32 inputs in double and rounds every prefix to fp32."""
    return _gf_diff(_gf_diff(t.cumsum(dim=2), r, 2).cumsum(dim=3), r, 3)


@torch.no_grad()
def fast_guided_filter(x_lr: torch.Tensor, y_lr: torch.Tensor, x_hr: torch.Tensor, radius: int = 1,
                       eps: float = 1e-8, return_coeffs: bool = False):
    """FastGuidedFilter.forward(x_lr, y_lr, x_hr), src/mon/vision/filtering/guided_filter.py:171-203: the linear model
    y = A x + b fitted in (2r+1)^2 windows at low resolution, A and b resized bicubically (align_corners=True) to the
    full resolution and applied to the full-resolution guide."""
    h, w = x_lr.shape[-2:]
    n = box_filter_cumsum(x_lr.new_ones(1, 1, h, w), radius)
    mean_x = box_filter_cumsum(x_lr, radius) / n
    mean_y = box_filter_cumsum(y_lr, radius) / n
    cov_xy = box_filter_cumsum(x_lr * y_lr, radius) / n - mean_x * mean_y
    var_x = box_filter_cumsum(x_lr * x_lr, radius) / n - mean_x * mean_x
    a = cov_xy / (var_x + eps)
    b = mean_y - a * mean_x
    if return_coeffs:
        return a, b
    mean_a = F.interpolate(a, x_hr.shape[-2:], mode="bicubic", align_corners=True)
    mean_b = F.interpolate(b, x_hr.shape[-2:], mode="bicubic", align_corners=True)
    return mean_a * x_hr + mean_b


@torch.no_grad()
def guided_filter(x: torch.Tensor, y: torch.Tensor, radius: int, eps: float = 1e-8) -> torch.Tensor:
    """GuidedFilter(radius, eps).forward(image=x, guide=y), src/mon/vision/filtering/guided_filter.py:99-146, as GCENet
    calls it (`self.gf(image, enhanced)`, gcenet.py:485, GuidedFilter(radius=3, eps=1e-4), :391).  The class builds its
    box filter as `BoxFilter(self.radius)` -- positionally, i.e. as KERNEL SIZE (box_filter.py:161-171) -- so the window
    radius is int((radius - 1) / 2): 1 for GCENet's radius 3."""
    r = int((radius - 1) / 2)
    h, w = x.shape[-2:]
    n = box_filter_cumsum(x.new_ones(1, 1, h, w), r)
    mean_x = box_filter_cumsum(x, r) / n
    mean_y = box_filter_cumsum(y, r) / n
    cov_xy = box_filter_cumsum(x * y, r) / n - mean_x * mean_y
    var_x = box_filter_cumsum(x * x, r) / n - mean_x * mean_x
    a = cov_xy / (var_x + eps)
    b = mean_y - a * mean_x
    mean_a = box_filter_cumsum(a, r) / n
    mean_b = box_filter_cumsum(b, r) / n
    return mean_a * x + mean_b


def rgb_to_hsv_kornia(image: torch.Tensor, eps: float = 1e-8) -> torch.Tensor:
    """kornia.color.rgb_to_hsv as published (kornia is a dependency of the reference, unpinned in env/*.yaml and absent
    here: PARITY UNPINNED for this function).  H in [0, 2 pi], S = delta / (max + eps), V = max."""
    max_rgb, argmax_rgb = image.max(-3)
    min_rgb = image.min(-3)[0]
    deltac = max_rgb - min_rgb
    v = max_rgb
    s = deltac / (max_rgb + eps)
    deltac = torch.where(deltac == 0, torch.ones_like(deltac), deltac)
    rc, gc, bc = torch.unbind(max_rgb.unsqueeze(-3) - image, dim=-3)
    h1 = bc - gc
    h2 = (rc - bc) + 2.0 * deltac
    h3 = (gc - rc) + 4.0 * deltac
    h = torch.stack((h1, h2, h3), dim=-3) / deltac.unsqueeze(-3)
    h = torch.gather(h, dim=-3, index=argmax_rgb.unsqueeze(-3)).squeeze(-3)
    h = (h / 6.0) % 1.0
    h = 2.0 * math.pi * h
    return torch.stack((h, s, v), dim=-3)


def hsv_to_rgb_kornia(image: torch.Tensor) -> torch.Tensor:
    """kornia.color.hsv_to_rgb as published (PARITY UNPINNED, see rgb_to_hsv_kornia)."""
    h = image[..., 0, :, :] / (2 * math.pi)
    s = image[..., 1, :, :]
    v = image[..., 2, :, :]
    hi = torch.floor(h * 6) % 6
    f = ((h * 6) % 6) - hi
    one = torch.tensor(1.0)
    p = v * (one - s)
    q = v * (one - f * s)
    t = v * (one - (one - f) * s)
    hi = hi.long()
    indices = torch.stack([hi, hi + 6, hi + 12], dim=-3)
    out = torch.stack((v, q, p, p, t, v, t, v, v, q, p, p, p, p, t, v, v, q), dim=-3)
    return torch.gather(out, -3, indices)


@torch.no_grad()
def zero_dce_v_forward(x: torch.Tensor, sd: dict, down_size: int = 256, num_iters: int = 15) -> dict:
    """ZeroDCEV.forward, zero_dce_v.py:172-203, all six outputs.  The reference runs with batch 1 only
    (`replace_v_component`, :221-230); a batch is treated here as a loop of batch-1 calls, so the final division is by
    each frame's own maximum.  `image_v_fixed` is pinned against the reference's FastGuidedFilter; `enhanced` goes
    through the kornia HSV restatements above and is unpinned."""
    core = zero_dce_v_core(x, sd, down_size, num_iters)
    v_fixed = torch.clip(fast_guided_filter(core["image_v_lr"], core["image_v_fixed_lr"], core["image_v"], radius=1), 0, 1)
    hsv = rgb_to_hsv_kornia(x)
    hsv = torch.cat([hsv[:, :2], v_fixed], dim=1)
    rgb = hsv_to_rgb_kornia(hsv)
    rgb = rgb / rgb.amax(dim=(1, 2, 3), keepdim=True)
    return {**core, "image_v_fixed": v_fixed, "enhanced": rgb}


@torch.no_grad()
def zero_dce_forward_bf16_operands(x: torch.Tensor, sd: dict, num_iters: int = 8) -> dict:
    """`zero_dce_forward` with the OPERAND rounding of the tensor-core mode: every conv input and every
    weight is rounded to bf16 (round-to-nearest-even), products/accumulation/bias/activation stay fp32.
    Not a reference path — a test helper that separates layout/indexing bugs (large errors against this)
    from the expected bf16 quantisation noise (the PSNR contract is checked against `zero_dce_forward`)."""
    r = lambda t: t.to(torch.bfloat16).to(torch.float32)  # noqa: E731

    def conv(t, i):
        return F.conv2d(r(t), r(sd[f"e_conv{i}.weight"]), sd[f"e_conv{i}.bias"], stride=1, padding=1)
    x1 = F.relu(conv(x, 1))
    x2 = F.relu(conv(x1, 2))
    x3 = F.relu(conv(x2, 3))
    x4 = F.relu(conv(x3, 4))
    x5 = F.relu(conv(torch.cat([x3, x4], 1), 5))
    x6 = F.relu(conv(torch.cat([x2, x5], 1), 6))
    x_r = torch.tanh(conv(torch.cat([x1, x6], 1), 7))
    return {"adjust": x_r, "enhanced": curve_per_iter(x, x_r, num_iters)}


@torch.no_grad()
def zero_dcepp_lowres_net(x_down: torch.Tensor, sd: dict) -> torch.Tensor:
    """The 7 depthwise-separable layers + tanh (zero_dcepp.py:190-196)."""
    def ds(t, i):
        c = t.shape[1]
        t = F.conv2d(t, sd[f"e_conv{i}.dw_conv.weight"], sd[f"e_conv{i}.dw_conv.bias"], stride=1, padding=1, groups=c)
        return F.conv2d(t, sd[f"e_conv{i}.pw_conv.weight"], sd[f"e_conv{i}.pw_conv.bias"])
    x1 = F.relu(ds(x_down, 1))
    x2 = F.relu(ds(x1, 2))
    x3 = F.relu(ds(x2, 3))
    x4 = F.relu(ds(x3, 4))
    x5 = F.relu(ds(torch.cat([x3, x4], 1), 5))
    x6 = F.relu(ds(torch.cat([x2, x5], 1), 6))
    return torch.tanh(ds(torch.cat([x1, x6], 1), 7))


@torch.no_grad()
def zero_dcepp_forward(x: torch.Tensor, sd: dict, scale_factor: float = 1.0, num_iters: int = 8) -> dict:
    """ZeroDCEpp_RE.forward (zero_dcepp.py:182-208) with upstream's working
    upsample for scale_factor != 1 (zero_dce++/model.py:39,68,80)."""
    x_down = x
    if scale_factor != 1:
        x_down = F.interpolate(x, scale_factor=1 / scale_factor, mode="bilinear")
    x_r = zero_dcepp_lowres_net(x_down, sd)
    if scale_factor != 1:
        x_r = torch.nn.UpsamplingBilinear2d(scale_factor=scale_factor)(x_r)
    y = curve_shared(x, x_r, num_iters)
    return {"adjust": x_r, "enhanced": y}


@torch.no_grad()
def sci_forward(x: torch.Tensor, sd: dict, eps: float = 1e-5) -> tuple:
    """Finetunemodel.forward (sci/model.py:155-159) over EnhanceNetwork.forward
    (:36-45), BatchNorm2d in eval mode.  Returns (illu, r)."""
    fea = F.relu(F.conv2d(x, sd["in_conv.0.weight"], sd["in_conv.0.bias"], padding=1))
    t = F.conv2d(fea, sd["conv.0.weight"], sd["conv.0.bias"], padding=1)
    t = F.batch_norm(t, sd["conv.1.running_mean"], sd["conv.1.running_var"], sd["conv.1.weight"], sd["conv.1.bias"],
                     training=False, eps=eps)
    fea = fea + F.relu(t)
    fea = torch.sigmoid(F.conv2d(fea, sd["out_conv.0.weight"], sd["out_conv.0.bias"], padding=1))
    illu = torch.clamp(fea + x, 0.0001, 1)
    r = torch.clamp(x / illu, 0, 1)
    return illu, r


def _ruas_op(x, sd, key, op):
    """operations.py:5-23 -> ConvBlock (:25-33) / ResBlock (:35-45) / Identity (:48-54)."""
    if op == "skip_connect":
        return x
    w, b = sd[key + ".op.weight"], sd[key + ".op.bias"]
    dil = 2 if "dil" in op else 1
    pad = (w.shape[-1] - 1) // 2 * dil
    y = F.conv2d(x, w, b, padding=pad, dilation=dil)
    return y + x if op.startswith("res") else y


def _ruas_cell(x, sd, prefix, geno):
    """SearchBlock.forward (ruas/model.py:46-66)."""
    act = lambda v: F.leaky_relu(v, 0.05)
    d1 = act(_ruas_op(x, sd, prefix + ".c1_d", geno[0]))
    r1 = act(_ruas_op(x, sd, prefix + ".c1_r", geno[1]) + x)
    d2 = act(_ruas_op(r1, sd, prefix + ".c2_d", geno[2]))
    r2 = act(_ruas_op(r1, sd, prefix + ".c2_r", geno[3]) + r1)
    d3 = act(_ruas_op(r2, sd, prefix + ".c3_d", geno[4]))
    r3 = act(_ruas_op(r2, sd, prefix + ".c3_r", geno[5]) + r2)
    r4 = act(_ruas_op(r3, sd, prefix + ".c4", geno[6]))
    return F.conv2d(torch.cat([d1, d2, d3, r4], dim=1), sd[prefix + ".c5.weight"], sd[prefix + ".c5.bias"])


def ruas_max_operation(x):
    """IEM.max_operation (ruas/model.py:77-82): 2x2 max over (r, r+1) x (c, c+1), zero beyond the last row/column."""
    x = F.pad(x, (0, 1, 0, 1))
    x = torch.max(x[:, :, :-1, :], x[:, :, 1:, :])
    return torch.max(x[:, :, :, :-1], x[:, :, :, 1:])


@torch.no_grad()
def ruas_forward(x: torch.Tensor, sd: dict, denoise: bool = True) -> tuple:
    """Network.forward (ruas/model.py:202-206): returns (u_list, t_list); with `denoise` the denoised u is appended to
    u_list and the noise map returned as a third value."""
    u_list, t_list = [], []
    u = torch.ones_like(x)
    for k in range(3):  # EnhanceNetwork.forward :117-125, IEM.forward :84-95
        t_hat = ruas_max_operation(x) if k == 0 else ruas_max_operation(u) - 0.5 * (u - x)
        t = torch.sigmoid(_ruas_cell(t_hat, sd, f"enhance_net.iems.{k}.cell", RUAS_IEM))
        t = torch.clamp(t, 0.001, 1.0)
        u = torch.clamp(x / t, 0.0, 1.0)
        u_list.append(u)
        t_list.append(t)
    if not denoise:
        return u_list, t_list, None
    feat = F.conv2d(u, sd["denoise_net.stem.weight"], sd["denoise_net.stem.bias"], padding=1)  # :142-149
    for i in range(3):
        feat = _ruas_cell(feat, sd, f"denoise_net.nrms.{i}", RUAS_NRM)
    n = F.conv2d(feat, sd["denoise_net.activate.0.weight"], sd["denoise_net.activate.0.bias"], padding=1)
    u_list.append(u - n)
    return u_list, t_list, n


# --------------------------------------------------------------------------- #
# independent explicit-index restatements of ATen's bilinear resize (numpy)
# --------------------------------------------------------------------------- #
def bilinear_down_ref(x: np.ndarray, sf: int) -> np.ndarray:
    """F.interpolate(x, scale_factor=1/sf, mode='bilinear') for integer sf:
    align_corners=False, scale used directly: src = sf*(dst+0.5)-0.5, clamped
    at 0 (ATen area_pixel_compute_source_index).  x: [B,C,H,W] float32."""
    B, C, H, W = x.shape
    h, w = int(np.floor(H * (1.0 / sf))), int(np.floor(W * (1.0 / sf)))
    scale = np.float32(1.0 / (1.0 / sf))  # ATen: 1/scale_factor when given
    out = np.empty((B, C, h, w), np.float32)

    def idx(d, size):
        src = np.float32(scale) * (np.float32(d) + np.float32(0.5)) - np.float32(0.5)
        src = max(np.float32(0), src)
        i0 = int(src)
        i1 = i0 + (1 if i0 < size - 1 else 0)
        l1 = np.float32(src - np.float32(i0))
        return i0, i1, np.float32(1) - l1, l1

    for yy in range(h):
        y0, y1, wy0, wy1 = idx(yy, H)
        for xx in range(w):
            x0, x1, wx0, wx1 = idx(xx, W)
            out[:, :, yy, xx] = wy0 * (wx0 * x[:, :, y0, x0] + wx1 * x[:, :, y0, x1]) + \
                wy1 * (wx0 * x[:, :, y1, x0] + wx1 * x[:, :, y1, x1])
    return out


def bilinear_up_ref(a: np.ndarray, H: int, W: int) -> np.ndarray:
    """nn.UpsamplingBilinear2d(scale_factor=sf) == interpolate(size=(H,W),
    align_corners=True): src = dst*(in-1)/(out-1) in fp32."""
    B, C, h, w = a.shape
    out = np.empty((B, C, H, W), np.float32)
    sh = np.float32(h - 1) / np.float32(H - 1) if H > 1 else np.float32(0)
    sw = np.float32(w - 1) / np.float32(W - 1) if W > 1 else np.float32(0)

    def idx(d, scale, size):
        src = np.float32(scale * np.float32(d))
        i0 = int(src)
        i1 = i0 + (1 if i0 < size - 1 else 0)
        l1 = np.float32(src - np.float32(i0))
        return i0, i1, np.float32(1) - l1, l1

    xs = [idx(X, sw, w) for X in range(W)]
    for Y in range(H):
        y0, y1, wy0, wy1 = idx(Y, sh, h)
        for X in range(W):
            x0, x1, wx0, wx1 = xs[X]
            out[:, :, Y, X] = wy0 * (wx0 * a[:, :, y0, x0] + wx1 * a[:, :, y0, x1]) + \
                wy1 * (wx0 * a[:, :, y1, x0] + wx1 * a[:, :, y1, x1])
    return out


def infer_resize(x: torch.Tensor, size=None, divisible_by: int = 32) -> torch.Tensor:
    """core.resize(image, size | divisible_by=32) for tensors (core/image/geometry/transform.py:104-235 with
    core/image/utils.py:381-382: ceil to the next multiple) -> kornia.geometry.transform.resize ->
    F.interpolate(size=..., mode="bilinear", align_corners=None, antialias=False)."""
    if size is None:
        h, w = x.shape[-2:]
        size = (int(math.ceil(h / divisible_by) * divisible_by), int(math.ceil(w / divisible_by) * divisible_by))
    return F.interpolate(x, size=tuple(size), mode="bilinear", align_corners=None, antialias=False)


def infer(forward, x: torch.Tensor) -> dict:
    """VisionModel.infer (vision/model.py:76-127) around a `forward(x) -> dict` restatement: resize up to a multiple
    of 32, forward, resize every image output back to the input size."""
    h0, w0 = x.shape[-2:]
    out = forward(infer_resize(x))
    return {k: (infer_resize(v, (h0, w0)) if isinstance(v, torch.Tensor) and v.dim() == 4 and
                tuple(v.shape[-2:]) != (h0, w0) else v) for k, v in out.items()}


# --------------------------------------------------------------------------- #
# frame format: u8 HWC <-> fp32 NCHW
# --------------------------------------------------------------------------- #
def u8hwc_to_f32nchw(frames: torch.Tensor) -> torch.Tensor:
    """`to_image_tensor(image, normalize=True)` (src/mon/core/image/utils.py:801-848) for a batch of HWC uint8 frames:
    channel-first, then `normalize_image` = `normalize_image_by_range(min=0, max=255, new_min=0, new_max=1)`
    (src/mon/core/image/photometry.py:232-288): cast to the default dtype (fp32), `ratio = (1 - 0) / (255 - 0)` as a
    Python double, `image = (image - 0.0) * ratio + 0.0` -- i.e. ONE fp32 multiply by fl32(1/255), not a division
    (the two differ in the last bit for 126 of the 256 byte values).  [B,H,W,3] uint8 -> [B,3,H,W] fp32."""
    image = frames.permute(0, 3, 1, 2)
    image = image.to(dtype=torch.float32)
    ratio = (1.0 - 0.0) / (255.0 - 0.0)
    return ((image - 0.0) * ratio + 0.0).contiguous()


def f32nchw_to_u8hwc(image: torch.Tensor) -> torch.Tensor:
    """`to_image_nparray(image, denormalize=True)` (src/mon/core/image/utils.py:766-800): numpy fp32,
    `denormalize_image` = `(image - 0.0) * 255.0 + 0.0` in fp32 (photometry.py:266-270,277-283), `.astype(np.uint8)`
    (truncation toward zero; no rounding, no clamp), channel-last.  Values outside [0, 1] are implementation-defined
    in the reference (C float -> uint8 cast); the curve's outputs never leave [0, 1].  [B,3,H,W] fp32 -> [B,H,W,3] u8."""
    a = image.detach().cpu().numpy().astype(np.float32)
    ratio = (255.0 - 0.0) / (1.0 - 0.0)
    a = ((a - 0.0) * ratio + 0.0).astype(np.uint8)
    return torch.from_numpy(np.ascontiguousarray(a.transpose(0, 2, 3, 1)))


def psnr(a: torch.Tensor, b: torch.Tensor) -> float:
    mse = torch.mean((a.double() - b.double()) ** 2).item()
    if mse == 0:
        return float("inf")
    return 10.0 * np.log10(1.0 / mse)
