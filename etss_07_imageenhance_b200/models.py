"""Drop-in model classes for the `mon` Zero-DCE family (and upstream SCI) backed by the sm_100a kernels.

Mirrors, for this path only, the reference's plugin surface:
  * registry entries `MODELS["zero_dce"]["zero_dce_re"]`, `["zero_dce++_re"]`
    (zero_dce.py:86, zero_dcepp.py:87)
  * constructor signatures / attribute names / defaults (zero_dce.py:107-150, zero_dcepp.py:111-155,
    base ctor src/mon/nn/model.py:237-277) and `weights` = None | path | dict handling (model.py:99-142,381-426)
  * `forward(datapoint: dict) -> dict` with the reference's keys in the reference's insertion order
    (zero_dce.py:173-191, zero_dcepp.py:182-208); `infer` adds "time" (src/mon/vision/model.py:76-127)
  * `state_dict` layout incl. the four fixed `loss.loss_spa.weight_*` tensors
    (src/mon/nn/loss/image.py:916-945), so `mon` checkpoints load with strict=True.

The arithmetic is NOT torch: `forward` calls `torch.ops.b200dce.*` (ops.py -> C ABI -> CUDA).  A CPU
tensor, a missing library or an unsupported hyper-parameter raises; nothing falls back.
"""
from __future__ import annotations

import time
from pathlib import Path
from typing import Any

import torch
from torch import nn

from . import ops
from ._lib import PRECISIONS
from .registry import MODELS

__all__ = ["DSConv2d", "DepthwiseSeparableConv2d", "ZeroDCELoss", "ZeroDCE_RE", "ZeroDCEV", "ZeroDCEpp_RE", "SCI_RE", "RUAS_RE",
           "remap_upstream_dcepp_keys", "WEIGHTS_EXTENSIONS"]

WEIGHTS_EXTENSIONS = (".pt", ".pth", ".onnx")  # src/mon/globals.py:709


# --------------------------------------------------------------------------------------------- layers
class DepthwiseSeparableConv2d(nn.Module):
    """Parameter container with the reference's names (src/mon/nn/modules/conv/dsconv.py:106-149):
    `dw_conv` = Conv2d(C, C, k, groups=C), `pw_conv` = Conv2d(C, O, 1)."""

    def __init__(self, in_channels: int, out_channels: int, kernel_size: int = 3, stride: int = 1, padding: int = 1,
                 bias: bool = True):
        super().__init__()
        self.dw_conv = nn.Conv2d(in_channels, in_channels, kernel_size, stride, padding, groups=in_channels, bias=bias)
        self.pw_conv = nn.Conv2d(in_channels, out_channels, 1, 1, 0, bias=bias)


DSConv2d = DepthwiseSeparableConv2d  # dsconv.py:229


class _SpatialConsistencyKernels(nn.Module):
    """Only the four fixed, non-trainable 3x3 kernels that `SpatialConsistencyLoss` registers as
    parameters (src/mon/nn/loss/image.py:916-945); they are part of every `mon` checkpoint."""

    def __init__(self):
        super().__init__()
        def k(rows):
            return nn.Parameter(torch.tensor(rows, dtype=torch.float32).view(1, 1, 3, 3), requires_grad=False)
        self.weight_left = k([[0, 0, 0], [-1, 1, 0], [0, 0, 0]])
        self.weight_right = k([[0, 0, 0], [0, 1, -1], [0, 0, 0]])
        self.weight_up = k([[0, -1, 0], [0, 1, 0], [0, 0, 0]])
        self.weight_down = k([[0, 0, 0], [0, 1, 0], [0, -1, 0]])


class ZeroDCELoss(nn.Module):
    """state_dict-compatible stand-in for the training loss (zero_dce.py:34-79).  Training is out of
    scope for this path; calling it raises."""

    def __init__(self, *args, **kwargs):
        super().__init__()
        self.loss_spa = _SpatialConsistencyKernels()

    def forward(self, *args, **kwargs):
        raise NotImplementedError("the B200 drop-in covers the inference forward only (training loss is out of scope)")


# --------------------------------------------------------------------------------------------- base
def _is_weights_file(p) -> bool:
    p = Path(p)
    return p.is_file() and p.suffix.lower() in WEIGHTS_EXTENSIONS


class _EnhancementModel(nn.Module):
    """The slice of mon's `Model` / `VisionModel` / `ImageEnhancementModel` contract this path needs
    (src/mon/nn/model.py:172-426, src/mon/vision/model.py:76-127, src/mon/vision/enhance/base.py:29-63)."""

    arch: str = ""
    tasks: list = ["llie"]
    schemes: list = ["unsupervised", "zero_reference"]
    zoo: dict = {}
    model_dir: Path = Path(__file__).resolve().parent

    def __init__(self, name: str = None, fullname: str = None, root: Any = Path(), in_channels: int = 3,
                 out_channels: int = None, num_classes: int = None, weights: Any = None, loss: Any = None,
                 metrics: Any = None, optimizers: Any = None, debug: bool = True, verbose: bool = True,
                 *args, **kwargs):
        super().__init__()
        self.debug = debug
        self.verbose = verbose
        self.name = name
        self.fullname = fullname or name
        self.root = root
        self.in_channels = in_channels
        self.out_channels = out_channels or num_classes or self.in_channels
        self.num_classes = num_classes
        self.weights = None
        self.assign_weights(weights)
        self.loss = None
        self.train_metrics = self.val_metrics = self.test_metrics = None
        self.optims = optimizers
        self.predicting = True
        self._packed = None
        self._packed_key = None

    # ---- weights (model.py:381-426, 99-142)
    def assign_weights(self, weights: Any, overwrite: bool = False):
        if weights is None:
            pass
        elif isinstance(weights, str) and weights in self.zoo:
            weights = self.zoo[weights]
        elif isinstance(weights, (str, Path)):
            if not _is_weights_file(weights):
                raise ValueError(f"`weights` must be a valid path to a weight file, but got {weights}.")
            weights = Path(weights)
        elif isinstance(weights, dict):
            pass
        self.weights = weights or self.weights

    def _resolve_state_dict(self, weights: Any) -> dict | None:
        path, sd = None, None
        if isinstance(weights, dict):
            if "path" in weights:
                path = Path(weights["path"])
            else:
                sd = weights
        elif isinstance(weights, (str, Path)) and _is_weights_file(weights):
            path = Path(weights)
        if path is not None:
            if not _is_weights_file(path):
                raise ValueError(f"Cannot load weights from: {path}")
            if path.suffix.lower() == ".onnx":
                # globals.py:709 lists the suffix, but the reference's loader is `torch.load` too (nn/model.py:99-131)
                # and cannot read an ONNX graph either: fail with a clear message instead of an unpickling error
                raise ValueError(f"Cannot load weights from an ONNX file ({path}): export a .pt / .pth state_dict")
            dev = next(self.parameters()).device
            sd = torch.load(str(path), weights_only=True, map_location=dev)
        if sd is not None and "state_dict" in sd:
            sd = sd["state_dict"]
        return sd

    def load_weights(self, weights: Any = None, overwrite: bool = False):
        self.assign_weights(weights, overwrite)
        sd = self._resolve_state_dict(self.weights) if self.weights else None
        if sd:
            self.load_state_dict(sd)  # strict, like the reference

    # ---- contract checks (enhance/base.py:32-47)
    def assert_datapoint(self, datapoint: dict) -> None:
        if "image" not in datapoint:
            raise ValueError("The key ``'image'`` must be defined in the `datapoint`.")

    def assert_outputs(self, outputs: dict) -> None:
        if "enhanced" not in outputs:
            raise ValueError("The key ``'enhanced'`` must be defined in the `outputs`.")

    def forward_loss(self, datapoint: dict, *args, **kwargs) -> dict:
        raise NotImplementedError("the B200 drop-in covers the inference forward only (training is out of scope)")

    @property
    def device(self) -> torch.device:
        return next(self.parameters()).device

    def infer(self, datapoint: dict, image_size: Any = 512, resize: bool = False, *args, **kwargs) -> dict:
        """`VisionModel.infer` (vision/model.py:76-127), statement by statement: every image tensor of the caller's
        `datapoint` is replaced IN PLACE by its resized version -- to `image_size` (`resize=True`) or up to the next
        multiple of 32 (`core.resize(v, divisible_by=32)`, core/image/utils.py:381-382 rounds with ceil) -- and moved
        to the model's device (:101-109), `forward` runs between `Timer.tick()` / `tock()` (:112-115), every image in
        the outputs whose size differs from the input's is resized back to (h0, w0) (:119-123) and `"time"` is
        appended LAST (:126; predict.py:91 pops it by name, then `popitem()` yields "enhanced").  The resize is
        `torch.ops.b200dce.resize_bilinear` (F.interpolate bilinear, align_corners=False -- what kornia's resize
        calls); the tensor goes to the device first because the kernel lives there (the reference resizes on the
        host, same arithmetic).  "time" is the forward's wall time; the reference's `Timer` does not synchronise the
        device, this one does so that the number means something."""
        self.assert_datapoint(datapoint)
        h0, w0 = datapoint["image"].shape[-2:]
        if resize:
            th, tw = (image_size, image_size) if isinstance(image_size, int) else tuple(image_size)[-2:]
        else:
            th, tw = -(-h0 // 32) * 32, -(-w0 // 32) * 32
        for k, v in datapoint.items():
            if isinstance(v, torch.Tensor):
                v = v.to(self.device)
                if v.dim() == 4 and v.is_floating_point() and tuple(v.shape[-2:]) != (th, tw):
                    v = torch.ops.b200dce.resize_bilinear(v.float(), th, tw)
                datapoint[k] = v
        torch.cuda.synchronize(self.device)
        t0 = time.time()
        outputs = self.forward(datapoint, *args, **kwargs)
        torch.cuda.synchronize(self.device)
        elapsed = time.time() - t0
        self.assert_outputs(outputs)
        for k, v in outputs.items():
            if isinstance(v, torch.Tensor) and v.dim() == 4 and tuple(v.shape[-2:]) != (h0, w0):
                outputs[k] = torch.ops.b200dce.resize_bilinear(v.contiguous(), h0, w0)
        outputs["time"] = elapsed
        return outputs

    # ---- packed-weight cache: rebuilt when any parameter's storage / version / device changes
    def _params_key(self, tensors) -> tuple:
        # (a move to another device changes the pointer too: str(t.device) per tensor cost 10 us per call here)
        return tuple([t.data_ptr() for t in tensors] + [t._version for t in tensors])

    def _image(self, datapoint: dict) -> torch.Tensor:
        self.assert_datapoint(datapoint)
        x = datapoint.get("image")
        if not isinstance(x, torch.Tensor):
            raise TypeError(f"datapoint['image'] must be a torch.Tensor, got {type(x)}")
        if not x.is_cuda:
            raise RuntimeError(f"{type(self).__name__}: input is on {x.device}; the B200 drop-in has no CPU fallback "
                               "(the CPU restatement lives in oracle/ and is test infrastructure)")
        if x.device != self.device:
            raise RuntimeError(f"input on {x.device} but model on {self.device}")
        if torch.is_grad_enabled() and x.requires_grad:
            raise NotImplementedError("the B200 drop-in is inference-only; call under torch.no_grad()")
        return x


_E_CONVS = tuple(f"e_conv{i}" for i in range(1, 8))


def _init_conv_weights(module: nn.Module) -> None:
    """Reference random init: N(0, 0.02) on every conv weight, biases keep torch's default
    (zero_dce.py:152-158, zero_dcepp.py:157-167)."""
    for m in module.modules():
        if isinstance(m, nn.Conv2d):
            m.weight.data.normal_(0.0, 0.02)


# --------------------------------------------------------------------------------------------- Zero-DCE
@MODELS.register(name="zero_dce_re", arch="zero_dce")
class ZeroDCE_RE(_EnhancementModel):
    """Zero-DCE (zero_dce.py:86-191).  `precision`: "fp32" (default: the reference's fp32 contract, max-abs <= 1e-3,
    on the tcgen05 tensor cores as a 3-term split-bf16 product with fp32 accumulation -- alias "bf16x3"), "bf16"
    (tcgen05, PSNR >= 50 dB, 2.6x faster) or "fp32_exact" (every product in fp32 on the CUDA cores, ~1e-6, 10x slower)."""

    arch = "zero_dce"

    def __init__(self, in_channels: int = 3, num_channels: int = 32, num_iters: int = 8, weights: Any = None,
                 precision: str = "fp32", *args, **kwargs):
        kwargs.pop("name", None)
        super().__init__(name="zero_dce_re", in_channels=in_channels, weights=weights, *args, **kwargs)
        if isinstance(self.weights, dict):
            in_channels = self.weights.get("in_channels", in_channels)
            num_channels = self.weights.get("num_channels", num_channels)
            num_iters = self.weights.get("num_iters", num_iters)
        self.in_channels = in_channels or self.in_channels
        self.num_channels = num_channels
        self.num_iters = num_iters
        self.out_channels = self.in_channels * self.num_iters
        if precision not in PRECISIONS:
            raise ValueError(f"precision must be one of {list(PRECISIONS)}, got {precision!r}")
        self.precision = precision
        nc = self.num_channels
        self.relu = nn.ReLU(inplace=True)
        self.e_conv1 = nn.Conv2d(self.in_channels, nc, 3, 1, 1, bias=True)
        self.e_conv2 = nn.Conv2d(nc, nc, 3, 1, 1, bias=True)
        self.e_conv3 = nn.Conv2d(nc, nc, 3, 1, 1, bias=True)
        self.e_conv4 = nn.Conv2d(nc, nc, 3, 1, 1, bias=True)
        self.e_conv5 = nn.Conv2d(nc * 2, nc, 3, 1, 1, bias=True)
        self.e_conv6 = nn.Conv2d(nc * 2, nc, 3, 1, 1, bias=True)
        self.e_conv7 = nn.Conv2d(nc * 2, self.out_channels, 3, 1, 1, bias=True)
        self.maxpool = nn.MaxPool2d(2, 2, return_indices=False, ceil_mode=False)  # dead code in the reference too
        self.upsample = nn.UpsamplingBilinear2d(scale_factor=2)
        self.loss = ZeroDCELoss(reduction="mean")
        if self.weights:
            self.load_weights()
        else:
            _init_conv_weights(self)

    def _convs(self):
        return [getattr(self, f"e_conv{i}") for i in range(1, 8)]

    def packed_weights(self) -> torch.Tensor:
        prec = PRECISIONS[self.precision]
        ps = [self._modules[n]._parameters for n in _E_CONVS]   # dictionary reads instead of nn.Module.__getattr__
        ws, bs = [d["weight"] for d in ps], [d["bias"] for d in ps]
        key = (prec, self.num_iters) + self._params_key(ws + bs)
        if self._packed is None or self._packed_key != key:
            self._packed = torch.ops.b200dce.dce_pack([w.detach() for w in ws], [b.detach() for b in bs],
                                                      self.num_iters, prec)
            self._packed_key = key
        return self._packed

    @torch.no_grad()
    def forward(self, datapoint: dict, *args, **kwargs) -> dict:
        x = self._image(datapoint)
        adjust, enhanced = torch.ops.b200dce.dce_forward(x, self.packed_weights(), self.num_iters,
                                                         PRECISIONS[self.precision], True)
        outputs = {}
        for i, a_i in enumerate(torch.split(adjust, 3, dim=1)):  # views of `adjust`, like the reference
            outputs[f"adjust_{i}"] = a_i
        outputs["adjust"] = adjust
        outputs["enhanced"] = enhanced
        return outputs

    @torch.no_grad()
    def enhance(self, x: torch.Tensor) -> tuple:
        """Tuple API of the upstream model (mon_extra/.../zero_dce/model.py:27-50): returns
        (enhanced, adjust)."""
        out = self.forward({"image": x})
        return out["enhanced"], out["adjust"]


# --------------------------------------------------------------------------------------------- ZeroDCE-V (core)
@MODELS.register(name="zero_dce_v", arch="zero_dce")
class ZeroDCEV(_EnhancementModel):
    """ZeroDCE-V (zero_dce_v.py:92-230): constructor, attribute names and state_dict layout of the
    reference class.  `forward_core` = its forward up to `image_v_fixed_lr` (:172-193: V channel, nearest resize to
    `down_size`, DCE-Net with in_channels=1, `num_iters` curve steps on V) on the fp32 kernels; `forward` adds the
    `FastGuidedFilter` upsample (:194), the HSV recomposition (kornia's formulas) and the division by the frame maximum
    (:195-197) and returns the reference's six outputs."""

    arch = "zero_dce"
    schemes: list = ["unsupervised", "zero_reference", "instance"]

    def __init__(self, in_channels: int = 1, num_channels: int = 32, num_iters: int = 15, down_size: int = 256,
                 weights: Any = None, *args, **kwargs):
        kwargs.pop("name", None)
        super().__init__(name="zero_dce_v", in_channels=in_channels, weights=weights, *args, **kwargs)
        if isinstance(self.weights, dict):
            in_channels = self.weights.get("in_channels", in_channels)
            num_channels = self.weights.get("num_channels", num_channels)
            num_iters = self.weights.get("num_iters", num_iters)
            down_size = self.weights.get("down_size", down_size)
        self.in_channels = in_channels or self.in_channels
        self.num_channels = num_channels
        self.num_iters = num_iters
        self.out_channels = self.in_channels * self.num_iters
        self.down_size = down_size
        nc = self.num_channels
        self.relu = nn.ReLU(inplace=True)
        self.e_conv1 = nn.Conv2d(self.in_channels, nc, 3, 1, 1, bias=True)
        self.e_conv2 = nn.Conv2d(nc, nc, 3, 1, 1, bias=True)
        self.e_conv3 = nn.Conv2d(nc, nc, 3, 1, 1, bias=True)
        self.e_conv4 = nn.Conv2d(nc, nc, 3, 1, 1, bias=True)
        self.e_conv5 = nn.Conv2d(nc * 2, nc, 3, 1, 1, bias=True)
        self.e_conv6 = nn.Conv2d(nc * 2, nc, 3, 1, 1, bias=True)
        self.e_conv7 = nn.Conv2d(nc * 2, self.out_channels, 3, 1, 1, bias=True)
        self.loss = ZeroDCELoss()
        if self.weights:
            self.load_weights()
        else:
            _init_conv_weights(self)

    def packed_weights(self) -> torch.Tensor:
        ps = [self._modules[n]._parameters for n in _E_CONVS]   # dictionary reads instead of nn.Module.__getattr__
        ws, bs = [d["weight"] for d in ps], [d["bias"] for d in ps]
        key = (self.num_iters,) + self._params_key(ws + bs)
        if self._packed is None or self._packed_key != key:
            self._packed = torch.ops.b200dce.dcev_pack([w.detach() for w in ws], [b.detach() for b in bs],
                                                       self.num_iters)
            self._packed_key = key
        return self._packed

    @torch.no_grad()
    def forward_core(self, datapoint: dict, *args, **kwargs) -> dict:
        """The first four entries of the reference's output dict (zero_dce_v.py:198-203), same keys and order."""
        x = self._image(datapoint)
        v, v_lr, adjust, fixed = torch.ops.b200dce.dcev_core(x, self.packed_weights(), int(self.down_size),
                                                             self.num_iters)
        return {"image_v": v, "image_v_lr": v_lr, "adjust": adjust, "image_v_fixed_lr": fixed}

    @torch.no_grad()
    def forward(self, datapoint: dict, *args, **kwargs) -> dict:
        """All six entries of the reference's output dict (zero_dce_v.py:198-205), same keys and order.  The reference
        only runs with batch 1 (`replace_v_component`, :221-230); a batch here is a loop of batch-1 calls, i.e. `enhanced`
        is divided by each frame's own maximum.  `enhanced` goes through kornia's HSV conversions, restated from their
        published form (kornia is not installed where this was built: that output is not pinned to the reference)."""
        out = self.forward_core(datapoint)
        x = self._image(datapoint)
        v_fixed, enhanced = torch.ops.b200dce.dcev_tail(x, out["image_v"], out["image_v_lr"], out["image_v_fixed_lr"])
        out["image_v_fixed"] = v_fixed
        out["enhanced"] = enhanced
        return out


# --------------------------------------------------------------------------------------------- Zero-DCE++
@MODELS.register(name="zero_dce++_re", arch="zero_dce")
class ZeroDCEpp_RE(_EnhancementModel):
    """Zero-DCE++ (zero_dcepp.py:87-208).  For scale_factor != 1 the reference class raises at run time
    (zero_dcepp.py:139 passes the factor as `size`); this class implements the intended semantics of the
    upstream model (mon_extra/.../zero_dce++/model.py:39,64-82): bilinear down by 1/sf (align_corners=False),
    bilinear up by sf (align_corners=True).  `return_adjust=False` skips materialising the full-resolution
    curve map (non-default fast mode; "adjust" is then omitted from the outputs)."""

    arch = "zero_dce"

    def __init__(self, in_channels: int = 3, num_channels: int = 32, num_iters: int = 8, scale_factor: float = 1.0,
                 weights: Any = None, return_adjust: bool = True, *args, **kwargs):
        kwargs.pop("name", None)
        super().__init__(name="zero_dce++_re", in_channels=in_channels, weights=weights, *args, **kwargs)
        if isinstance(self.weights, dict):
            in_channels = self.weights.get("in_channels", in_channels)
            num_channels = self.weights.get("num_channels", num_channels)
            num_iters = self.weights.get("num_iters", num_iters)
            scale_factor = self.weights.get("scale_factor", scale_factor)
        self.in_channels = in_channels
        self.num_channels = num_channels
        self.num_iters = num_iters
        self.scale_factor = scale_factor
        self.return_adjust = return_adjust
        nc = self.num_channels
        self.relu = nn.ReLU(inplace=True)
        self.upsample = nn.UpsamplingBilinear2d(scale_factor=self.scale_factor)
        self.e_conv1 = DSConv2d(self.in_channels, nc, 3, 1, 1)
        self.e_conv2 = DSConv2d(nc, nc, 3, 1, 1)
        self.e_conv3 = DSConv2d(nc, nc, 3, 1, 1)
        self.e_conv4 = DSConv2d(nc, nc, 3, 1, 1)
        self.e_conv5 = DSConv2d(nc * 2, nc, 3, 1, 1)
        self.e_conv6 = DSConv2d(nc * 2, nc, 3, 1, 1)
        self.e_conv7 = DSConv2d(nc * 2, self.out_channels, 3, 1, 1)
        self.loss = ZeroDCELoss()
        if self.weights:
            self.load_weights()
        else:
            _init_conv_weights(self)

    def _layers(self):
        return [getattr(self, f"e_conv{i}") for i in range(1, 8)]

    def _int_scale(self) -> int:
        sf = float(self.scale_factor)
        if sf < 1 or sf != int(sf):
            raise NotImplementedError(f"the sm_100a Zero-DCE++ kernels implement integer scale_factor >= 1, got {sf}")
        return int(sf)

    def packed_weights(self) -> torch.Tensor:
        # straight dictionary reads: nn.Module.__getattr__ on 7 x 2 submodules x 2 parameters cost ~10 us per call
        mods = self._modules
        dw = [mods[n]._modules["dw_conv"]._parameters for n in _E_CONVS]
        pw = [mods[n]._modules["pw_conv"]._parameters for n in _E_CONVS]
        ts = [d["weight"] for d in dw] + [d["bias"] for d in dw] + [d["weight"] for d in pw] + [d["bias"] for d in pw]
        key = self._params_key(ts)
        if self._packed is None or self._packed_key != key:
            d = [t.detach() for t in ts]
            self._packed = torch.ops.b200dce.dcepp_pack(d[0:7], d[7:14], d[14:21], d[21:28])
            self._packed_key = key
        return self._packed

    @torch.no_grad()
    def forward(self, datapoint: dict, *args, **kwargs) -> dict:
        x = self._image(datapoint)
        adjust, enhanced = torch.ops.b200dce.dcepp_forward(x, self.packed_weights(), self._int_scale(),
                                                           self.num_iters, self.return_adjust)
        if self.return_adjust:
            return {"adjust": adjust, "enhanced": enhanced}
        return {"enhanced": enhanced}

    @torch.no_grad()
    def enhance(self, x: torch.Tensor) -> tuple:
        """Tuple API of the upstream model (mon_extra/.../zero_dce++/model.py:64-82): (enhanced, adjust)."""
        out = self.forward({"image": x})
        return out["enhanced"], out.get("adjust")


def remap_upstream_dcepp_keys(state_dict: dict) -> dict:
    """Upstream Zero-DCE++ checkpoints name the DS-conv halves `depth_conv` / `point_conv`
    (mon_extra/.../zero_dce++/model.py:10-25); `mon` names them `dw_conv` / `pw_conv`."""
    out = {}
    for k, v in state_dict.items():
        out[k.replace(".depth_conv.", ".dw_conv.").replace(".point_conv.", ".pw_conv.")] = v
    return out


# --------------------------------------------------------------------------------------------- SCI
class _SCIEnhanceNetwork(nn.Module):
    """Parameter container with upstream's names (sci/model.py:7-34, layers=1, channels=3); the conv
    block is also reachable as `blocks.0` (same module object), as upstream appends it to a ModuleList."""

    def __init__(self, layers: int = 1, channels: int = 3):
        super().__init__()
        if layers != 1 or channels != 3:
            raise NotImplementedError("the sm_100a SCI kernel implements layers=1, channels=3 (upstream Finetunemodel)")
        self.in_conv = nn.Sequential(nn.Conv2d(3, channels, 3, 1, 1), nn.ReLU())
        self.conv = nn.Sequential(nn.Conv2d(channels, channels, 3, 1, 1), nn.BatchNorm2d(channels), nn.ReLU())
        self.blocks = nn.ModuleList([self.conv])
        self.out_conv = nn.Sequential(nn.Conv2d(channels, 3, 3, 1, 1), nn.Sigmoid())


@MODELS.register(name="sci_re", arch="sci")
class SCI_RE(_EnhancementModel):
    """SCI inference model = upstream `Finetunemodel` (sci/model.py:133-159): state_dict keys
    `enhance.{in_conv.0,conv.0,conv.1,blocks.0.0,blocks.0.1,out_conv.0}.*`.  BatchNorm runs in eval mode
    (running statistics), which is what the upstream predict script uses (`model.eval()`)."""

    arch = "sci"

    def __init__(self, weights: Any = None, *args, **kwargs):
        kwargs.pop("name", None)
        super().__init__(name="sci_re", in_channels=3, weights=weights, *args, **kwargs)
        self.enhance = _SCIEnhanceNetwork(layers=1, channels=3)
        for m in self.modules():  # upstream weights_init: N(0,0.02) convs, zero bias; BN N(1,0.02)
            if isinstance(m, nn.Conv2d):
                m.weight.data.normal_(0, 0.02)
                m.bias.data.zero_()
            elif isinstance(m, nn.BatchNorm2d):
                m.weight.data.normal_(1.0, 0.02)
        self.eval()
        if self.weights:
            self.load_weights()

    def packed_weights(self) -> torch.Tensor:
        sd = self.enhance.state_dict()
        names = ["in_conv.0.weight", "in_conv.0.bias", "conv.0.weight", "conv.0.bias", "conv.1.weight", "conv.1.bias",
                 "conv.1.running_mean", "conv.1.running_var", "out_conv.0.weight", "out_conv.0.bias"]
        key = self._params_key([sd[n] for n in names])
        if self._packed is None or self._packed_key != key:
            self._packed = ops.sci_pack(sd, eps=self.enhance.conv[1].eps)
            self._packed_key = key
        return self._packed

    @torch.no_grad()
    def forward(self, datapoint: dict, *args, **kwargs) -> dict:
        if self.enhance.conv[1].training:
            raise NotImplementedError("SCI_RE implements eval-mode BatchNorm only; call .eval()")
        x = self._image(datapoint)
        illu, r = torch.ops.b200dce.sci_forward(x, self.packed_weights(), True)
        return {"illumination": illu, "enhanced": r}

    @torch.no_grad()
    def enhance_tuple(self, x: torch.Tensor) -> tuple:
        """Upstream tuple API: `Finetunemodel.forward` returns (i, r) (sci/model.py:155-159)."""
        out = self.forward({"image": x})
        return out["illumination"], out["enhanced"]


# --------------------------------------------------------------------------------------------- RUAS
_RUAS_IEM = ("skip_connect", "resconv_1x1", "resdilconv_3x3", "conv_3x3", "conv_3x3", "skip_connect", "conv_3x3")
_RUAS_NRM = ("resconv_1x1", "resconv_1x1", "resdilconv_3x3", "skip_connect", "resconv_1x1", "resconv_1x1",
             "skip_connect")


class _RuasOp(nn.Module):
    """Parameter holder named like upstream's ConvBlock / ResBlock (`.op`, ruas/operations.py:25-45)."""

    def __init__(self, channels: int, op: str):
        super().__init__()
        k, dil = (1 if op.endswith("1x1") else 3), (2 if "dil" in op else 1)
        self.op = nn.Conv2d(channels, channels, k, 1, padding=(k - 1) // 2 * dil, dilation=dil)


class _RuasSearchBlock(nn.Module):
    """Parameter container with upstream SearchBlock's names (ruas/model.py:24-44)."""

    def __init__(self, channels: int, genotype: tuple):
        super().__init__()
        for slot, op in zip(("c1_d", "c1_r", "c2_d", "c2_r", "c3_d", "c3_r", "c4"), genotype):
            setattr(self, slot, nn.Identity() if op == "skip_connect" else _RuasOp(channels, op))
        self.c5 = nn.Conv2d(4 * channels, channels, 1)


class _RuasIEM(nn.Module):
    def __init__(self):
        super().__init__()
        self.cell = _RuasSearchBlock(3, _RUAS_IEM)


class _RuasEnhanceNetwork(nn.Module):
    def __init__(self):
        super().__init__()
        self.iems = nn.ModuleList([_RuasIEM() for _ in range(3)])


class _RuasDenoiseNetwork(nn.Module):
    def __init__(self):
        super().__init__()
        self.stem = nn.Conv2d(3, 6, 3, 1, 1)
        self.nrms = nn.ModuleList([_RuasSearchBlock(6, _RUAS_NRM) for _ in range(3)])
        self.activate = nn.Sequential(nn.Conv2d(6, 3, 3, 1, 1))


@MODELS.register(name="ruas_re", arch="ruas")
class RUAS_RE(_EnhancementModel):
    """RUAS inference model = upstream `Network` (ruas/model.py:152-206): state_dict keys `enhance_net.iems.{i}.cell.*`
    and `denoise_net.{stem,nrms.{i}.*,activate.0}.*`, iem_nums = nrm_nums = 3, IEM / NRM genotypes.

    `denoise=True` returns the denoised result as `enhanced` (what upstream's predict saves for the `lol` weights,
    `u_list[-1]`); `denoise=False` skips the denoise net and returns `u_list[-2]` (the `upe` / `dark` weights)."""

    arch = "ruas"

    def __init__(self, denoise: bool = True, weights: Any = None, *args, **kwargs):
        kwargs.pop("name", None)
        super().__init__(name="ruas_re", in_channels=3, weights=weights, *args, **kwargs)
        self.denoise = bool(denoise)
        self.enhance_net = _RuasEnhanceNetwork()
        self.denoise_net = _RuasDenoiseNetwork()
        self.eval()
        if self.weights:
            self.load_weights()

    def packed_weights(self) -> torch.Tensor:
        key = self._params_key(list(self.parameters()))
        if self._packed is None or self._packed_key != key:
            self._packed = ops.ruas_pack(self.state_dict())
            self._packed_key = key
        return self._packed

    @torch.no_grad()
    def forward_lists(self, x: torch.Tensor) -> tuple:
        """Upstream tuple API: `Network.forward` returns (u_list, t_list) (ruas/model.py:202-206)."""
        u, t, den, _ = torch.ops.b200dce.ruas_forward(x, self.packed_weights(), self.denoise)
        u_list = list(u.unbind(0)) + ([den] if self.denoise else [])
        return u_list, list(t.unbind(0))

    @torch.no_grad()
    def forward(self, datapoint: dict, *args, **kwargs) -> dict:
        x = self._image(datapoint)
        u, t, den, noise = torch.ops.b200dce.ruas_forward(x, self.packed_weights(), self.denoise)
        out = {"illumination": t[2], "enhanced_no_denoise": u[2]}
        if self.denoise:
            out["noise"] = noise
            out["enhanced"] = den
        else:
            out["enhanced"] = out.pop("enhanced_no_denoise")
        return out
