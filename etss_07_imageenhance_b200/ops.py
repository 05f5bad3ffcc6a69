"""torch.library custom ops (`torch.ops.b200dce.*`) over the C ABI (include/b200dce.h).

Each op unpacks tensors to raw device pointers + sizes and calls the C function on
torch's current CUDA stream.  CUDA-only: a CPU tensor raises (there is no CPU
fallback by design — the CPU restatement lives in `oracle/` and is test
infrastructure).  Inference-only: no autograd formula is registered.

Replaces the ATen op sequence of
  src/mon/vision/enhance/llie/zero_dce/zero_dce.py:176-188      (dce_forward)
  src/mon/vision/enhance/llie/zero_dce/zero_dcepp.py:186-203    (dcepp_forward, curve_shared)
  src/mon/vision/enhance/llie/gcenet/gcenet.py:469-484          (gce_curve)
  src/mon_extra/vision/enhance/llie/sci/model.py:36-45,155-159  (sci_forward)
  src/mon_extra/vision/enhance/llie/ruas/model.py:46-149,202-206  (ruas_forward)
"""
from __future__ import annotations

from typing import List, Tuple

import torch
from torch import Tensor

from . import _lib
from ._lib import PRECISIONS, check, lib, ptr_array

__all__ = [
    "curve_shared", "curve_per_iter", "gce_curve", "guided_filter", "dcepp_pack", "dcepp_forward", "dcepp_lowres", "dcepp_apply",
    "dce_pack", "dce_forward", "dcev_pack", "dcev_core", "dcev_tail", "sci_pack", "sci_forward", "ruas_pack", "ruas_forward", "u8hwc_to_f32nchw", "f32nchw_to_u8hwc", "resize_bilinear",
    "current_stream_ptr", "require_device",
]


def current_stream_ptr(device=None) -> int:
    return torch.cuda.current_stream(device).cuda_stream


class _NoSwitch:
    def __enter__(self):
        return None

    def __exit__(self, *exc):
        return False


_NO_SWITCH = _NoSwitch()


def _on(device):
    """`torch.cuda.device(device)` only when the current device is another one (the context manager costs ~8 us per
    call, a third of the host time of a small forward)."""
    return _NO_SWITCH if torch.cuda.current_device() == device.index else torch.cuda.device(device)


def require_device() -> None:
    """Raise unless a B200 (sm_100) is the current CUDA device."""
    if not torch.cuda.is_available():
        raise RuntimeError("b200dce: CUDA is not available and there is no CPU fallback")
    check(lib().b200_check_device(), "b200_check_device")


def _img(t: Tensor, name: str, channels: int | None = 3) -> Tensor:
    if not isinstance(t, Tensor):
        raise TypeError(f"b200dce: `{name}` must be a torch.Tensor, got {type(t)}")
    if not t.is_cuda:
        raise RuntimeError(f"b200dce: `{name}` is on {t.device}; the sm_100a kernels have no CPU fallback")
    if t.dtype != torch.float32:
        raise TypeError(f"b200dce: `{name}` must be float32, got {t.dtype}")
    if t.dim() != 4:
        raise ValueError(f"b200dce: `{name}` must be [B,C,H,W], got shape {tuple(t.shape)}")
    if channels is not None and t.shape[1] != channels:
        raise ValueError(f"b200dce: `{name}` must have {channels} channels, got {t.shape[1]}")
    return t.contiguous()


def _no_grad(*ts: Tensor) -> None:
    if torch.is_grad_enabled() and any(t.requires_grad for t in ts):
        raise NotImplementedError("b200dce ops are inference-only (no autograd formula); "
                                  "call under torch.no_grad() / detach the inputs")


# ----------------------------------------------------------------------------- curve
@torch.library.custom_op("b200dce::curve_shared", mutates_args=(), device_types="cuda")
def curve_shared(x: Tensor, a: Tensor, iters: int) -> Tensor:
    x, a = _img(x, "x", None), _img(a, "a", None)
    if a.shape != x.shape:
        raise ValueError(f"b200dce::curve_shared: a {tuple(a.shape)} must match x {tuple(x.shape)}")
    y = torch.empty_like(x)
    B, C, H, W = x.shape
    if x.numel():
        with _on(x.device):
            check(lib().b200_curve_shared(x.data_ptr(), a.data_ptr(), y.data_ptr(), B, C, H, W, iters,
                                          current_stream_ptr()), "b200_curve_shared")
    return y


@curve_shared.register_fake
def _(x, a, iters):
    return torch.empty_like(x)


@torch.library.custom_op("b200dce::gce_curve", mutates_args=(), device_types="cuda")
def gce_curve(image: Tensor, adjust: Tensor, iters: int, gamma: float, debug: bool) -> Tuple[Tensor, Tensor, Tensor, Tensor]:
    """GCENet's enhancement loop (gcenet.py:469-484): (enhanced, bam, bright, dark).  `enhanced` is the loop's result
    before the guided filter; bam [B,1,H,W] = (1 - max(R,G,B))^gamma (prior/image.py:269-272) and the last iteration's
    bright / dark [B,3,H,W] are returned when `debug` (the reference's debug dict) and gamma != 0, else empty [0]
    tensors (the reference returns None for them when `bam_gamma in [None, 0.0]`)."""
    image, adjust = _img(image, "image"), _img(adjust, "adjust")
    if adjust.shape != image.shape:
        raise ValueError(f"b200dce::gce_curve: adjust {tuple(adjust.shape)} must match image {tuple(image.shape)}")
    B, _, H, W = image.shape
    enhanced = torch.empty_like(image)
    extra = debug and gamma != 0.0
    bam = image.new_empty(B, 1, H, W) if extra else image.new_empty(0)
    bright = torch.empty_like(image) if extra else image.new_empty(0)
    dark = torch.empty_like(image) if extra else image.new_empty(0)
    if image.numel():
        with _on(image.device):
            check(lib().b200_gce_curve(image.data_ptr(), adjust.data_ptr(), enhanced.data_ptr(),
                                       bam.data_ptr() if extra else None, bright.data_ptr() if extra else None,
                                       dark.data_ptr() if extra else None, B, H, W, iters, float(gamma),
                                       current_stream_ptr()), "b200_gce_curve")
    return enhanced, bam, bright, dark


@gce_curve.register_fake
def _(image, adjust, iters, gamma, debug):
    B, _, H, W = image.shape
    extra = debug and gamma != 0.0
    e = image.new_empty(0)
    return (torch.empty_like(image), image.new_empty(B, 1, H, W) if extra else e,
            torch.empty_like(image) if extra else e, torch.empty_like(image) if extra else e)


@torch.library.custom_op("b200dce::guided_filter", mutates_args=(), device_types="cuda")
def guided_filter(image: Tensor, guide: Tensor, radius: int, eps: float) -> Tensor:
    """filtering.GuidedFilter(radius, eps)(image, guide) (guided_filter.py:99-146), the step GCENet applies after its
    curve loop (gcenet.py:391,485).  `radius` is the CLASS's argument: it reaches BoxFilter as kernel size, so the window
    radius is int((radius - 1) / 2) (1 for GCENet's radius 3)."""
    image, guide = _img(image, "image", None), _img(guide, "guide", None)
    B, Cx, H, W = image.shape
    C = guide.shape[1]
    if guide.shape[0] != B or tuple(guide.shape[2:]) != (H, W) or Cx not in (1, C):
        raise ValueError(f"b200dce::guided_filter: image {tuple(image.shape)} / guide {tuple(guide.shape)}")
    if not (H > 2 * radius + 1 and W > 2 * radius + 1):   # the reference's assert (guided_filter.py:126)
        raise ValueError("b200dce::guided_filter: H and W must exceed 2 * radius + 1")
    out = torch.empty_like(guide)
    if out.numel():
        L = lib()
        ws_bytes = L.b200_guided_filter_workspace_bytes(B, C, H, W)
        ws = torch.empty(ws_bytes, dtype=torch.uint8, device=image.device)
        with _on(image.device):
            check(L.b200_guided_filter(image.data_ptr(), guide.data_ptr(), out.data_ptr(), B, Cx, C, H, W,
                                       int((radius - 1) / 2), float(eps), ws.data_ptr(), ws_bytes, current_stream_ptr()),
                  "b200_guided_filter")
    return out


@guided_filter.register_fake
def _(image, guide, radius, eps):
    return torch.empty_like(guide)


@torch.library.custom_op("b200dce::curve_per_iter", mutates_args=(), device_types="cuda")
def curve_per_iter(x: Tensor, a: Tensor, iters: int) -> Tensor:
    x, a = _img(x, "x", None), _img(a, "a", None)
    B, C, H, W = x.shape
    if tuple(a.shape) != (B, C * iters, H, W):
        raise ValueError(f"b200dce::curve_per_iter: a must be {(B, C * iters, H, W)}, got {tuple(a.shape)}")
    y = torch.empty_like(x)
    if x.numel():
        with _on(x.device):
            check(lib().b200_curve_per_iter(x.data_ptr(), a.data_ptr(), y.data_ptr(), B, C, H, W, iters,
                                            current_stream_ptr()), "b200_curve_per_iter")
    return y


@curve_per_iter.register_fake
def _(x, a, iters):
    return torch.empty_like(x)


# ----------------------------------------------------------------------------- Zero-DCE++
def _f32c(ts: List[Tensor], dev) -> List[Tensor]:
    return [t.detach().to(device=dev, dtype=torch.float32).contiguous() for t in ts]


@torch.library.custom_op("b200dce::dcepp_pack", mutates_args=(), device_types="cuda")
def dcepp_pack(dw_w: List[Tensor], dw_b: List[Tensor], pw_w: List[Tensor], pw_b: List[Tensor]) -> Tensor:
    """Re-pack e_conv{1..7}.{dw_conv,pw_conv}.{weight,bias} (dsconv.py:122-149 layouts) for the kernels."""
    if not (len(dw_w) == len(dw_b) == len(pw_w) == len(pw_b) == 7):
        raise ValueError("b200dce::dcepp_pack expects 7 layers")
    dev = dw_w[0].device
    cin = [3, 32, 32, 32, 64, 64, 64]
    cout = [32, 32, 32, 32, 32, 32, 3]
    for i in range(7):
        exp = [(cin[i], 1, 3, 3), (cin[i],), (cout[i], cin[i], 1, 1), (cout[i],)]
        got = [tuple(dw_w[i].shape), tuple(dw_b[i].shape), tuple(pw_w[i].shape), tuple(pw_b[i].shape)]
        if exp != got:
            raise NotImplementedError(
                f"b200dce: the sm_100a Zero-DCE++ kernels implement in_channels=3, num_channels=32 only; "
                f"layer {i + 1} has shapes {got}, expected {exp}")
    dw_w, dw_b, pw_w, pw_b = _f32c(dw_w, dev), _f32c(dw_b, dev), _f32c(pw_w, dev), _f32c(pw_b, dev)
    packed = torch.empty(lib().b200_dcepp_packed_floats(), dtype=torch.float32, device=dev)
    with _on(dev):
        check(lib().b200_dcepp_pack_weights(ptr_array([t.data_ptr() for t in dw_w]),
                                            ptr_array([t.data_ptr() for t in dw_b]),
                                            ptr_array([t.data_ptr() for t in pw_w]),
                                            ptr_array([t.data_ptr() for t in pw_b]),
                                            packed.data_ptr(), current_stream_ptr()), "b200_dcepp_pack_weights")
    return packed


@dcepp_pack.register_fake
def _(dw_w, dw_b, pw_w, pw_b):
    return dw_w[0].new_empty(lib().b200_dcepp_packed_floats(), dtype=torch.float32)  # host-only query


def _check_sf(H: int, W: int, sf: int) -> None:
    if sf < 1:
        raise ValueError(f"b200dce: scale_factor must be a positive integer, got {sf}")
    if H % sf or W % sf:
        # the upstream model fails the same way: the upsampled map no longer broadcasts against x
        raise RuntimeError(f"b200dce::dcepp_forward: H={H}, W={W} must be divisible by scale_factor={sf} "
                           f"(the reference raises a size-mismatch error here)")


@torch.library.custom_op("b200dce::dcepp_forward", mutates_args=(), device_types="cuda")
def dcepp_forward(x: Tensor, packed: Tensor, scale_factor: int, iters: int,
                  return_adjust: bool) -> Tuple[Tensor, Tensor]:
    """x [B,3,H,W] -> (adjust [B,3,H,W] (empty [0] if not requested), enhanced [B,3,H,W])."""
    x = _img(x, "x")
    B, _, H, W = x.shape
    _check_sf(H, W, scale_factor)
    enhanced = torch.empty_like(x)
    adjust = torch.empty_like(x) if return_adjust else x.new_empty(0)
    if x.numel() == 0:
        return adjust, enhanced
    L = lib()
    ws_bytes = L.b200_dcepp_workspace_bytes(B, H, W, scale_factor)
    ws = torch.empty(ws_bytes, dtype=torch.uint8, device=x.device)
    with _on(x.device):
        check(L.b200_dcepp_forward(x.data_ptr(), packed.data_ptr(), enhanced.data_ptr(),
                                   adjust.data_ptr() if return_adjust else None, B, H, W, scale_factor, iters,
                                   ws.data_ptr(), ws_bytes, current_stream_ptr()), "b200_dcepp_forward")
    return adjust, enhanced


@dcepp_forward.register_fake
def _(x, packed, scale_factor, iters, return_adjust):
    return (torch.empty_like(x) if return_adjust else x.new_empty(0)), torch.empty_like(x)


@torch.library.custom_op("b200dce::dcepp_lowres", mutates_args=(), device_types="cuda")
def dcepp_lowres(x: Tensor, packed: Tensor, scale_factor: int) -> Tensor:
    """Stages a3+a4 only: the low-resolution curve map [B,3,H/sf,W/sf]."""
    x = _img(x, "x")
    B, _, H, W = x.shape
    _check_sf(H, W, scale_factor)
    a = x.new_empty(B, 3, H // scale_factor, W // scale_factor)
    L = lib()
    ws_bytes = L.b200_dcepp_workspace_bytes(B, H, W, scale_factor)
    ws = torch.empty(ws_bytes, dtype=torch.uint8, device=x.device)
    with _on(x.device):
        check(L.b200_dcepp_lowres(x.data_ptr(), packed.data_ptr(), a.data_ptr(), B, H, W, scale_factor,
                                  ws.data_ptr(), ws_bytes, current_stream_ptr()), "b200_dcepp_lowres")
    return a


@dcepp_lowres.register_fake
def _(x, packed, scale_factor):
    B, _, H, W = x.shape
    return x.new_empty(B, 3, H // scale_factor, W // scale_factor)


@torch.library.custom_op("b200dce::dcepp_apply", mutates_args=(), device_types="cuda")
def dcepp_apply(x: Tensor, a_lr: Tensor, iters: int, return_adjust: bool) -> Tuple[Tensor, Tensor]:
    """Stages a5+a6 only: bilinear(align_corners=True) upsample of a_lr to x's size + curve."""
    x, a_lr = _img(x, "x"), _img(a_lr, "a_lr")
    B, _, H, W = x.shape
    if a_lr.shape[0] != B:
        raise ValueError("b200dce::dcepp_apply: batch mismatch")
    enhanced = torch.empty_like(x)
    adjust = torch.empty_like(x) if return_adjust else x.new_empty(0)
    with _on(x.device):
        check(lib().b200_dcepp_apply(x.data_ptr(), a_lr.data_ptr(), enhanced.data_ptr(),
                                     adjust.data_ptr() if return_adjust else None, B, H, W, a_lr.shape[2],
                                     a_lr.shape[3], iters, current_stream_ptr()), "b200_dcepp_apply")
    return adjust, enhanced


@dcepp_apply.register_fake
def _(x, a_lr, iters, return_adjust):
    return (torch.empty_like(x) if return_adjust else x.new_empty(0)), torch.empty_like(x)


# ----------------------------------------------------------------------------- Zero-DCE
@torch.library.custom_op("b200dce::dce_pack", mutates_args=(), device_types="cuda")
def dce_pack(weights: List[Tensor], biases: List[Tensor], iters: int, precision: int) -> Tensor:
    """Re-pack e_conv{1..7}.{weight,bias} (OIHW fp32, zero_dce.py:133-139) for the selected precision."""
    if not (len(weights) == len(biases) == 7):
        raise ValueError("b200dce::dce_pack expects 7 layers")
    cin = [3, 32, 32, 32, 64, 64, 64]
    cout = [32] * 6 + [3 * iters]
    for i in range(7):
        if tuple(weights[i].shape) != (cout[i], cin[i], 3, 3) or tuple(biases[i].shape) != (cout[i],):
            raise NotImplementedError(
                f"b200dce: the sm_100a Zero-DCE kernels implement in_channels=3, num_channels=32, 1<=num_iters<=8 "
                f"only; layer {i + 1} weight shape {tuple(weights[i].shape)}")
    dev = weights[0].device
    w, b = _f32c(weights, dev), _f32c(biases, dev)
    L = lib()
    nbytes = L.b200_dce_packed_bytes(precision, iters)
    if nbytes == 0:
        raise NotImplementedError(f"b200dce::dce_pack: precision {precision} / iters {iters} not supported")
    packed = torch.empty(nbytes, dtype=torch.uint8, device=dev)
    with _on(dev):
        check(L.b200_dce_pack_weights(ptr_array([t.data_ptr() for t in w]), ptr_array([t.data_ptr() for t in b]),
                                      packed.data_ptr(), precision, iters, current_stream_ptr()),
              "b200_dce_pack_weights")
    return packed


@dce_pack.register_fake
def _(weights, biases, iters, precision):
    return weights[0].new_empty(lib().b200_dce_packed_bytes(precision, iters), dtype=torch.uint8)  # host-only query


@torch.library.custom_op("b200dce::dce_forward", mutates_args=(), device_types="cuda")
def dce_forward(x: Tensor, packed: Tensor, iters: int, precision: int,
                return_adjust: bool) -> Tuple[Tensor, Tensor]:
    """x [B,3,H,W] -> (adjust [B,3*iters,H,W] (empty [0] if not requested), enhanced [B,3,H,W])."""
    x = _img(x, "x")
    B, _, H, W = x.shape
    enhanced = torch.empty_like(x)
    adjust = x.new_empty(B, 3 * iters, H, W) if return_adjust else x.new_empty(0)
    if x.numel() == 0:
        return adjust, enhanced
    L = lib()
    ws_bytes = max(L.b200_dce_workspace_bytes(B, H, W, precision), 256)
    ws = torch.empty(ws_bytes, dtype=torch.uint8, device=x.device)
    with _on(x.device):
        check(L.b200_dce_forward(x.data_ptr(), packed.data_ptr(), adjust.data_ptr() if return_adjust else None,
                                 enhanced.data_ptr(), B, H, W, iters, precision, ws.data_ptr(), ws_bytes,
                                 current_stream_ptr()), "b200_dce_forward")
    return adjust, enhanced


@dce_forward.register_fake
def _(x, packed, iters, precision, return_adjust):
    B, _, H, W = x.shape
    return (x.new_empty(B, 3 * iters, H, W) if return_adjust else x.new_empty(0)), torch.empty_like(x)


# ----------------------------------------------------------------------------- ZeroDCE-V core (partial)
@torch.library.custom_op("b200dce::dcev_pack", mutates_args=(), device_types="cuda")
def dcev_pack(weights: List[Tensor], biases: List[Tensor], iters: int) -> Tensor:
    """Re-pack ZeroDCEV's e_conv{1..7}.{weight,bias} (zero_dce_v.py:127-133: in_channels=1, num_channels=32,
    out = num_iters maps) for the fp32 kernels."""
    if not (len(weights) == len(biases) == 7):
        raise ValueError("b200dce::dcev_pack expects 7 layers")
    cin = [1, 32, 32, 32, 64, 64, 64]
    cout = [32] * 6 + [iters]
    for i in range(7):
        if tuple(weights[i].shape) != (cout[i], cin[i], 3, 3) or tuple(biases[i].shape) != (cout[i],):
            raise NotImplementedError(
                f"b200dce: the ZeroDCE-V kernels implement in_channels=1, num_channels=32, num_iters<=32 only; "
                f"layer {i + 1} weight shape {tuple(weights[i].shape)}")
    dev = weights[0].device
    w, b = _f32c(weights, dev), _f32c(biases, dev)
    packed = torch.empty(lib().b200_dcev_packed_floats(), dtype=torch.float32, device=dev)
    with _on(dev):
        check(lib().b200_dcev_pack_weights(ptr_array([t.data_ptr() for t in w]), ptr_array([t.data_ptr() for t in b]),
                                           packed.data_ptr(), iters, current_stream_ptr()), "b200_dcev_pack_weights")
    return packed


@dcev_pack.register_fake
def _(weights, biases, iters):
    return weights[0].new_empty(lib().b200_dcev_packed_floats(), dtype=torch.float32)


@torch.library.custom_op("b200dce::dcev_core", mutates_args=(), device_types="cuda")
def dcev_core(image: Tensor, packed: Tensor, down_size: int, iters: int) -> Tuple[Tensor, Tensor, Tensor, Tensor]:
    """image [B,3,H,W] -> (image_v [B,1,H,W], image_v_lr [B,1,S,S], adjust [B,iters,S,S], image_v_fixed_lr [B,1,S,S]):
    ZeroDCEV.forward up to `image_v_fixed_lr` (zero_dce_v.py:172-193)."""
    image = _img(image, "image")
    B, _, H, W = image.shape
    S = int(down_size)
    if S <= 0:
        raise ValueError(f"b200dce::dcev_core: bad down_size {down_size}")
    v = image.new_empty(B, 1, H, W)
    v_lr = image.new_empty(B, 1, S, S)
    adjust = image.new_empty(B, iters, S, S)
    fixed = image.new_empty(B, 1, S, S)
    if B:
        L = lib()
        ws_bytes = L.b200_dcev_workspace_bytes(B, S)
        ws = torch.empty(ws_bytes, dtype=torch.uint8, device=image.device)
        with _on(image.device):
            check(L.b200_dcev_core(image.data_ptr(), packed.data_ptr(), v.data_ptr(), v_lr.data_ptr(), adjust.data_ptr(),
                                   fixed.data_ptr(), B, H, W, S, iters, ws.data_ptr(), ws_bytes, current_stream_ptr()),
                  "b200_dcev_core")
    return v, v_lr, adjust, fixed


@dcev_core.register_fake
def _(image, packed, down_size, iters):
    B, _, H, W = image.shape
    S = int(down_size)
    return (image.new_empty(B, 1, H, W), image.new_empty(B, 1, S, S), image.new_empty(B, iters, S, S),
            image.new_empty(B, 1, S, S))


@torch.library.custom_op("b200dce::dcev_tail", mutates_args=(), device_types="cuda")
def dcev_tail(image: Tensor, image_v: Tensor, image_v_lr: Tensor, image_v_fixed_lr: Tensor) -> Tuple[Tensor, Tensor]:
    """(image_v_fixed [B,1,H,W], enhanced [B,3,H,W]): the tail of ZeroDCEV.forward (zero_dce_v.py:193-197):
    FastGuidedFilter(radius=1) upsample of the corrected low-resolution V (guided_filter.py:171-203), clip, V replaced
    in kornia's HSV, back to RGB, division by the frame's maximum (each frame = one batch-1 call of the reference)."""
    image = _img(image, "image")
    image_v, image_v_lr, image_v_fixed_lr = (_img(t, n, 1) for t, n in ((image_v, "image_v"), (image_v_lr, "image_v_lr"),
                                                                        (image_v_fixed_lr, "image_v_fixed_lr")))
    B, _, H, W = image.shape
    S = image_v_lr.shape[-1]
    if tuple(image_v.shape) != (B, 1, H, W) or tuple(image_v_lr.shape) != (B, 1, S, S) or image_v_fixed_lr.shape != image_v_lr.shape:
        raise ValueError("b200dce::dcev_tail: shapes must be [B,3,H,W], [B,1,H,W], [B,1,S,S], [B,1,S,S]")
    v_fixed = torch.empty_like(image_v)
    enhanced = torch.empty_like(image)
    if B:
        L = lib()
        ws_bytes = L.b200_dcev_tail_workspace_bytes(B, S)
        ws = torch.empty(ws_bytes, dtype=torch.uint8, device=image.device)
        with _on(image.device):
            check(L.b200_dcev_tail(image.data_ptr(), image_v.data_ptr(), image_v_lr.data_ptr(), image_v_fixed_lr.data_ptr(),
                                   v_fixed.data_ptr(), enhanced.data_ptr(), B, H, W, S, ws.data_ptr(), ws_bytes,
                                   current_stream_ptr()), "b200_dcev_tail")
    return v_fixed, enhanced


@dcev_tail.register_fake
def _(image, image_v, image_v_lr, image_v_fixed_lr):
    return torch.empty_like(image_v), torch.empty_like(image)


# ----------------------------------------------------------------------------- SCI
def sci_pack(sd: dict, eps: float = 1e-5, prefix: str = "") -> Tensor:
    """Fold EnhanceNetwork(layers=1, channels=3) weights (sci/model.py:16-34; BatchNorm eval) into the
    252-float HOST block that `sci_forward` passes to the kernel by value."""
    keys = ["in_conv.0.weight", "in_conv.0.bias", "conv.0.weight", "conv.0.bias", "conv.1.weight", "conv.1.bias",
            "conv.1.running_mean", "conv.1.running_var", "out_conv.0.weight", "out_conv.0.bias"]
    ts = [sd[prefix + k].detach().to("cpu", torch.float32).contiguous() for k in keys]
    for t, k in zip(ts, keys):
        exp = (3, 3, 3, 3) if k.endswith("0.weight") else (3,)
        if tuple(t.shape) != exp:
            raise NotImplementedError(f"b200dce::sci_pack: {k} has shape {tuple(t.shape)}, expected {exp} "
                                      "(the kernel implements layers=1, channels=3)")
    packed = torch.empty(_lib.SCI_PACKED_FLOATS, dtype=torch.float32)
    p = [t.data_ptr() for t in ts]
    check(lib().b200_sci_pack_weights_host(p[0], p[1], p[2], p[3], p[4], p[5], p[6], p[7], float(eps), p[8], p[9],
                                           packed.data_ptr()), "b200_sci_pack_weights_host")
    return packed


@torch.library.custom_op("b200dce::sci_forward", mutates_args=(), device_types="cuda")
def sci_forward(x: Tensor, packed_host: Tensor, return_illu: bool) -> Tuple[Tensor, Tensor]:
    """x [B,3,H,W] -> (illu (empty [0] if not requested), r)."""
    x = _img(x, "x")
    if packed_host.is_cuda or packed_host.numel() != _lib.SCI_PACKED_FLOATS or packed_host.dtype != torch.float32:
        raise ValueError("b200dce::sci_forward: packed_host must be the 252-float CPU tensor from sci_pack()")
    B, _, H, W = x.shape
    r = torch.empty_like(x)
    illu = torch.empty_like(x) if return_illu else x.new_empty(0)
    if x.numel():
        with _on(x.device):
            check(lib().b200_sci_forward(x.data_ptr(), packed_host.data_ptr(),
                                         illu.data_ptr() if return_illu else None, r.data_ptr(), B, H, W,
                                         current_stream_ptr()), "b200_sci_forward")
    return illu, r


@sci_forward.register_fake
def _(x, packed_host, return_illu):
    return (torch.empty_like(x) if return_illu else x.new_empty(0)), torch.empty_like(x)


# ----------------------------------------------------------------------------- RUAS
_RUAS_IEM_SLOTS = ("c1_r.op", "c2_d.op", "c2_r.op", "c3_d.op", "c4.op", "c5")          # genotypes.py:15 (skips own
_RUAS_NRM_SLOTS = ("c1_d.op", "c1_r.op", "c2_d.op", "c3_d.op", "c3_r.op", "c5")        # :16       no parameters)
_RUAS_IEM_SHAPES = ((3, 3, 1, 1), (3, 3, 3, 3), (3, 3, 3, 3), (3, 3, 3, 3), (3, 3, 3, 3), (3, 12, 1, 1))
_RUAS_NRM_SHAPES = ((6, 6, 1, 1), (6, 6, 1, 1), (6, 6, 3, 3), (6, 6, 1, 1), (6, 6, 1, 1), (6, 24, 1, 1))


def ruas_pack(sd: dict, prefix: str = "") -> Tensor:
    """Concatenate upstream `Network` weights (ruas/model.py:152-200 key names) into the 3438-float HOST block of
    include/b200dce.h (order: 3 IEM cells, stem, 3 NRM cells, activate.0; weight then bias of every conv)."""
    parts = []

    def take(key, shape):
        w = sd[prefix + key + ".weight"].detach().to("cpu", torch.float32).contiguous()
        b = sd[prefix + key + ".bias"].detach().to("cpu", torch.float32).contiguous()
        if tuple(w.shape) != shape or tuple(b.shape) != shape[:1]:
            raise NotImplementedError(f"b200dce::ruas_pack: {key} has shape {tuple(w.shape)}, expected {shape} "
                                      "(the kernels implement the IEM / NRM genotypes with 3 / 6 channels)")
        parts.extend((w.reshape(-1), b))

    for i in range(3):
        for slot, shape in zip(_RUAS_IEM_SLOTS, _RUAS_IEM_SHAPES):
            take(f"enhance_net.iems.{i}.cell.{slot}", shape)
    take("denoise_net.stem", (6, 3, 3, 3))
    for i in range(3):
        for slot, shape in zip(_RUAS_NRM_SLOTS, _RUAS_NRM_SHAPES):
            take(f"denoise_net.nrms.{i}.{slot}", shape)
    take("denoise_net.activate.0", (3, 6, 3, 3))
    packed = torch.cat(parts).contiguous()
    assert packed.numel() == _lib.RUAS_PACKED_FLOATS
    return packed


@torch.library.custom_op("b200dce::ruas_forward", mutates_args=(), device_types="cuda")
def ruas_forward(x: Tensor, packed_host: Tensor, denoise: bool) -> Tuple[Tensor, Tensor, Tensor, Tensor]:
    """x [B,3,H,W] -> (u [3,B,3,H,W], t [3,B,3,H,W], denoised [B,3,H,W], noise [B,3,H,W]); the last two are empty
    [0] tensors without `denoise`."""
    x = _img(x, "x")
    if packed_host.is_cuda or packed_host.numel() != _lib.RUAS_PACKED_FLOATS or packed_host.dtype != torch.float32:
        raise ValueError("b200dce::ruas_forward: packed_host must be the 3438-float CPU tensor from ruas_pack()")
    packed_host = packed_host.contiguous()
    B, _, H, W = x.shape
    u = x.new_empty(3, B, 3, H, W)
    t = x.new_empty(3, B, 3, H, W)
    den = torch.empty_like(x) if denoise else x.new_empty(0)
    noise = torch.empty_like(x) if denoise else x.new_empty(0)
    if x.numel():
        with _on(x.device):
            nbytes = lib().b200_ruas_workspace_bytes(B, H, W, int(denoise))
            ws = torch.empty(max(nbytes, 16), dtype=torch.uint8, device=x.device)
            check(lib().b200_ruas_forward(x.data_ptr(), packed_host.data_ptr(), u.data_ptr(), t.data_ptr(),
                                          den.data_ptr() if denoise else None, noise.data_ptr() if denoise else None,
                                          B, H, W, ws.data_ptr(), nbytes, current_stream_ptr()), "b200_ruas_forward")
    return u, t, den, noise


@ruas_forward.register_fake
def _(x, packed_host, denoise):
    B, _, H, W = x.shape
    e = torch.empty_like(x) if denoise else x.new_empty(0)
    return x.new_empty(3, B, 3, H, W), x.new_empty(3, B, 3, H, W), e, torch.empty_like(e)


# ----------------------------------------------------------------------------- frame format
@torch.library.custom_op("b200dce::u8hwc_to_f32nchw", mutates_args=(), device_types="cuda")
def u8hwc_to_f32nchw(frames: Tensor) -> Tensor:
    """[B,H,W,3] uint8 -> [B,3,H,W] float32 in [0,1] (to_image_tensor(normalize=True), core/image/utils.py:801-848)."""
    if not frames.is_cuda or frames.dtype != torch.uint8 or frames.dim() != 4 or frames.shape[3] != 3:
        raise ValueError("b200dce::u8hwc_to_f32nchw expects a CUDA uint8 [B,H,W,3] tensor")
    frames = frames.contiguous()
    B, H, W, _ = frames.shape
    out = torch.empty(B, 3, H, W, dtype=torch.float32, device=frames.device)
    with _on(frames.device):
        check(lib().b200_u8hwc_to_f32nchw(frames.data_ptr(), out.data_ptr(), B, H, W, current_stream_ptr()),
              "b200_u8hwc_to_f32nchw")
    return out


@u8hwc_to_f32nchw.register_fake
def _(frames):
    B, H, W, _ = frames.shape
    return frames.new_empty(B, 3, H, W, dtype=torch.float32)


@torch.library.custom_op("b200dce::f32nchw_to_u8hwc", mutates_args=(), device_types="cuda")
def f32nchw_to_u8hwc(image: Tensor) -> Tensor:
    """[B,3,H,W] float32 -> [B,H,W,3] uint8 (to_image_nparray(denormalize=True), core/image/utils.py:766-800)."""
    image = _img(image, "image")
    B, _, H, W = image.shape
    out = torch.empty(B, H, W, 3, dtype=torch.uint8, device=image.device)
    with _on(image.device):
        check(lib().b200_f32nchw_to_u8hwc(image.data_ptr(), out.data_ptr(), B, H, W, current_stream_ptr()),
              "b200_f32nchw_to_u8hwc")
    return out


@f32nchw_to_u8hwc.register_fake
def _(image):
    B, _, H, W = image.shape
    return image.new_empty(B, H, W, 3, dtype=torch.uint8)


# ----------------------------------------------------------------------------- infer()'s resize
@torch.library.custom_op("b200dce::resize_bilinear", mutates_args=(), device_types="cuda")
def resize_bilinear(image: Tensor, height: int, width: int) -> Tensor:
    """[B,C,H,W] float32 -> [B,C,height,width]: core.resize (core/image/geometry/transform.py:104-235) ->
    F.interpolate(size=..., mode="bilinear", align_corners=False, antialias=False)."""
    image = _img(image, "image", channels=None)
    if height <= 0 or width <= 0:
        raise ValueError(f"b200dce::resize_bilinear: bad target size {(height, width)}")
    B, C, H, W = image.shape
    out = torch.empty(B, C, height, width, dtype=torch.float32, device=image.device)
    if B * C:
        with _on(image.device):
            check(lib().b200_resize_bilinear(image.data_ptr(), out.data_ptr(), B * C, H, W, height, width,
                                             current_stream_ptr()), "b200_resize_bilinear")
    return out


@resize_bilinear.register_fake
def _(image, height, width):
    return image.new_empty(image.shape[0], image.shape[1], height, width)
