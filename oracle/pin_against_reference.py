#!/usr/bin/env python
"""Pin the oracle against the reference's OWN code and write tests/golden/*.npz.

TEST INFRASTRUCTURE.  Runs only in the build container (needs /root/reference).

What is executed from the reference, unmodified, read from where it lies:
  * src/mon/vision/enhance/llie/zero_dce/zero_dce.py     (class ZeroDCE_RE)
  * src/mon/vision/enhance/llie/zero_dce/zero_dcepp.py   (class ZeroDCEpp_RE)
  * src/mon/nn/modules/conv/dsconv.py                    (DSConv2d -> state_dict keys)
  * src/mon_extra/vision/enhance/llie/zero_dce/model.py      (upstream Zero-DCE)
  * src/mon_extra/vision/enhance/llie/zero_dce++/model.py    (upstream Zero-DCE++, sf != 1)
  * src/mon_extra/vision/enhance/llie/sci/model.py           (upstream SCI EnhanceNetwork)
  * src/mon_extra/vision/enhance/llie/ruas/model.py          (upstream RUAS EnhanceNetwork / DenoiseNetwork, with
    operations.py and genotypes.py; `Network.__init__` itself loads a zoo checkpoint that is not in the tree, so the
    two sub-networks are built the way Network.__init__ :168-177 builds them and chained as Network.forward :202-206)
`import mon` itself is impossible here (plum/humps/lightning/fvcore/... are not
installed, no network), so the three `mon` files are exec'd under a STUB `mon`
package that provides only the base-class plumbing they subclass
(Model ctor: src/mon/nn/model.py:237-277; assert_datapoint:
src/mon/vision/enhance/base.py:32-47).  The arithmetic (`forward`) and the
parameter names are the reference's own code.

Assertions (all on CPU, fp32): oracle == reference bit-for-bit.
  * src/mon/vision/enhance/llie/zero_dce/zero_dce_v.py   (class ZeroDCEV; kornia's HSV conversions and the
    plum-dispatched guided filter are stubbed, only the four outputs that do not depend on them are pinned)
  * src/mon/core/image/photometry.py, src/mon/core/image/utils.py  (to_image_tensor(normalize=True) /
    to_image_nparray(denormalize=True): the u8 HWC <-> fp32 NCHW frame format)
Outputs: tests/golden/{dce,dcepp_sf1,dcepp_sf2,dcepp_sf12,sci,ruas}_{init,stress}.npz
holding input, weights and the REFERENCE outputs; tests/golden/u8_wire.npz.
"""
from __future__ import annotations

import importlib.util
import os
import sys
import types
from enum import Enum
from pathlib import Path

import numpy as np
import torch

ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))
from oracle import oracle as O  # noqa: E402

REF = Path(os.environ.get("B200DCE_REFERENCE", "/root/reference"))
GOLD = ROOT / "tests" / "golden"


def _load_file(name: str, path: Path):
    spec = importlib.util.spec_from_file_location(name, str(path))
    mod = importlib.util.module_from_spec(spec)
    sys.modules[name] = mod
    spec.loader.exec_module(mod)
    return mod


def install_stub_mon():
    """A minimal `mon` so the reference model files import.  Only plumbing."""
    mon = types.ModuleType("mon")
    core = types.ModuleType("mon.core")
    nn = types.ModuleType("mon.nn")
    glob = types.ModuleType("mon.globals")
    vision = types.ModuleType("mon.vision")
    enhance = types.ModuleType("mon.vision.enhance")
    base = types.ModuleType("mon.vision.enhance.base")
    nn_modules = types.ModuleType("mon.nn.modules")
    activation = types.ModuleType("mon.nn.modules.activation")

    class _Console:
        def log(self, *a, **k):
            pass
    core.console = _Console()
    core.Path = Path

    for k in dir(torch.nn):
        if not k.startswith("_"):
            setattr(nn, k, getattr(torch.nn, k))

    class Loss(torch.nn.Module):  # stands in for mon.nn.Loss (nn/loss/base.py:82)
        def __init__(self, *a, **k):
            super().__init__()
    nn.Loss = Loss

    class _SpaLoss(torch.nn.Module):
        """Only the 4 fixed parameters SpatialConsistencyLoss registers
        (src/mon/nn/loss/image.py:925-945,1082-1085)."""
        def __init__(self, *a, **k):
            super().__init__()
            def p(rows):
                return torch.nn.Parameter(torch.tensor(rows, dtype=torch.float32)[None, None], requires_grad=False)
            self.weight_left = p([[0, 0, 0], [-1, 1, 0], [0, 0, 0]])
            self.weight_right = p([[0, 0, 0], [0, 1, -1], [0, 0, 0]])
            self.weight_up = p([[0, -1, 0], [0, 1, 0], [0, 0, 0]])
            self.weight_down = p([[0, 0, 0], [0, 1, 0], [0, -1, 0]])

    class _NoLoss(torch.nn.Module):
        def __init__(self, *a, **k):
            super().__init__()
    nn.SpatialConsistencyLoss = _SpaLoss
    nn.ExposureControlLoss = _NoLoss
    nn.ColorConstancyLoss = _NoLoss
    nn.TotalVariationLoss = _NoLoss

    class Task(Enum):
        LLIE = "llie"

    class Scheme(Enum):
        UNSUPERVISED = "unsupervised"
        ZERO_REFERENCE = "zero_reference"
        SUPERVISED = "supervised"
        INSTANCE = "instance"

    class _Models(dict):
        def register(self, name=None, arch=None, replace=False):
            def deco(cls):
                self.setdefault(arch, {})[name] = cls
                return cls
            return deco
    glob.MODELS, glob.Task, glob.Scheme = _Models(), Task, Scheme

    class ImageEnhancementModel(torch.nn.Module):
        """Model ctor contract of src/mon/nn/model.py:237-277 (subset)."""
        schemes: list = []
        zoo: dict = {}

        def __init__(self, name=None, in_channels=3, out_channels=None, num_classes=None, weights=None, **kw):
            super().__init__()
            self.name = name
            self.in_channels = in_channels
            self.out_channels = out_channels or num_classes or self.in_channels
            self.weights = weights
            self.predicting = True

        def assert_datapoint(self, datapoint):
            if "image" not in datapoint:
                raise ValueError("The key ``'image'`` must be defined in the `datapoint`.")

        def assert_outputs(self, outputs):
            if "enhanced" not in outputs:
                raise ValueError("The key ``'enhanced'`` must be defined in the `outputs`.")

        def load_weights(self):
            raise NotImplementedError
    base.ImageEnhancementModel = ImageEnhancementModel

    mon.core, mon.nn, mon.globals, mon.vision = core, nn, glob, vision
    vision.enhance, enhance.base, nn.modules, nn_modules.activation = enhance, base, nn_modules, activation
    for n, m in [("mon", mon), ("mon.core", core), ("mon.nn", nn), ("mon.globals", glob), ("mon.vision", vision),
                 ("mon.vision.enhance", enhance), ("mon.vision.enhance.base", base), ("mon.nn.modules", nn_modules),
                 ("mon.nn.modules.activation", activation)]:
        sys.modules[n] = m
    ds = _load_file("mon.nn.modules.conv.dsconv", REF / "src/mon/nn/modules/conv/dsconv.py")
    nn.DSConv2d = ds.DSConv2d
    zd = _load_file("ref_zero_dce", REF / "src/mon/vision/enhance/llie/zero_dce/zero_dce.py")
    zp = _load_file("ref_zero_dcepp", REF / "src/mon/vision/enhance/llie/zero_dce/zero_dcepp.py")
    return zd.ZeroDCE_RE, zp.ZeroDCEpp_RE, glob.MODELS


def load_upstream():
    up = REF / "src/mon_extra/vision/enhance/llie"
    dce = _load_file("up_zero_dce_model", up / "zero_dce/model.py")
    dcepp = _load_file("up_zero_dcepp_model", up / "zero_dce++/model.py")
    sys.path.insert(0, str(up / "sci"))  # sci/model.py does `from loss import LossFunction`
    sci = _load_file("up_sci_model", up / "sci/model.py")
    return dce, dcepp, sci


def load_upstream_ruas():
    """ruas/model.py does `from operations import *` and `import genotypes`: its directory goes first on sys.path."""
    d = REF / "src/mon_extra/vision/enhance/llie/ruas"
    sys.path.insert(0, str(d))
    return _load_file("up_ruas_model", d / "model.py")


def load_reference_image_utils():
    """src/mon/core/image/{photometry,utils}.py, unmodified, under the stub `mon` (they need only `mon.core.pathlib`
    for an isinstance check and `mon.core.image.color` for an unrelated function; cv2 / torchvision are installed)."""
    import pathlib as _pl
    core = sys.modules["mon.core"]
    cpath = types.ModuleType("mon.core.pathlib")
    cpath.Path = _pl.Path
    image = types.ModuleType("mon.core.image")
    color = types.ModuleType("mon.core.image.color")
    core.pathlib, core.image, image.color = cpath, image, color
    for n, m in [("mon.core.pathlib", cpath), ("mon.core.image", image), ("mon.core.image.color", color)]:
        sys.modules[n] = m
    d = REF / "src/mon/core/image"
    phot = _load_file("mon.core.image.photometry", d / "photometry.py")
    utils = _load_file("mon.core.image.utils", d / "utils.py")
    image.photometry, image.utils = phot, utils
    return phot, utils


def load_reference_zero_dce_v():
    """src/mon/vision/enhance/llie/zero_dce/zero_dce_v.py, unmodified, under the stub `mon`.  Its forward calls three
    things that are not available here and do not feed the outputs we pin: `core.rgb_to_hsv` / `core.hsv_to_rgb`
    (kornia) and `filtering.FastGuidedFilter` (needs `plum`).  The stubs: rgb_to_hsv returns (0, 0, V) with V = the
    channel maximum -- kornia's own definition of V (`v = image.max(-3)`), which is all `core.rgb_to_v` takes from it
    (src/mon/core/image/color.py:216-219) --, hsv_to_rgb is the identity, the guided filter returns a nearest resize.
    Pinned: `image_v`, `image_v_lr`, `adjust`, `image_v_fixed_lr`; NOT pinned: `image_v_fixed`, `enhanced`."""
    core = sys.modules["mon.core"]

    def rgb_to_hsv(image, eps=1e-8):
        v = image.max(dim=-3, keepdim=True)[0]
        return torch.cat([torch.zeros_like(v), torch.zeros_like(v), v], dim=-3)
    core.rgb_to_hsv = rgb_to_hsv
    core.rgb_to_v = lambda image, eps=1e-8: rgb_to_hsv(image, eps)[:, 2:3, :, :]   # color.py:216-219
    core.hsv_to_rgb = lambda image: image
    vision = sys.modules["mon.vision"]
    filtering = types.ModuleType("mon.vision.filtering")

    class FastGuidedFilter(torch.nn.Module):
        def __init__(self, radius=1, **kw):
            super().__init__()

        def forward(self, x_lr, y_lr, x_hr):
            return torch.nn.functional.interpolate(y_lr, size=x_hr.shape[-2:])
    filtering.FastGuidedFilter = FastGuidedFilter
    vision.filtering = filtering
    sys.modules["mon.vision.filtering"] = filtering
    zv = _load_file("ref_zero_dce_v", REF / "src/mon/vision/enhance/llie/zero_dce/zero_dce_v.py")
    return zv.ZeroDCEV


def load_reference_gce_loop(iutils):
    """The reference's own statements for GCENet's enhancement loop, run here without the rest of the class:
      * `BrightnessAttentionMap` from src/mon/nn/modules/prior/image.py, the file loaded unmodified under the stub
        `mon` with `core.get_image_channel` = the reference's own (core/image/utils.py:180) and a stand-in `kornia`
        (not installed here) whose `color.rgb_to_hsv` returns (0, 0, V), V = the channel maximum -- kornia's
        definition of the value channel, the only one the map reads;
      * the `if self.bam_gamma in [None, 0.0]: ... else: ...` statement of `GCENet.forward`
        (src/mon/vision/enhance/llie/gcenet/gcenet.py:469-484), cut out of the parsed file by its AST node and compiled
        into `loop(self, image, adjust)`; nothing is retyped, the EnhanceNet / guided filter around it are not run."""
    import ast
    core = sys.modules["mon.core"]
    core.get_image_channel = iutils.get_image_channel
    if "kornia" not in sys.modules:
        kornia = types.ModuleType("kornia")
        kornia.color = types.ModuleType("kornia.color")
        kornia.filters = types.ModuleType("kornia.filters")

        def rgb_to_hsv(image, eps=1e-8):
            v = image.max(dim=-3, keepdim=True)[0]
            return torch.cat([torch.zeros_like(v), torch.zeros_like(v), v], dim=-3)
        kornia.color.rgb_to_hsv = rgb_to_hsv
        sys.modules["kornia"], sys.modules["kornia.color"], sys.modules["kornia.filters"] = kornia, kornia.color, kornia.filters
    prior = _load_file("ref_prior_image", REF / "src/mon/nn/modules/prior/image.py")
    src = (REF / "src/mon/vision/enhance/llie/gcenet/gcenet.py").read_text()
    tree = ast.parse(src)
    cls = next(n for n in tree.body if isinstance(n, ast.ClassDef) and n.name == "GCENet")
    fwd = next(n for n in cls.body if isinstance(n, ast.FunctionDef) and n.name == "forward")
    node = next(n for n in fwd.body if isinstance(n, ast.If) and "bam_gamma" in ast.unparse(n.test))
    assert (node.lineno, node.end_lineno) == (469, 484), (node.lineno, node.end_lineno)
    fn = ast.parse("def loop(self, image, adjust):\n    pass\n    return enhanced, bam, bright, dark")
    fn.body[0].body[0] = node
    ast.fix_missing_locations(fn)
    ns = {"torch": torch}
    exec(compile(fn, "gcenet.py:469-484", "exec"), ns)
    return ns["loop"], prior.BrightnessAttentionMap


def load_reference_guided_filter():
    """src/mon/vision/filtering/{box_filter,guided_filter}.py, unmodified, under the stub `mon`.  Stand-ins for what is
    not installed here and not on the path being pinned: `plum.dispatch` (overloads told apart by their number of
    positional arguments -- FastGuidedFilter.forward has a 3- and a 2-tensor form), `cv2.ximgproc` (numpy branch only),
    `mon.nn.functional` / `mon.nn.init` = torch's."""
    import cv2
    if not hasattr(cv2, "ximgproc"):
        cv2.ximgproc = types.ModuleType("cv2.ximgproc")
        sys.modules["cv2.ximgproc"] = cv2.ximgproc
    if "plum" not in sys.modules:
        plum = types.ModuleType("plum")
        registry = {}

        def dispatch(fn):
            import inspect
            n = len([p_ for p_ in inspect.signature(fn).parameters.values() if p_.name != "self"])
            registry.setdefault(fn.__qualname__, {})[n] = fn

            def call(self, *args):
                return registry[fn.__qualname__][len(args)](self, *args)
            return call
        plum.dispatch = dispatch
        sys.modules["plum"] = plum
    nn = sys.modules["mon.nn"]
    nn.functional = torch.nn.functional
    nn.init = torch.nn.init
    if not hasattr(nn, "AdaptiveBatchNorm2d"):  # a default argument of DeepGuidedFilter (not on the pinned path)
        nn.AdaptiveBatchNorm2d = torch.nn.BatchNorm2d
    sys.modules["mon.nn.functional"] = torch.nn.functional
    sys.modules["mon.nn.init"] = torch.nn.init
    vision = sys.modules["mon.vision"]
    filtering = types.ModuleType("mon.vision.filtering")
    vision.filtering = filtering
    sys.modules["mon.vision.filtering"] = filtering
    d = REF / "src/mon/vision/filtering"
    box = _load_file("mon.vision.filtering.box_filter", d / "box_filter.py")
    filtering.box_filter = box
    gf = _load_file("mon.vision.filtering.guided_filter", d / "guided_filter.py")
    return gf, box


def to_upstream_keys(sd: dict) -> dict:
    return {k.replace("dw_conv", "depth_conv").replace("pw_conv", "point_conv"): v for k, v in sd.items()}


def eq(a, b, what):
    assert a.shape == b.shape, (what, a.shape, b.shape)
    assert torch.equal(a, b), f"{what}: max abs diff {(a - b).abs().max().item():.3e}"


def save(name, x, sd, **outs):
    GOLD.mkdir(parents=True, exist_ok=True)
    arrs = {"x": x.numpy()}
    for k, v in sd.items():
        arrs["w:" + k] = v.numpy()
    for k, v in outs.items():
        arrs["out:" + k] = v.numpy()
    np.savez_compressed(GOLD / f"{name}.npz", **arrs)


def main():
    torch.manual_seed(0)
    torch.set_num_threads(4)
    ZeroDCE_RE, ZeroDCEpp_RE, MODELS = install_stub_mon()
    up_dce, up_dcepp, up_sci = load_upstream()
    assert MODELS["zero_dce"]["zero_dce_re"] is ZeroDCE_RE
    assert MODELS["zero_dce"]["zero_dce++_re"] is ZeroDCEpp_RE
    report = []

    # ---- state_dict layouts straight from the reference classes (F6) ----
    m = ZeroDCE_RE()
    keys = list(m.state_dict().keys())
    exp = [f"e_conv{i}.{p}" for i in range(1, 8) for p in ("weight", "bias")] + \
          [f"loss.loss_spa.weight_{d}" for d in ("left", "right", "up", "down")]
    assert keys == exp, keys
    assert m.out_channels == 24
    mp = ZeroDCEpp_RE()
    keys = list(mp.state_dict().keys())
    exp = [f"e_conv{i}.{c}.{p}" for i in range(1, 8) for c in ("dw_conv", "pw_conv") for p in ("weight", "bias")] + \
          [f"loss.loss_spa.weight_{d}" for d in ("left", "right", "up", "down")]
    assert keys == exp, keys
    assert mp.out_channels == 3
    # init_weights: every conv weight ~ N(0, 0.02) (zero_dcepp.py:157-167)
    for k, v in mp.state_dict().items():
        if k.endswith("conv.weight") and v.numel() > 200:
            assert 0.012 < v.std().item() < 0.028, (k, v.std().item())
    # sf != 1 raises in the mon class as written
    try:
        ZeroDCEpp_RE(scale_factor=12).eval()({"image": torch.rand(1, 3, 48, 48)})
        raise SystemExit("expected the reference ZeroDCEpp_RE(scale_factor=12) to raise")
    except RuntimeError:
        report.append("mon ZeroDCEpp_RE(scale_factor=12).forward raises RuntimeError (F3 confirmed)")

    for regime in ("init", "stress"):
        # ---------------- Zero-DCE ----------------
        x = O.synthetic_lowlight(2, 40, 52, seed=101)
        sd = O.make_dce_state_dict(7, regime)
        m = ZeroDCE_RE().eval()
        m.load_state_dict(sd, strict=False)
        with torch.no_grad():
            ref = m({"image": x})
        assert list(ref.keys()) == [f"adjust_{i}" for i in range(8)] + ["adjust", "enhanced"]
        got = O.zero_dce_forward(x, sd)
        assert list(got.keys()) == list(ref.keys())
        for k in ref:
            eq(got[k], ref[k], f"dce/{regime}/{k}")
        um = up_dce.enhance_net_nopool().eval()
        um.load_state_dict(sd)
        with torch.no_grad():
            e1, e, r = um(x)
        eq(got["enhanced"], e, f"dce-upstream/{regime}/enhanced")
        eq(got["adjust"], r, f"dce-upstream/{regime}/adjust")
        save(f"dce_{regime}", x, sd, adjust=ref["adjust"], enhanced=ref["enhanced"])
        report.append(f"dce/{regime}: oracle == mon class == upstream; |adjust|max={ref['adjust'].abs().max():.3f}")

        # ---------------- Zero-DCE++ ----------------
        for sf, (H, W) in ((1, (36, 44)), (2, (36, 44)), (12, (96, 132))):
            x = O.synthetic_lowlight(2, H, W, seed=200 + sf)
            sd = O.make_dcepp_state_dict(11, regime)
            got = O.zero_dcepp_forward(x, sd, scale_factor=sf)
            um = up_dcepp.enhance_net_nopool(sf).eval()
            um.load_state_dict(to_upstream_keys(sd))
            with torch.no_grad():
                e, r = um(x)
            eq(got["enhanced"], e, f"dcepp-upstream/sf{sf}/{regime}/enhanced")
            eq(got["adjust"], r, f"dcepp-upstream/sf{sf}/{regime}/adjust")
            if sf == 1:
                m = ZeroDCEpp_RE().eval()
                m.load_state_dict(sd, strict=False)
                with torch.no_grad():
                    ref = m({"image": x})
                assert list(ref.keys()) == ["adjust", "enhanced"]
                eq(got["enhanced"], ref["enhanced"], f"dcepp-mon/sf1/{regime}/enhanced")
                eq(got["adjust"], ref["adjust"], f"dcepp-mon/sf1/{regime}/adjust")
            else:
                # explicit-index restatements of ATen's resize agree with ATen
                xd = torch.nn.functional.interpolate(x, scale_factor=1 / sf, mode="bilinear")
                d = np.abs(O.bilinear_down_ref(x.numpy(), sf) - xd.numpy()).max()
                assert d <= 1e-6, d
                a_lr = O.zero_dcepp_lowres_net(xd, sd)
                d = np.abs(O.bilinear_up_ref(a_lr.numpy(), H, W) - r.numpy()).max()
                assert d <= 2e-6, d
            save(f"dcepp_sf{sf}_{regime}", x, sd, adjust=r, enhanced=e)
            report.append(f"dcepp/sf{sf}/{regime}: oracle == upstream"
                          f"{' == mon class' if sf == 1 else ''}; |adjust|max={r.abs().max():.3f}")

        # ---------------- SCI ----------------
        x = O.synthetic_lowlight(2, 40, 52, seed=301)
        sd = O.make_sci_state_dict(13, regime)
        net = up_sci.EnhanceNetwork(layers=1, channels=3).eval()
        full = dict(sd)
        for k in list(sd):
            if k.startswith("conv."):
                full["blocks.0." + k[len("conv."):]] = sd[k]  # aliased module (sci/model.py:27-29)
        net.load_state_dict(full)
        with torch.no_grad():
            illu_ref = net(x)
            r_ref = torch.clamp(x / illu_ref, 0, 1)  # Finetunemodel.forward, sci/model.py:155-159
        illu, r = O.sci_forward(x, sd)
        eq(illu, illu_ref, f"sci/{regime}/illu")
        eq(r, r_ref, f"sci/{regime}/r")
        save(f"sci_{regime}", x, sd, illu=illu_ref, r=r_ref)
        report.append(f"sci/{regime}: oracle == upstream EnhanceNetwork + divide")

    # ---------------- RUAS ----------------
    up_ruas = load_upstream_ruas()
    import genotypes as up_geno  # noqa: E402  (ruas/genotypes.py, on sys.path since load_upstream_ruas)
    assert tuple(n for n, _ in up_geno.IEM.normal) == O.RUAS_IEM and tuple(n for n, _ in up_geno.NRM.normal) == O.RUAS_NRM
    for regime in ("init", "stress"):
        x = O.synthetic_lowlight(2, 40, 52, seed=401)
        sd = O.make_ruas_state_dict(17, regime)
        enh = up_ruas.EnhanceNetwork(iteratioin=3, channel=3, genotype=up_geno.IEM).eval()
        den = up_ruas.DenoiseNetwork(layers=3, channel=6, genotype=up_geno.NRM).eval()
        enh.load_state_dict({k[len("enhance_net."):]: v for k, v in sd.items() if k.startswith("enhance_net.")})
        den.load_state_dict({k[len("denoise_net."):]: v for k, v in sd.items() if k.startswith("denoise_net.")})
        with torch.no_grad():
            u_ref, t_ref = enh(x)
            ud_ref, n_ref = den(u_ref[-1])
        u_list, t_list, n = O.ruas_forward(x, sd)
        for i in range(3):
            eq(u_list[i], u_ref[i], f"ruas/{regime}/u{i}")
            eq(t_list[i], t_ref[i], f"ruas/{regime}/t{i}")
        eq(u_list[3], ud_ref, f"ruas/{regime}/denoised")
        eq(n, n_ref, f"ruas/{regime}/noise")
        save(f"ruas_{regime}", x, sd, u=torch.stack(u_ref), t=torch.stack(t_ref), denoised=ud_ref, noise=n_ref)
        report.append(f"ruas/{regime}: oracle == upstream EnhanceNetwork -> DenoiseNetwork; "
                      f"t in [{t_ref[-1].min():.3f}, {t_ref[-1].max():.3f}], |noise|max={n_ref.abs().max():.3f}")

    # ---------------- ZeroDCE-V core (partial) ----------------
    ZeroDCEV = load_reference_zero_dce_v()
    assert MODELS["zero_dce"]["zero_dce_v"] is ZeroDCEV
    mv = ZeroDCEV()
    assert list(mv.state_dict().keys()) == [f"e_conv{i}.{p}" for i in range(1, 8) for p in ("weight", "bias")] + \
        [f"loss.loss_spa.weight_{d}" for d in ("left", "right", "up", "down")]
    assert (mv.in_channels, mv.num_iters, mv.out_channels, mv.down_size) == (1, 15, 15, 256)
    assert mv.e_conv1.weight.shape == (32, 1, 3, 3) and mv.e_conv7.weight.shape == (15, 64, 3, 3)
    for regime in ("init", "stress"):
        # batch 1 per call: the reference's replace_v_component (:221-230) assigns a [B,1,H,W] tensor into a [B,H,W]
        # slice, which only broadcasts for B == 1 -- the class does not run on larger batches as written
        x = O.synthetic_lowlight(2, 70, 90, seed=601)
        sd = O.make_dce_state_dict(19, regime, num_iters=15, in_channels=1)
        m = ZeroDCEV(down_size=48).eval()
        m.load_state_dict(sd, strict=False)
        refs = []
        with torch.no_grad():
            for i in range(2):
                refs.append(m({"image": x[i:i + 1].clone()}))
        assert list(refs[0].keys()) == ["image_v", "image_v_lr", "adjust", "image_v_fixed_lr", "image_v_fixed", "enhanced"]
        got = O.zero_dce_v_core(x, sd, down_size=48, num_iters=15)
        ref = {k: torch.cat([r[k] for r in refs]) for k in got}
        for k in got:
            eq(got[k], ref[k], f"dcev/{regime}/{k}")
        save(f"dcev_{regime}", x, sd, **ref)
        report.append(f"dcev/{regime}: oracle == mon ZeroDCEV.forward for image_v, image_v_lr, adjust, image_v_fixed_lr "
                      f"(HSV / guided filter stubbed, their outputs not pinned); |adjust|max={ref['adjust'].abs().max():.3f}")

    # ---------------- u8 HWC <-> fp32 NCHW frame format ----------------
    phot, iutils = load_reference_image_utils()
    g = torch.Generator().manual_seed(501)
    frames = torch.randint(0, 256, (2, 36, 52, 3), dtype=torch.uint8, generator=g)
    frames[0, 0, :, 0] = torch.arange(52, dtype=torch.uint8)          # every byte value appears
    frames[0, :32, :8, 1] = torch.arange(256, dtype=torch.uint8).view(32, 8)
    ref_f = torch.cat([iutils.to_image_tensor(f.numpy(), keepdim=False, normalize=True) for f in frames])
    got_f = O.u8hwc_to_f32nchw(frames)
    eq(got_f, ref_f, "u8hwc_to_f32nchw")
    n_div = int((got_f != frames.permute(0, 3, 1, 2).float() / 255.0).sum())
    all_bytes = torch.arange(256, dtype=torch.uint8).view(1, 16, 16, 1).expand(1, 16, 16, 3).contiguous()
    nb = int((O.u8hwc_to_f32nchw(all_bytes)[0, 0] != all_bytes[0, :, :, 0].float() / 255.0).sum())
    e_img = torch.rand(2, 3, 36, 52, generator=g)
    e_img[0, :, 0, :4] = torch.tensor([0.0, 1.0, 0.5, 254.999 / 255.0])
    ref_u = torch.stack([torch.from_numpy(iutils.to_image_nparray(e_img[i:i + 1], keepdim=False, denormalize=True))
                         for i in range(2)])
    got_u = O.f32nchw_to_u8hwc(e_img)
    eq(got_u, ref_u, "f32nchw_to_u8hwc")
    GOLD.mkdir(parents=True, exist_ok=True)
    np.savez_compressed(GOLD / "u8_wire.npz", frames=frames.numpy(), f32=ref_f.numpy(), image=e_img.numpy(),
                        u8=ref_u.numpy())
    report.append(f"u8 wire: oracle == reference to_image_tensor(normalize=True) / to_image_nparray(denormalize=True); "
                  f"multiply by fl32(1/255) differs from /255.0 for {nb} of the 256 byte values ({n_div} elements of the fixture)")

    # ---------------- GCENet's curve loop with the brightness attention map (partial) ----------------
    gce_loop, BAM = load_reference_gce_loop(iutils)
    g = torch.Generator().manual_seed(777)
    gimg = O.synthetic_lowlight(2, 37, 52, seed=702)
    gimg[0, :, 0, :3] = torch.tensor([0.0, 1.0, 0.25]).view(1, 3)       # V = 0 (bam 1), V = 1 (bam 0)
    gadj = torch.rand(2, 3, 37, 52, generator=g) * 2 - 1                # GCENet's EnhanceNet ends in tanh
    arrs = {"image": gimg.numpy(), "adjust": gadj.numpy()}
    for gamma, iters in ((2.5, 8), (1.0, 4), (0.0, 8)):
        self_ = types.SimpleNamespace(bam_gamma=gamma, num_iters=iters, bam=BAM(gamma=gamma, denoise_ksize=None))
        with torch.no_grad():
            re, rb, rbr, rdk = gce_loop(self_, gimg, gadj)
        got = O.gce_curve_loop(gimg, gadj, iters, gamma)
        eq(got["enhanced"], re, f"gce/{gamma}/enhanced")
        tag = f"g{gamma}_i{iters}"
        arrs[f"{tag}:enhanced"] = re.numpy()
        if gamma:
            for k, r_ in (("bam", rb), ("bright", rbr), ("dark", rdk)):
                eq(got[k], r_, f"gce/{gamma}/{k}")
                arrs[f"{tag}:{k}"] = r_.numpy()
        else:
            assert rb is None and rbr is None and rdk is None and got["bam"] is None
    np.savez_compressed(GOLD / "gce_curve.npz", **arrs)
    report.append("gce curve loop: oracle == the reference's own statements (gcenet.py:469-484 cut out by AST, "
                  "prior/image.py BrightnessAttentionMap with kornia's V = channel maximum) for gamma 2.5 / 1.0 / 0.0")

    # ---------------- FastGuidedFilter of the ZeroDCE-V tail ----------------
    gfmod, boxmod = load_reference_guided_filter()
    ZV = load_reference_zero_dce_v.__globals__["sys"].modules["ref_zero_dce_v"].ZeroDCEV
    garrs = {}
    for regime in ("init", "stress"):
        x = O.synthetic_lowlight(2, 70, 90, seed=601)
        sd = O.make_dce_state_dict(19, regime, num_iters=15, in_channels=1)
        core_out = O.zero_dce_v_core(x, sd, down_size=48, num_iters=15)
        with torch.no_grad():
            # the reference's own static method (zero_dce_v.py:208-219) on top of its own FastGuidedFilter; the stubbed
            # `filtering` the class saw at import time is replaced by the real module for this call
            sys.modules["ref_zero_dce_v"].filtering = types.SimpleNamespace(FastGuidedFilter=gfmod.FastGuidedFilter)
            ref_v = ZV.filter_up(core_out["image_v_lr"], core_out["image_v_fixed_lr"], core_out["image_v"])
            ref_box = boxmod.BoxFilter(radius=1)(core_out["image_v_lr"])
        got = O.zero_dce_v_forward(x, sd, down_size=48, num_iters=15)
        eq(O.box_filter_cumsum(core_out["image_v_lr"], 1), ref_box, f"gf/{regime}/box")
        eq(got["image_v_fixed"], ref_v, f"gf/{regime}/image_v_fixed")
        a_, b_ = O.fast_guided_filter(core_out["image_v_lr"], core_out["image_v_fixed_lr"], core_out["image_v"], return_coeffs=True)
        garrs[f"{regime}:image_v_fixed"] = ref_v.numpy()
        garrs[f"{regime}:enhanced_unpinned"] = got["enhanced"].numpy()
        garrs[f"{regime}:A"], garrs[f"{regime}:b"] = a_.numpy(), b_.numpy()
    # GCENet's post-network path: curve loop + GuidedFilter(radius=3, eps=1e-4) (gcenet.py:391,469-485)
    ggen = torch.Generator().manual_seed(778)
    gx = O.synthetic_lowlight(2, 41, 58, seed=703)
    gadj2 = torch.rand(2, 3, 41, 58, generator=ggen) * 2 - 1
    gy = O.gce_curve_loop(gx, gadj2, 15, 2.6)["enhanced"]
    for rad, eps_ in ((3, 1e-4), (5, 1e-8), (9, 1e-2)):
        with torch.no_grad():
            ref_g = gfmod.GuidedFilter(radius=rad, eps=eps_)(gx, gy)
        eq(O.guided_filter(gx, gy, rad, eps_), ref_g, f"guided_filter/r{rad}")
        garrs[f"gf_r{rad}:out"] = ref_g.numpy()
    garrs["gf:x"], garrs["gf:y"] = gx.numpy(), gy.numpy()
    np.savez_compressed(GOLD / "dcev_tail.npz", **garrs)
    report.append("guided filter: oracle == the reference's GuidedFilter(radius, eps) for (3, 1e-4) [GCENet], (5, 1e-8), (9, 1e-2) (bit-exact)")
    report.append("dcev tail: oracle == the reference's ZeroDCEV.filter_up on its own FastGuidedFilter / BoxFilter (bit-exact); "
                  "`enhanced` (kornia HSV, absent here) is the oracle's restatement: UNPINNED")

    (GOLD / "PIN_REPORT.txt").write_text(
        f"generated by oracle/pin_against_reference.py with torch {torch.__version__} (CPU, fp32)\n" +
        "\n".join(report) + "\n")
    print("\n".join(report))
    print("golden fixtures written to", GOLD)


if __name__ == "__main__":
    main()
