"""CPU: the oracle restatement against the golden vectors produced by the REFERENCE's own code
(oracle/pin_against_reference.py), plus the domain properties the GPU tests rely on."""
import numpy as np
import pytest
import torch

from pathlib import Path

from oracle import oracle as O

GOLDEN_DIR = Path(__file__).resolve().parent / "golden"

CASES_DCEPP = [("dcepp_sf1", 1), ("dcepp_sf2", 2), ("dcepp_sf12", 12)]


@pytest.mark.parametrize("regime", ["init", "stress"])
def test_oracle_dce_matches_reference_golden(golden, regime):
    x, sd, outs = golden(f"dce_{regime}")
    got = O.zero_dce_forward(x, sd)
    assert list(got.keys()) == [f"adjust_{i}" for i in range(8)] + ["adjust", "enhanced"]
    assert torch.equal(got["adjust"], outs["adjust"])
    assert torch.equal(got["enhanced"], outs["enhanced"])
    for i in range(8):  # adjust_i are views of adjust (zero_dce.py:183)
        assert got[f"adjust_{i}"].data_ptr() == got["adjust"][:, 3 * i:3 * i + 3].data_ptr()


@pytest.mark.parametrize("regime", ["init", "stress"])
@pytest.mark.parametrize("name,sf", CASES_DCEPP)
def test_oracle_dcepp_matches_reference_golden(golden, name, sf, regime):
    x, sd, outs = golden(f"{name}_{regime}")
    got = O.zero_dcepp_forward(x, sd, scale_factor=sf)
    assert torch.equal(got["adjust"], outs["adjust"])
    assert torch.equal(got["enhanced"], outs["enhanced"])


@pytest.mark.parametrize("regime", ["init", "stress"])
def test_oracle_sci_matches_reference_golden(golden, regime):
    x, sd, outs = golden(f"sci_{regime}")
    illu, r = O.sci_forward(x, sd)
    assert torch.equal(illu, outs["illu"])
    assert torch.equal(r, outs["r"])


def test_stress_regime_exercises_the_curve(golden):
    _, _, outs = golden("dce_stress")
    assert outs["adjust"].abs().max() > 0.5
    _, _, outs = golden("dcepp_sf12_stress")
    assert outs["adjust"].abs().max() > 0.5


def test_curve_sign_follows_the_code_not_the_paper():
    # zero_dce.py:187: y + a*(y^2 - y) == y - a*y*(1-y); positive a DARKENS
    x = torch.full((1, 3, 2, 2), 0.5)
    a = torch.full((1, 3, 2, 2), 0.5)
    y = O.curve_shared(x, a, 1)
    assert torch.allclose(y, torch.full_like(y, 0.5 + 0.5 * (0.25 - 0.5)))


def test_curve_fixed_points_and_identity():
    g = torch.Generator().manual_seed(3)
    a = torch.rand(2, 3, 5, 7, generator=g) * 2 - 1
    for v in (0.0, 1.0):
        x = torch.full((2, 3, 5, 7), v)
        assert torch.equal(O.curve_shared(x, a, 8), x)
    x = torch.rand(2, 3, 5, 7, generator=g)
    assert torch.equal(O.curve_shared(x, torch.zeros_like(x), 8), x)
    a24 = torch.zeros(2, 24, 5, 7)
    assert torch.equal(O.curve_per_iter(x, a24, 8), x)


def test_zero_weights_give_identity():
    sd = {k: torch.zeros_like(v) for k, v in O.make_dcepp_state_dict(1).items()}
    x = O.synthetic_lowlight(1, 24, 36, seed=5)
    out = O.zero_dcepp_forward(x, sd, scale_factor=12)
    assert torch.equal(out["adjust"], torch.zeros_like(x))
    assert torch.equal(out["enhanced"], x)


@pytest.mark.parametrize("sf,H,W", [(2, 12, 20), (3, 18, 27), (12, 48, 72)])
def test_explicit_bilinear_restatements_match_aten(sf, H, W):
    x = O.synthetic_lowlight(1, H, W, seed=9)
    xd = torch.nn.functional.interpolate(x, scale_factor=1 / sf, mode="bilinear")
    assert np.abs(O.bilinear_down_ref(x.numpy(), sf) - xd.numpy()).max() <= 1e-6
    up = torch.nn.UpsamplingBilinear2d(scale_factor=sf)(xd)
    assert np.abs(O.bilinear_up_ref(xd.numpy(), H, W) - up.numpy()).max() <= 2e-6


def test_sf12_downsample_is_the_2x2_mean_at_offsets_5_6():
    # SURVEY F4
    x = O.synthetic_lowlight(1, 48, 72, seed=4)
    xd = torch.nn.functional.interpolate(x, scale_factor=1 / 12, mode="bilinear")
    m = 0.25 * (x[..., 5::12, 5::12] + x[..., 5::12, 6::12] + x[..., 6::12, 5::12] + x[..., 6::12, 6::12])
    assert (xd - m).abs().max() < 1e-6


def test_synthetic_input_is_deterministic_and_in_range():
    a = O.synthetic_lowlight(2, 24, 36, seed=1)
    b = O.synthetic_lowlight(2, 24, 36, seed=1)
    assert torch.equal(a, b) and a.min() == 0.0 and a.max() == 1.0


def test_infer_resize_rounds_up_to_multiple_of_32():
    x = O.synthetic_lowlight(1, 50, 70, seed=3)
    assert O.infer_resize(x).shape[-2:] == (64, 96)                  # core/image/utils.py:381-382 (ceil)
    assert O.infer_resize(x[..., :32, :64]).shape[-2:] == (32, 64)   # already divisible: identity size
    assert torch.equal(O.infer_resize(x[..., :32, :64]), x[..., :32, :64])
    out = O.infer(lambda t: {"adjust": t * 0.5, "enhanced": t.clamp(0, 1)}, x)
    assert out["enhanced"].shape == x.shape and list(out.keys()) == ["adjust", "enhanced"]


def test_oracle_u8_wire_matches_reference_golden():
    """tests/golden/u8_wire.npz = outputs of the reference's own to_image_tensor(normalize=True) /
    to_image_nparray(denormalize=True); normalisation is a multiply by fl32(1/255), not a division."""
    from pathlib import Path
    z = np.load(Path(__file__).resolve().parent / "golden" / "u8_wire.npz")
    frames, image = torch.from_numpy(z["frames"]), torch.from_numpy(z["image"])
    f = O.u8hwc_to_f32nchw(frames)
    assert torch.equal(f, torch.from_numpy(z["f32"]))
    assert torch.equal(O.f32nchw_to_u8hwc(image), torch.from_numpy(z["u8"]))
    assert (f != frames.permute(0, 3, 1, 2).float() / 255.0).any()   # the division is a different function


@pytest.mark.parametrize("regime", ["init", "stress"])
def test_oracle_zero_dce_v_core_matches_reference_golden(golden, regime):
    """tests/golden/dcev_*.npz = the first four outputs of the reference's own ZeroDCEV.forward (batch 1 per call)."""
    x, sd, outs = golden(f"dcev_{regime}")
    got = O.zero_dce_v_core(x, sd, down_size=48, num_iters=15)
    assert list(got.keys()) == ["image_v", "image_v_lr", "adjust", "image_v_fixed_lr"]
    for k in got:
        assert torch.equal(got[k], outs[k]), k


def test_oracle_gce_curve_loop_matches_reference_golden():
    """tests/golden/gce_curve.npz = GCENet.forward's own loop statements (gcenet.py:469-484) and the reference's
    BrightnessAttentionMap, run by oracle/pin_against_reference.py."""
    z = np.load(GOLDEN_DIR / "gce_curve.npz")
    image, adjust = torch.from_numpy(z["image"]), torch.from_numpy(z["adjust"])
    for gamma, iters in ((2.5, 8), (1.0, 4), (0.0, 8)):
        got = O.gce_curve_loop(image, adjust, iters, gamma)
        tag = f"g{gamma}_i{iters}"
        assert torch.equal(got["enhanced"], torch.from_numpy(z[f"{tag}:enhanced"]))
        for k in ("bam", "bright", "dark"):
            if gamma:
                assert torch.equal(got[k], torch.from_numpy(z[f"{tag}:{k}"])), k
            else:
                assert got[k] is None


@pytest.mark.parametrize("regime", ["init", "stress"])
def test_oracle_zero_dce_v_tail_matches_reference_golden(golden, regime):
    """tests/golden/dcev_tail.npz: `image_v_fixed` = the reference's ZeroDCEV.filter_up on its own FastGuidedFilter /
    BoxFilter (bit-exact pin); `enhanced` is the oracle's own kornia-HSV restatement (unpinned: regression value)."""
    x, sd, _ = golden(f"dcev_{regime}")
    z = np.load(GOLDEN_DIR / "dcev_tail.npz")
    got = O.zero_dce_v_forward(x, sd, down_size=48, num_iters=15)
    assert list(got.keys()) == ["image_v", "image_v_lr", "adjust", "image_v_fixed_lr", "image_v_fixed", "enhanced"]
    assert torch.equal(got["image_v_fixed"], torch.from_numpy(z[f"{regime}:image_v_fixed"]))
    assert torch.equal(got["enhanced"], torch.from_numpy(z[f"{regime}:enhanced_unpinned"]))
    # the conversions are each other's inverse up to rounding
    assert (O.hsv_to_rgb_kornia(O.rgb_to_hsv_kornia(x)) - x).abs().max() <= 1e-6


@pytest.mark.parametrize("radius,eps", [(3, 1e-4), (5, 1e-8), (9, 1e-2)])
def test_oracle_guided_filter_matches_reference_golden(radius, eps):
    """tests/golden/dcev_tail.npz gf_r*: the reference's own GuidedFilter(radius, eps) (GCENet: 3, 1e-4)."""
    z = np.load(GOLDEN_DIR / "dcev_tail.npz")
    x, y = torch.from_numpy(z["gf:x"]), torch.from_numpy(z["gf:y"])
    assert torch.equal(O.guided_filter(x, y, radius, eps), torch.from_numpy(z[f"gf_r{radius}:out"]))
