"""CPU: the drop-in boundary — registry, constructor/state_dict contract, error behaviour, the C-ABI
library's exports, and the frame-sharding host logic (world_size-2 gloo).  No compute calls."""
import ctypes
import os
import re
import sys
from pathlib import Path

import pytest
import torch
import torch.multiprocessing as mp

import etss_07_imageenhance_b200 as P
from etss_07_imageenhance_b200 import _lib
from oracle import oracle as O

ROOT = Path(__file__).resolve().parents[1]


# ------------------------------------------------------------------ C ABI
def test_library_exports_every_symbol_the_header_declares(lib_built):
    header = (ROOT / "include" / "b200dce.h").read_text()
    declared = set(re.findall(r"\b(b200_[a-z0-9_]+)\s*\(", header))
    assert len(declared) >= 19
    h = ctypes.CDLL(str(lib_built))
    for name in sorted(declared):
        assert hasattr(h, name), f"{name} declared in include/b200dce.h but not exported"
    assert declared == set(_lib.EXPORTS), declared ^ set(_lib.EXPORTS)


def test_abi_argument_validation_without_a_gpu(lib_built):
    L = _lib.lib()
    assert L.b200_version() >= 100
    assert L.b200_strerror(0) == b"ok"
    assert b"NULL" in L.b200_strerror(_lib.ERR_NULL)
    # argument checks come before any CUDA call, so they are testable on a CPU box
    assert L.b200_curve_shared(None, None, None, 1, 3, 4, 4, 8, None) == _lib.ERR_NULL
    assert L.b200_curve_shared(16, 16, 16, 0, 3, 4, 4, 8, None) == _lib.ERR_BAD_SHAPE
    assert L.b200_curve_shared(16, 20, 16, 1, 3, 4, 4, 8, None) == _lib.ERR_MISALIGNED
    assert L.b200_dcepp_forward(16, 16, 16, None, 1, 50, 48, 12, 8, 16, 1 << 30, None) == _lib.ERR_BAD_SHAPE
    assert L.b200_dcepp_forward(16, 16, 16, None, 1, 48, 48, 12, 8, 16, 8, None) == _lib.ERR_WORKSPACE
    assert L.b200_dce_forward(16, 16, None, 16, 1, 8, 8, 9, 0, 16, 1 << 30, None) == _lib.ERR_UNSUPPORTED
    assert L.b200_dcepp_workspace_bytes(8, 2160, 3840, 12) > 0
    assert L.b200_dce_packed_bytes(_lib.PREC_FP32, 8) == (291 * 288 + 7 * 32) * 4
    if not torch.cuda.is_available():
        assert L.b200_check_device() == _lib.ERR_NO_DEVICE


def test_sci_host_pack_folds_batchnorm(lib_built):
    sd = O.make_sci_state_dict(13, "stress")
    packed = P.ops.sci_pack(sd)
    assert packed.shape == (252,) and not packed.is_cuda
    s = sd["conv.1.weight"].double() / torch.sqrt(sd["conv.1.running_var"].double() + 1e-5)
    w2 = (sd["conv.0.weight"].double() * s.view(3, 1, 1, 1)).float().reshape(-1)
    b2 = ((sd["conv.0.bias"].double() - sd["conv.1.running_mean"].double()) * s + sd["conv.1.bias"].double()).float()
    assert torch.equal(packed[0:81], sd["in_conv.0.weight"].reshape(-1))
    assert torch.allclose(packed[84:165], w2, rtol=0, atol=1e-7)
    assert torch.allclose(packed[165:168], b2, rtol=0, atol=1e-7)


def test_missing_library_fails_loudly(monkeypatch, tmp_path):
    monkeypatch.setenv("B200DCE_LIB", str(tmp_path / "nope.so"))
    monkeypatch.setattr(_lib, "_lib", None)
    with pytest.raises(RuntimeError, match="no fallback"):
        _lib.lib()


# ------------------------------------------------------------------ registry (factory.py:233-436)
def test_registry_entries_and_name_spellings():
    assert P.MODELS["zero_dce"]["zero_dce_re"] is P.ZeroDCE_RE
    assert P.MODELS["zero_dce"]["zero_dce++_re"] is P.ZeroDCEpp_RE
    assert set(P.MODELS.archs) >= {"zero_dce", "sci", "ruas"}
    m = P.MODELS.build(name="zero_dce++_re")
    assert isinstance(m, P.ZeroDCEpp_RE) and m.name == "zero_dce++_re"
    m = P.MODELS.build(config={"name": "zero_dce_re", "num_iters": 4})
    assert isinstance(m, P.ZeroDCE_RE) and m.num_iters == 4 and m.out_channels == 12
    assert isinstance(P.MODELS.build(name="ZeroDceRe"), P.ZeroDCE_RE)   # Pascal -> snake lookup
    assert P.MODELS.build() is None and P.MODELS.build(config={"x": 1}) is None
    with pytest.raises(ValueError):
        P.MODELS.build(name="no_such_model")


def test_second_registration_is_ignored_unless_replace():
    f = P.ModelFactory("t")

    class A:
        pass

    class B:
        pass
    f.register(name="m", arch="a", module=A)
    f.register(name="m", arch="a", module=B)
    assert f["a"]["m"] is A
    f.register(name="m", arch="a", module=B, replace=True)
    assert f["a"]["m"] is B

    @f.register(name="d", arch="a")
    class C:
        pass
    assert f["a"]["d"] is C and f.models == ["m", "d"]


_MON_DROPIN_SCRIPT = r"""
import importlib.util, sys, types
from pathlib import Path
sys.path.insert(0, {root!r})
_s = importlib.util.spec_from_file_location("_r", {root!r} + "/etss_07_imageenhance_b200/registry.py")
R = importlib.util.module_from_spec(_s); _s.loader.exec_module(R)   # by path: the package itself is imported below
ref_factory = Path({ref!r}) / "src/mon/core/factory.py"
if ref_factory.exists():
    # the REFERENCE's own ModelFactory (src/mon/core/factory.py:233-436), unmodified; `humps` is not installed, so a
    # stub with the three spellings it calls stands in for it
    humps = types.ModuleType("humps")
    humps.kebabize, humps.pascalize = R.kebabize, R.pascalize
    humps.depascalize = R.snakeize
    sys.modules["humps"] = humps
    spec = importlib.util.spec_from_file_location("ref_factory", str(ref_factory))
    fac = importlib.util.module_from_spec(spec); spec.loader.exec_module(fac)
    Factory, kind = fac.ModelFactory, "reference"
else:
    Factory, kind = R.ModelFactory, "local"
mon = types.ModuleType("mon")
mon.MODELS = Factory("Models")
sys.modules["mon"] = mon

@mon.MODELS.register(name="zero_dce_re", arch="zero_dce")        # what zero_dce.py:86 / zero_dcepp.py:87 do at import
class RefZeroDCE: pass
@mon.MODELS.register(name="zero_dce++_re", arch="zero_dce")
class RefZeroDCEpp: pass
class Other: pass
mon.MODELS.register(name="zero_dce++_re", arch="zero_dce", module=Other)          # no replace: ignored
assert mon.MODELS["zero_dce"]["zero_dce++_re"] is RefZeroDCEpp

import etss_07_imageenhance_b200 as P                                            # the documented one-liner
assert P.MON_REGISTERED is True
assert mon.MODELS["zero_dce"]["zero_dce++_re"] is P.ZeroDCEpp_RE
assert mon.MODELS["zero_dce"]["zero_dce_re"] is P.ZeroDCE_RE
m = mon.MODELS.build(config={{"name": "zero_dce++_re", "scale_factor": 12}})       # predict.py:46
assert type(m) is P.ZeroDCEpp_RE and m.scale_factor == 12 and m.name == "zero_dce++_re"
assert type(mon.MODELS.build(name="zero_dce_re")) is P.ZeroDCE_RE
print("ok", kind)
"""


def test_import_registers_the_dropins_into_mon_with_replace():
    """with `mon` importable, `import etss_07_imageenhance_b200` re-registers the two classes
    over the reference's entries (factory.py:300-328 ignores a second registration unless replace=True).  Runs in a
    fresh interpreter with a stub `mon` whose MODELS is the reference's own ModelFactory when /root/reference is
    there (this container), else the local clone of it."""
    import subprocess
    ref = os.environ.get("B200DCE_REFERENCE", "/root/reference")
    r = subprocess.run([sys.executable, "-c", _MON_DROPIN_SCRIPT.format(root=str(ROOT), ref=ref)],
                       capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stderr[-2000:]
    assert r.stdout.strip().startswith("ok")
    if Path(ref).exists():
        assert r.stdout.strip() == "ok reference"


def test_without_mon_the_import_is_a_no_op():
    assert P.MON_REGISTERED is False or "mon" in sys.modules


# ------------------------------------------------------------------ state_dict contract
SPA = [f"loss.loss_spa.weight_{d}" for d in ("left", "right", "up", "down")]


def test_zero_dce_state_dict_layout_and_strict_load():
    m = P.ZeroDCE_RE()
    exp = [f"e_conv{i}.{p}" for i in range(1, 8) for p in ("weight", "bias")] + SPA
    assert list(m.state_dict().keys()) == exp
    assert sum(p.numel() for p in m.parameters() if p.requires_grad) == 79416
    assert m.e_conv7.weight.shape == (24, 64, 3, 3) and m.out_channels == 24
    sd = O.make_dce_state_dict(7, "stress")
    with pytest.raises(RuntimeError):          # strict: the loss kernels are part of mon checkpoints
        m.load_state_dict(sd)
    full = {**sd, **{k: m.state_dict()[k] for k in SPA}}
    m2 = P.ZeroDCE_RE(weights={"state_dict": full, "num_iters": 8})   # nested dict + hyper-params from weights
    assert torch.equal(m2.e_conv5.weight, sd["e_conv5.weight"])
    assert torch.equal(m2.loss.loss_spa.weight_left.view(3, 3), torch.tensor([[0., 0, 0], [-1, 1, 0], [0, 0, 0]]))


def test_zero_dcepp_state_dict_layout_init_and_file_weights(tmp_path):
    m = P.ZeroDCEpp_RE(scale_factor=12)
    exp = [f"e_conv{i}.{c}.{p}" for i in range(1, 8) for c in ("dw_conv", "pw_conv") for p in ("weight", "bias")] + SPA
    assert list(m.state_dict().keys()) == exp
    assert sum(p.numel() for p in m.parameters() if p.requires_grad) == 10561
    assert m.out_channels == 3 and m.scale_factor == 12 and m.e_conv7.pw_conv.weight.shape == (3, 64, 1, 1)
    for k, v in m.state_dict().items():        # N(0, 0.02) on every conv weight
        if k.endswith("conv.weight") and v.numel() > 200:
            assert 0.012 < v.std().item() < 0.028
    path = tmp_path / "w.pt"
    torch.save({"state_dict": m.state_dict()}, path)
    m2 = P.ZeroDCEpp_RE(weights=str(path))
    for k, v in m.state_dict().items():
        assert torch.equal(v, m2.state_dict()[k])
    with pytest.raises(ValueError):
        P.ZeroDCEpp_RE(weights=str(tmp_path / "w.txt"))
    (tmp_path / "w.onnx").write_bytes(b"\x08\x07")   # accepted suffix (globals.py:709), but not a torch.load-able file
    with pytest.raises(ValueError, match="ONNX"):
        P.ZeroDCEpp_RE(weights=str(tmp_path / "w.onnx"))
    up = {k.replace("dw_conv", "depth_conv").replace("pw_conv", "point_conv"): v for k, v in m.state_dict().items()}
    assert set(P.remap_upstream_dcepp_keys(up)) == set(m.state_dict())


def test_zero_dce_v_constructor_and_state_dict_layout():
    """the reference class's registry entry, defaults, attribute names and keys (zero_dce_v.py:87-140)."""
    m = P.MODELS.build(name="zero_dce_v")
    assert isinstance(m, P.ZeroDCEV) and P.MODELS["zero_dce"]["zero_dce_v"] is P.ZeroDCEV
    assert (m.in_channels, m.num_channels, m.num_iters, m.out_channels, m.down_size) == (1, 32, 15, 15, 256)
    assert list(m.state_dict().keys()) == [f"e_conv{i}.{p}" for i in range(1, 8) for p in ("weight", "bias")] + SPA
    assert m.e_conv1.weight.shape == (32, 1, 3, 3) and m.e_conv7.weight.shape == (15, 64, 3, 3)
    m2 = P.ZeroDCEV(weights={"state_dict": m.state_dict(), "num_iters": 15, "down_size": 128})
    assert m2.down_size == 128 and torch.equal(m2.e_conv3.weight, m.e_conv3.weight)
    for call in (m, m.forward_core):  # no CPU fallback for either
        with pytest.raises(RuntimeError, match="no CPU fallback"):
            call({"image": torch.rand(1, 3, 8, 8)})


def test_numa_helpers(tmp_path):
    from etss_07_imageenhance_b200 import sched
    assert sched.parse_cpulist("0-3,8,10-11\n") == [0, 1, 2, 3, 8, 10, 11]
    assert sched.parse_cpulist("") == []
    before = os.sched_getaffinity(0)
    with sched.numa_local(0) as cpus:  # no CUDA device here: unknown topology -> no-op
        assert cpus is None or set(cpus) <= before
    assert os.sched_getaffinity(0) == before


def test_ruas_state_dict_and_pack_layout(lib_built):
    """Key names = upstream Network (ruas/model.py:152-200; pinned by oracle/pin_against_reference.py loading the same
    dict into the upstream modules); pack order = include/b200dce.h."""
    sd = O.make_ruas_state_dict(17, "init")
    m = P.RUAS_RE()
    assert set(m.state_dict().keys()) == set(sd.keys())
    m.load_state_dict(sd)
    packed = m.packed_weights()
    assert packed.shape == (3438,) and not packed.is_cuda
    assert torch.equal(packed[:9], sd["enhance_net.iems.0.cell.c1_r.op.weight"].reshape(-1))
    assert torch.equal(packed[9:12], sd["enhance_net.iems.0.cell.c1_r.op.bias"])
    assert torch.equal(packed[3 * 387:3 * 387 + 162], sd["denoise_net.stem.weight"].reshape(-1))
    assert torch.equal(packed[-3:], sd["denoise_net.activate.0.bias"])
    with pytest.raises(RuntimeError):
        m({"image": torch.rand(1, 3, 8, 8)})  # CPU tensor: no CPU fallback


def test_ruas_oracle_matches_golden(golden):
    for regime in ("init", "stress"):
        x, sd, outs = golden(f"ruas_{regime}")
        u, t, n = O.ruas_forward(x, sd)
        assert torch.equal(torch.stack(u[:3]), outs["u"]) and torch.equal(torch.stack(t), outs["t"])
        assert torch.equal(u[3], outs["denoised"]) and torch.equal(n, outs["noise"])


def test_sci_state_dict_matches_upstream_finetunemodel_keys():
    m = P.SCI_RE()
    keys = set(m.state_dict().keys())
    assert {"enhance.in_conv.0.weight", "enhance.conv.0.weight", "enhance.conv.1.running_var",
            "enhance.blocks.0.0.weight", "enhance.blocks.0.1.num_batches_tracked", "enhance.out_conv.0.bias"} <= keys
    assert m.enhance.blocks[0] is m.enhance.conv
    assert sum(p.numel() for p in set(m.parameters())) == 258


# ------------------------------------------------------------------ error behaviour
def test_forward_contract_errors_on_cpu():
    m = P.ZeroDCEpp_RE()
    with pytest.raises(ValueError, match="image"):
        m({"input": torch.rand(1, 3, 8, 8)})
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        m({"image": torch.rand(1, 3, 8, 8)})
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        m(datapoint={"image": torch.rand(1, 3, 8, 8)})   # keyword form (core/thop.py:96-97)
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        P.GraphedForward(m, torch.rand(1, 3, 24, 24))    # the graph helper does not run on the CPU either
    with pytest.raises(NotImplementedError):
        torch.ops.b200dce.curve_shared(torch.rand(1, 3, 4, 4), torch.rand(1, 3, 4, 4), 8)   # CUDA-only dispatch
    with pytest.raises(NotImplementedError):
        P.ZeroDCEpp_RE(scale_factor=1.5)._int_scale()
    with pytest.raises(ValueError):
        P.ZeroDCE_RE(precision="fp8")


def test_fake_impls_give_shapes_without_a_gpu():
    from torch._subclasses.fake_tensor import FakeTensorMode
    with FakeTensorMode():
        x = torch.empty(2, 3, 24, 36, device="cuda")
        packed = torch.empty(100, device="cuda")
        adj, enh = torch.ops.b200dce.dcepp_forward(x, packed, 12, 8, True)
        assert adj.shape == x.shape and enh.shape == x.shape
        adj, enh = torch.ops.b200dce.dce_forward(x, packed, 8, 0, True)
        assert adj.shape == (2, 24, 24, 36)
        a = torch.ops.b200dce.dcepp_lowres(x, packed, 12)
        assert a.shape == (2, 3, 2, 3)
        # the pack ops' fake sizes are the library's own (host-only queries), not literals
        ws = [torch.empty(1, device="cuda")] * 7
        assert torch.ops.b200dce.dcepp_pack(ws, ws, ws, ws).shape == (_lib.lib().b200_dcepp_packed_floats(),)
        for prec in (_lib.PREC_FP32, _lib.PREC_BF16, _lib.PREC_BF16X3):
            assert torch.ops.b200dce.dce_pack(ws, ws, 8, prec).shape == (_lib.lib().b200_dce_packed_bytes(prec, 8),)


# ------------------------------------------------------------------ frame sharding (multi-GPU host logic)
def test_shard_range_partitions_exactly():
    for n in (0, 1, 7, 16, 257):
        for w in (1, 2, 3, 8):
            rs = [P.shard_range(n, r, w) for r in range(w)]
            assert [i for r in rs for i in r] == list(range(n))
            assert max(len(r) for r in rs) - min(len(r) for r in rs) <= 1
    with pytest.raises(ValueError):
        P.shard_range(4, 2, 2)


def _gloo_worker(rank, world, port, q):
    import torch.distributed as dist
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    frames = torch.arange(10, dtype=torch.float32)
    mine = P.shard_range(10, rank, world)
    checksum = float(frames[mine.start:mine.stop].sum())
    ms = P.gather_max_ms(5.0 + rank)
    objs = P.gather_objects({"rank": rank, "frames": list(mine), "sum": checksum})
    q.put((rank, ms, objs))
    dist.destroy_process_group()


def test_frame_sharding_two_ranks_gloo():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29500 + os.getpid() % 2000
    procs = [ctx.Process(target=_gloo_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = sorted(q.get(timeout=120) for _ in procs)
    for p in procs:
        p.join(60)
        assert p.exitcode == 0
    for rank, ms, objs in res:
        assert ms == 6.0                                  # MAX over ranks
        assert [o["rank"] for o in objs] == [0, 1]
        assert sorted(i for o in objs for i in o["frames"]) == list(range(10))
        assert sum(o["sum"] for o in objs) == 45.0        # no frame lost or duplicated, no data-path collective
