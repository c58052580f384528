"""CPU: host-side logic of the front end, and that the C-ABI library loads and exports every
symbol the header declares (no compute calls: there is no GPU here)."""
import ctypes
import os
import re

import pytest
import torch

import dlwp_benchmark_b200 as pkg
from dlwp_benchmark_b200 import native
from dlwp_benchmark_b200.plan import afno_kept_region, centre_crop, neuralop_row_bins
from oracle import afno as oafno, neuralop_fno as ofno, wrappers as owrap

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_library_exports_every_declared_symbol(built_lib):
    header = open(os.path.join(ROOT, "include", "dlwp_b200.h")).read()
    declared = set(re.findall(r"\b(dlwp_[a-z_0-9]+)\s*\(", header))
    declared -= {"dlwp_status", "dlwp_plan"}
    assert declared == set(native.SYMBOLS), declared ^ set(native.SYMBOLS)
    lib = ctypes.CDLL(built_lib)
    for name in declared:
        assert hasattr(lib, name), name
    assert native.lib().dlwp_abi_version() == native.ABI_VERSION


def test_model_desc_matches_header_field_order():
    header = open(os.path.join(ROOT, "include", "dlwp_b200.h")).read()
    body = header[header.index("typedef struct {"):header.index("} dlwp_model_desc;")]
    body = re.sub(r"/\*.*?\*/", "", body, flags=re.S)
    names = []
    for decl in body.split(";"):
        decl = decl.replace("typedef struct {", "").strip()
        if not decl:
            continue
        decl = re.sub(r"^(const\s+)?(int32_t|float)\s*\*?", "", decl).strip()
        names += [n.strip().lstrip("*") for n in decl.split(",")]
    assert names == [f[0] for f in native.ModelDesc._fields_]


@pytest.mark.skipif(torch.cuda.is_available(), reason="checks the no-GPU failure mode")
def test_plan_create_fails_loudly_without_gpu(built_lib):
    m = pkg.TFNO2DModule(context_size=1)
    with torch.no_grad(), pytest.raises(native.NativeError):
        m(prognostic=torch.randn(1, 2, 8, 32, 64), constants=torch.randn(1, 1, 4, 32, 64),
          prescribed=torch.randn(1, 2, 1, 32, 64))
    d = native.ModelDesc()
    h = ctypes.c_void_p()
    rc = native.lib().dlwp_plan_create(ctypes.byref(d), 0, ctypes.byref(h))
    assert rc == -5 and b"no CPU fallback" in native.lib().dlwp_last_error()


def test_row_bins_match_oracle():
    for H, m1 in ((32, 12), (64, 12), (721, 64), (15, 6), (8, 12), (33, 33)):
        assert neuralop_row_bins(H, m1) == tuple(ofno.neuralop_row_bins(H, m1))
    assert neuralop_row_bins(32, 12)[0] == [26, 27, 28, 29, 30, 31, 0, 1, 2, 3, 4, 5]


def test_afno_kept_region_matches_oracle():
    for h, wh, frac in ((32, 33, 1.0), (16, 17, 0.5), (8, 17, 1.0), (45, 91, 0.3)):
        rows, cols = afno_kept_region(h, wh, frac)
        orow, ocol = oafno.afno_kept_region(h, wh, frac)
        assert rows == orow and cols == len(ocol)
    assert afno_kept_region(32, 33, 1.0) == (list(range(32)), 17)


def test_centre_crop():
    assert centre_crop(12, 12) == slice(0, 12)
    assert centre_crop(16, 12) == slice(2, 14)
    assert centre_crop(15, 12) == slice(1, 13)


@pytest.mark.parametrize("cls,ocls,kw", [
    ("FNO2DModule", owrap.FNO2DModule, dict(context_size=2)),
    ("TFNO2DModule", owrap.TFNO2DModule, dict(rank=0.8, context_size=1)),
    ("TFNO2DModule", owrap.TFNO2DModule, dict(rank=0.5, hidden_channels=24, n_modes=[8, 10], context_size=1)),
])
def test_fno_state_dict_layout_matches_oracle(cls, ocls, kw):
    ours, ref = getattr(pkg, cls)(**kw), ocls(**kw)
    a = {k: tuple(v.shape) for k, v in ours.state_dict().items()}
    b = {k: tuple(v.shape) for k, v in ref.state_dict().items()}
    assert a == b
    ours.load_state_dict(ref.state_dict(), strict=True)


def test_afno_state_dict_layout_matches_oracle_and_yaml_kwargs():
    kw = dict(type="FourCastNet", name="test", img_height=32, img_width=64, patch_size=[1, 1],
              constant_channels=4, prescribed_channels=1, prognostic_channels=3, embed_dim=64, filter="AFNO2D",
              depth=4, mlp_ratio=4.0, drop_rate=0.0, drop_path_rate=0.0, num_blocks=4, sparsity_threshold=0.01,
              hard_thresholding_fraction=1.0, context_size=1, use_pos_embed=True)
    ours, ref = pkg.FourCastNet(**kw), oafno.AFNONet(**kw)
    a = {k: tuple(v.shape) for k, v in ours.state_dict().items()}
    b = {k: tuple(v.shape) for k, v in ref.state_dict().items()}
    assert a == b
    assert sum(p.numel() for p in ours.parameters()) == 282752
    with pytest.raises(AssertionError):
        pkg.FourCastNet(embed_dim=30, num_blocks=4)


def test_fourcastnet_fno2d_filter_layout_and_oracle_semantics():
    """``filter: FNO2D`` (fourcastnet.py:130-153, eval'd at :259): parameter names / shapes match the oracle's,
    and the oracle block is ``fno(norm1(x)) + norm1(x) + x`` followed by the Mlp branch."""
    kw = dict(img_height=16, img_width=32, patch_size=[1, 1], constant_channels=2, prescribed_channels=1,
              prognostic_channels=2, embed_dim=32, filter="FNO2D", depth=2, num_blocks=4, context_size=1)
    ours, ref = pkg.FourCastNet(**kw), oafno.AFNONet(**kw).eval()
    a = {k: tuple(v.shape) for k, v in ours.state_dict().items()}
    b = {k: tuple(v.shape) for k, v in ref.state_dict().items()}
    assert a == b
    assert "blocks.0.filter.fno.fno_blocks.convs.weight.0.tensor" in a
    assert a["blocks.0.filter.fno.fno_blocks.convs.weight.0.tensor"] == (32, 32, 12, 7, 2)
    fields, rows_in, rows_out = ours._desc(16, 32)
    assert fields["afno_filter"] == native.AFNO_FILTER_FNO2D and fields["fft_norm"] == native.FFT_NORM_FORWARD
    assert fields["n_cols"] == 7 and len(rows_in) == 12
    blk = ref.blocks[0]
    x = torch.randn(2, 16, 32, 32)
    with torch.no_grad():
        y = blk.norm1(x)
        t2 = blk.filter.fno(y.permute(0, 3, 1, 2)).permute(0, 2, 3, 1) + y + x
        want = blk.mlp(blk.norm2(t2)) + t2
        assert torch.allclose(blk(x), want, atol=1e-6)
    with pytest.raises(NotImplementedError):
        pkg.FourCastNet(filter="SFNO")


def test_forward_guards():
    m = pkg.TFNO2DModule(context_size=1, type="TFNO2DModule", name="test")
    prog = torch.randn(1, 2, 8, 8, 16)
    with pytest.raises(native.NativeError):             # grad mode on: the training path runs on CUDA tensors only, like inference
        m(prognostic=prog, constants=torch.randn(1, 1, 4, 8, 16), prescribed=torch.randn(1, 2, 1, 8, 16))
    fc = pkg.FourCastNet(img_height=8, img_width=16, patch_size=[1, 1], constant_channels=4, prescribed_channels=1, prognostic_channels=8,
                         embed_dim=64, depth=1, num_blocks=4, context_size=1)
    with pytest.raises(native.NativeError):      # the same for FourCastNet
        fc(prognostic=prog, constants=torch.randn(1, 1, 4, 8, 16), prescribed=torch.randn(1, 2, 1, 8, 16))
    with torch.no_grad():
        with pytest.raises(ValueError):
            m(prognostic=prog, constants=None, prescribed=torch.randn(1, 2, 1, 8, 16))
        with pytest.raises(ValueError):
            m(prognostic=prog[:, :1], constants=torch.randn(1, 1, 4, 8, 16), prescribed=torch.randn(1, 1, 1, 8, 16))
        with pytest.raises(native.NativeError):         # CPU tensors: no fallback
            m(prognostic=prog, constants=torch.randn(1, 1, 4, 8, 16), prescribed=torch.randn(1, 2, 1, 8, 16))


REF_CFG = "/root/reference/src/dlwpbench/configs"


@pytest.mark.skipif(not os.path.isdir(REF_CFG), reason="/root/reference is not on this machine")
@pytest.mark.parametrize("fname", ["fno.yaml", "fourcastnet.yaml"])
def test_modules_build_from_the_reference_model_yaml(fname):
    """scripts/train.py:54 / evaluate.py:140 do ``eval(cfg.model.type)(**cfg.model)``: every key of the reference's OWN yaml
    file (including ``type`` and ``name``) goes to the constructor.  The ${data.*} interpolations are resolved from
    configs/data/initial_variables.yaml as Hydra would."""
    import yaml
    data = yaml.safe_load(open(os.path.join(REF_CFG, "data", "initial_variables.yaml")))
    text = open(os.path.join(REF_CFG, "model", fname)).read()
    text = text.replace("${data.height}", str(data["height"])).replace("${data.width}", str(data["width"]))
    cfg = yaml.safe_load(text)
    cls = getattr(pkg, cfg["type"])
    mod = cls(**cfg)
    ref_cls = {"TFNO2DModule": owrap.TFNO2DModule, "FourCastNet": oafno.AFNONet}[cfg["type"]]
    torch.manual_seed(0)
    ref = ref_cls(**cfg)
    # same parameter names and shapes as the (oracle restatement of the) reference module built from the same file
    want = {k: tuple(v.shape) for k, v in ref.state_dict().items()}
    got = {k: tuple(v.shape) for k, v in mod.state_dict().items()}
    assert got == want
    mod.load_state_dict(ref.state_dict(), strict=True)
    assert mod.context_size == cfg["context_size"] and mod.prognostic_channels == cfg["prognostic_channels"]


def test_device_prefetcher_has_no_host_fallback():
    with pytest.raises(native.NativeError):
        pkg.DevicePrefetcher([(torch.zeros(1),)], "cpu")
    if not torch.cuda.is_available():
        with pytest.raises(native.NativeError):
            pkg.DevicePrefetcher([(torch.zeros(1),)], "cuda")
