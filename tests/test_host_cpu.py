"""CPU-only checks: the C-ABI library loads and exports every symbol include/dsdf.h declares, the drop-in Decoder
mirrors the reference's parameter names / shapes / attributes, NetSpec resolves layer sizes like the oracle."""
import ctypes as C
import os
import re

import numpy as np
import pytest
import torch

from deepsdf_b200 import _lib as L
from deepsdf_b200.spec import NetSpec
from oracle import decoder_oracle as O
from tests.util import build_module, load_cases

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
META, BLOB = load_cases()


def test_library_exports_every_declared_symbol():
    hdr = open(os.path.join(ROOT, "include", "dsdf.h")).read()
    names = set(re.findall(r"\b(dsdf_[a-z0-9_]+)\s*\(", hdr))
    assert len(names) >= 10
    lib = L.lib()
    for n in sorted(names):
        assert hasattr(lib, n), f"{n} declared in include/dsdf.h but not exported by libdsdf_b200.so"
    assert lib.dsdf_abi_version() == L.ABI_VERSION


def test_ctypes_binding_matches_the_header_prototypes():
    """Every function the Python side declares (deepsdf_b200/_lib.py) takes as many arguments as its prototype in
    include/dsdf.h: a drifted binding would pass garbage through the C ABI without any error."""
    hdr = open(os.path.join(ROOT, "include", "dsdf.h")).read()
    hdr = re.sub(r"/\*.*?\*/", " ", hdr, flags=re.S)                       # drop comments
    protos = {}
    for m in re.finditer(r"\b(dsdf_[a-z0-9_]+)\s*\(([^()]*)\)\s*;", hdr):
        args = m.group(2).strip()
        protos[m.group(1)] = 0 if args in ("", "void") else args.count(",") + 1
    lib = L.lib()
    checked = 0
    for name, n_args in sorted(protos.items()):
        fn = getattr(lib, name)
        if fn.argtypes is None:
            continue                                                       # not used from Python
        assert len(fn.argtypes) == n_args, f"{name}: include/dsdf.h has {n_args} parameters, _lib.py declares {len(fn.argtypes)}"
        checked += 1
    assert checked >= 40, checked


def test_no_gpu_means_loud_failure():
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    m = build_module(META["small_4x32_L1"]["spec"])
    with pytest.raises(L.DsdfError):
        m(torch.zeros(4, 4))


@pytest.mark.parametrize("name", sorted(META))
def test_spec_and_module_mirror_reference(name):
    kw = META[name]["spec"]
    o = O.OracleSpec(**kw)
    s = NetSpec(**kw)
    assert s.in_dims == o.in_dims and s.out_dims == o.out_dims
    m = build_module(kw)
    sd = m.state_dict()
    assert list(sd.keys()) == o.param_names() or sorted(sd.keys()) == sorted(o.param_names())
    for k, shp in o.param_shapes().items():
        assert tuple(sd[k].shape) == shp
    assert [n for n, _ in m.named_parameters()] == o.param_names()
    assert m.geom_dimension == kw["geom_dimension"] and m.num_layers == len(kw["dims"]) + 2
    # loads a reference-format checkpoint (DataParallel prefix stripped by the caller, as workspace.py does)
    params = O.make_params(o, META[name]["seed"])
    m.load_state_dict(params, strict=True)
    d = s.desc()
    assert d.num_linear == o.num_linear


def test_fused_support_matrix():
    lib = L.lib()
    exp = {"big_8x512_L256": 1, "big_8x512_L2": 1, "big_8x512_L16": 1, "small_4x64_L2_tanh": 0, "mid_6x128_L1": 0,
           "layernorm_4x48_L3": 0}
    for name, want in exp.items():
        d = NetSpec(**META[name]["spec"]).desc()
        assert lib.dsdf_fused_supported(C.byref(d)) == want, name


def test_reference_export_keys_load():
    z = np.load(os.path.join(ROOT, "tests", "golden", "small_net_export.npz"))
    params = {k[len("param::"):]: torch.from_numpy(z[k]) for k in z.files if k.startswith("param::")}
    m = build_module(META["small_4x64_L2_tanh"]["spec"], params)
    assert torch.equal(m.lin4.parametrizations.weight.original1, params["lin4.parametrizations.weight.original1"])


def test_sampler_host_logic(tmp_path):
    """deep_sdf/data.py:15-33 file list, :88-99 take counts, and the sampler's row bijection (host copy of the kernel's)."""
    from deepsdf_b200.data import get_instance_filenames, remove_nans_np, subsample_counts
    split = {"dsA": {"c1": ["x", "y"], "c2": ["z"]}}
    os.makedirs(tmp_path / "SdfSamples" / "dsA" / "c1")
    np.savez(tmp_path / "SdfSamples" / "dsA" / "c1" / "x.npz", pos=np.zeros((1, 4)), neg=np.zeros((1, 4)))
    assert get_instance_filenames(str(tmp_path), split) == [os.path.join("dsA", "c1", "x.npz"), os.path.join("dsA", "c1", "y.npz"),
                                                            os.path.join("dsA", "c2", "z.npz")]      # missing files only warn
    for n_pos, n_neg, sub in [(500, 700, 256), (77, 900, 256), (800, 64, 256), (128, 128, 256), (8000, 9000, 16000)]:
        half = int(sub / 2)                                      # data.py:88-99 restated
        if n_pos < half:
            want = (n_pos, 2 * half - n_pos)
        elif n_neg < half:
            want = (2 * half - n_neg, n_neg)
        else:
            want = (half, half)
        assert subsample_counts(n_pos, n_neg, sub) == want
    rows = np.array([[0, 0, 0, 1.0], [0, 0, 0, np.nan], [1, 1, 1, -2.0]])
    assert remove_nans_np(rows, 3).tolist() == [[0, 0, 0, 1.0], [1, 1, 1, -2.0]] and remove_nans_np(rows, 3).dtype == np.float32
    lib = L.lib()
    lib.dsdf_sample_perm.restype = C.c_uint32
    lib.dsdf_sample_perm.argtypes = [C.c_uint32, C.c_uint32, C.c_uint64, C.c_int64, C.c_int32]
    for n in (1, 2, 3, 17, 256, 1000, 65537):
        assert sorted(lib.dsdf_sample_perm(i, n, 2024, 5, 1) for i in range(n)) == list(range(n))
    assert [lib.dsdf_sample_perm(i, 1000, 1, 0, 0) for i in range(8)] != [lib.dsdf_sample_perm(i, 1000, 2, 0, 0) for i in range(8)]


def test_torchscript_export_reproduces_reference_cpp_model(tmp_path):
    """deepsdf_b200.export (create_libtorch_executable.py:4-23): the exported TorchScript graph of a decoder carrying the
    weights of the reference's experiments/corner_spheres_only_small_network/cpp_model.pt must give that file's outputs."""
    from deepsdf_b200 import export as E
    z = np.load(os.path.join(ROOT, "tests", "golden", "small_net_export.npz"))
    params = {k[len("param::"):]: torch.from_numpy(z[k]) for k in z.files if k.startswith("param::")}
    m = build_module(META["small_4x64_L2_tanh"]["spec"], params)
    x = torch.from_numpy(z["input"])
    sm = E.to_torchscript(m, x[:1], str(tmp_path / "cpp_model.pt"))
    assert np.abs(sm(x).numpy() - z["output"]).max() <= 1e-6
    back = torch.jit.load(str(tmp_path / "cpp_model.pt"))                     # what the C++ consumer does
    assert torch.equal(back(x), sm(x))
    # a LayerNorm network and a latent_in network go through the same graph builder
    for name in ("layernorm_4x48_L3", "mid_6x128_L1"):
        spec = META[name]["spec"]
        mod = build_module(spec)
        g = E.exported_module(mod)
        xin = torch.randn(5, spec["latent_size"] + 3)
        want = O.decoder_forward(O.OracleSpec(**spec), {k: v.detach() for k, v in mod.state_dict().items()}, xin, training=False)
        assert g(xin).shape == (5, 1) and float((g(xin) - want).abs().max()) <= 1e-6
