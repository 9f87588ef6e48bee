"""Device-resident dataset + sampler against the semantics of deep_sdf/data.py:61-110: NaN rows
dropped, half positive then half negative rows per scene, drawn WITHOUT replacement from the right scene, the
scarce-sign rule, determinism; and an epoch of the fused trainer fed from the store."""
import json
import os

import numpy as np
import pytest
import torch

from deepsdf_b200 import _lib as L
from deepsdf_b200.data import DeviceSDFSamples, subsample_counts

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _write_dataset(tmp_path, sizes, dtype=np.float32, nan_every=0):
    """npz files in the reference's layout; row = (scene, index within file array, 0, +-sdf) so that a sampled row names
    its own origin."""
    d = tmp_path / "SdfSamples" / "ds" / "cls"
    os.makedirs(d)
    names = []
    for s, (n_pos, n_neg) in enumerate(sizes):
        pos = np.stack([np.full(n_pos, s), np.arange(n_pos), np.zeros(n_pos), 0.01 + np.arange(n_pos) * 1e-6], 1).astype(dtype)
        neg = np.stack([np.full(n_neg, s), np.arange(n_neg), np.ones(n_neg), -0.01 - np.arange(n_neg) * 1e-6], 1).astype(dtype)
        if nan_every:
            pos[::nan_every, 3] = np.nan
        np.savez(d / f"inst{s}.npz", pos=pos, neg=neg)
        names.append(f"inst{s}")
    return str(tmp_path), {"ds": {"cls": names}}


def test_sampler_semantics(tmp_path):
    sizes = [(500, 700), (90, 900), (800, 64), (300, 300)]
    src, split = _write_dataset(tmp_path, sizes, dtype=np.float64, nan_every=7)
    sps = 256
    ds = DeviceSDFSamples(src, split, sps, 3, DEV)
    assert len(ds) == 4
    n_pos_eff = [n - len(range(0, n, 7)) for n, _ in sizes]             # NaN rows dropped (data.py:61-63)
    assert ds.meta_host[:, 1].tolist() == n_pos_eff and ds.meta_host[:, 3].tolist() == [m for _, m in sizes]
    ids = torch.tensor([2, 0, 1, 3], device=DEV)
    b = ds.sample_batch(ids, seed=99).cpu().numpy().reshape(4, sps, 4)
    assert not np.isnan(b).any()
    for slot, scene in enumerate([2, 0, 1, 3]):
        tp, tn = subsample_counts(n_pos_eff[scene], sizes[scene][1], sps)
        assert tp + tn == sps
        rows = b[slot]
        assert (rows[:, 0] == scene).all()                               # right scene
        assert (rows[:tp, 3] > 0).all() and (rows[tp:, 3] < 0).all()     # positives first, then negatives (data.py:105)
        assert len(set(rows[:tp, 1].tolist())) == tp and len(set(rows[tp:, 1].tolist())) == tn    # without replacement
        assert not (rows[:tp, 1] % 7 == 0).any()                         # the NaN rows (every 7th positive) never appear
    # scarce-sign rule (data.py:92-97): scene 1 has 77 positives < 128 -> all of them + 179 negatives; scene 2: 64 negatives
    assert subsample_counts(n_pos_eff[1], 900, sps) == (n_pos_eff[1], sps - n_pos_eff[1])
    assert subsample_counts(n_pos_eff[2], 64, sps) == (sps - 64, 64)
    # deterministic in the seed, different across seeds and scenes
    b2 = ds.sample_batch(ids, seed=99).cpu().numpy().reshape(4, sps, 4)
    b3 = ds.sample_batch(ids, seed=100).cpu().numpy().reshape(4, sps, 4)
    assert np.array_equal(b, b2) and not np.array_equal(b, b3)
    # the kernel's bijection is the exported host function
    lib = L.lib()
    tp0, _ = subsample_counts(n_pos_eff[0], 700, sps)
    kept = [i for i in range(500) if i % 7]
    want = [kept[lib.dsdf_sample_perm(i, len(kept), 99, 0, 0)] for i in range(tp0)]
    assert b[1][:tp0, 1].astype(int).tolist() == want
    # every row is reachable and equally likely: over many seeds the first positive pick covers the set evenly
    picks = np.array([lib.dsdf_sample_perm(0, 50, s, 3, 0) for s in range(20000)])
    cnt = np.bincount(picks, minlength=50)
    assert cnt.min() > 300 and cnt.max() < 500                           # expectation 400


def test_not_enough_rows_is_an_error(tmp_path):
    src, split = _write_dataset(tmp_path, [(10, 20)])
    with pytest.raises(L.DsdfError):
        DeviceSDFSamples(src, split, 64, 3, DEV)


def test_epoch_from_device_store_trains():
    from deepsdf_b200.trainer import SceneTrainer
    from oracle import decoder_oracle as O
    rng = np.random.RandomState(0)
    arrays = []
    for s in range(6):
        xyz = rng.uniform(-1, 1, (3000, 3)).astype(np.float32)
        sdf = 0.25 * O.corner_spheres_sdf(xyz, 0.4 + 0.1 * s, limit=0.9).astype(np.float32)
        rows = np.concatenate([xyz, sdf], 1)
        arrays.append((rows[rows[:, 3] > 0], rows[rows[:, 3] <= 0]))
    ds = DeviceSDFSamples(None, None, 1024, 3, DEV, arrays=arrays)
    specs = {"NetworkSpecs": dict(dims=[512] * 8, dropout=list(range(8)), dropout_prob=0.2, norm_layers=list(range(8)),
                                  latent_in=[4], weight_norm=True, xyz_in_all=False, use_tanh=False, latent_dropout=False,
                                  geom_dimension=3),
             "CodeLength": 16, "LearningRateSchedule": [{"Type": "Step", "Initial": 1e-4, "Interval": 500, "Factor": 0.5},
                                                        {"Type": "Step", "Initial": 1e-3, "Interval": 500, "Factor": 0.5}],
             "SamplesPerScene": 1024, "ScenesPerBatch": 3, "ClampingDistance": 0.1, "CodeRegularization": True,
             "CodeRegularizationLambda": 1e-4, "CodeBound": 1.0}
    tr = SceneTrainer(specs, 6, DEV, seed=1)
    g = torch.Generator().manual_seed(0)
    first = last = None
    for epoch in range(1, 16):
        losses = [float(l[0]) for l in tr.train_epoch(ds, epoch, generator=g)]
        assert len(losses) == 2 and all(np.isfinite(losses))            # 6 scenes / 3 per batch, drop_last
        first = first if first is not None else np.mean(losses)
        last = np.mean(losses)
    print(f"mean epoch loss {first:.5f} -> {last:.5f}")
    assert last < 0.9 * first


def test_train_entry_point_writes_reference_layout(tmp_path, monkeypatch):
    """deepsdf_b200.train on a tiny experiment: reference experiment layout in, reference checkpoint formats out
    (train_deep_sdf.py:96-143,179-200; deep_sdf/workspace.py:9-21) - loadable with DataParallel's ``module.`` prefix."""
    from deepsdf_b200 import train as T
    from oracle import decoder_oracle as O
    from tests.util import build_module
    rng = np.random.RandomState(0)
    d = tmp_path / "data" / "SdfSamples" / "micro" / "cs"
    os.makedirs(d)
    names = []
    for s in range(4):
        xyz = rng.uniform(-1, 1, (2000, 3)).astype(np.float32)
        sdf = 0.25 * O.corner_spheres_sdf(xyz, 0.4 + 0.1 * s, limit=0.9).astype(np.float32)
        rows = np.concatenate([xyz, sdf], 1)
        np.savez(d / f"cs_{s}.npz", pos=rows[rows[:, 3] > 0], neg=rows[rows[:, 3] <= 0])
        names.append(f"cs_{s}")
    os.makedirs(tmp_path / "data" / "splits")
    with open(tmp_path / "data" / "splits" / "cs.json", "w") as f:
        json.dump({"micro": {"cs": names}}, f)
    exp = tmp_path / "exp"
    os.makedirs(exp)
    net = dict(dims=[512] * 8, dropout=list(range(8)), dropout_prob=0.2, norm_layers=list(range(8)), latent_in=[4],
               weight_norm=True, xyz_in_all=False, use_tanh=False, latent_dropout=False, geom_dimension=3)
    specs = {"Description": "test", "DataSource": str(tmp_path / "data"), "NetworkArch": "deep_sdf_decoder",
             "TrainSplit": str(tmp_path / "data" / "splits" / "cs.json"), "NetworkSpecs": net, "CodeLength": 8, "NumEpochs": 3,
             "SnapshotFrequency": 2, "AdditionalSnapshots": [], "LogFrequency": 1,
             "LearningRateSchedule": [{"Type": "Step", "Initial": 1e-4, "Interval": 500, "Factor": 0.5},
                                      {"Type": "Step", "Initial": 1e-3, "Interval": 500, "Factor": 0.5}],
             "SamplesPerScene": 512, "ScenesPerBatch": 2, "ClampingDistance": 0.1, "CodeRegularization": True,
             "CodeRegularizationLambda": 1e-4, "CodeBound": 1.0}
    with open(exp / "specs.json", "w") as f:
        json.dump(specs, f)
    for k in ("WORLD_SIZE", "RANK", "LOCAL_RANK"):
        monkeypatch.delenv(k, raising=False)
    res = T.main_function(str(exp))
    assert res["epochs"] == 3 and len(res["loss"]) == 3 * 2              # 4 scenes / 2 per batch
    assert sorted(os.listdir(exp / "ModelParameters")) == ["2.pth", "latest.pth"]
    ck = torch.load(exp / "ModelParameters" / "latest.pth", weights_only=False)
    assert ck["epoch"] == 3 and all(k.startswith("module.") for k in ck["model_state_dict"])
    dp = torch.nn.DataParallel(build_module(dict(latent_size=8, **net)))
    dp.load_state_dict(ck["model_state_dict"], strict=True)               # what ws.load_model_parameters does
    lat = torch.load(exp / "LatentCodes" / "latest.pth", weights_only=False)
    assert lat["latent_codes"]["weight"].shape == (4, 8)                  # an nn.Embedding state_dict
    emb = torch.nn.Embedding(4, 8)
    emb.load_state_dict(lat["latent_codes"])
    logs = torch.load(exp / "Logs.pth", weights_only=False)
    assert set(logs) == {"epoch", "loss", "learning_rate", "timing", "latent_magnitude", "param_magnitude"}
    assert len(logs["loss"]) == 6 and len(logs["learning_rate"]) == 3 and "lin0.bias" in logs["param_magnitude"]
    assert os.path.isfile(exp / "OptimizerParameters" / "latest.pth")


def test_reconstruct_cli_writes_codes_and_improves(tmp_path, monkeypatch):
    """deepsdf_b200.reconstruct on the checkpoint of a short training run: Reconstructions/<epoch>/Codes/<dataset>/<class>/
    <instance>.pth ([1,1,L], workspace.py:137-149) for every instance of the split; the optimised code must fit its shape's
    samples better than the N(0, 0.01) starting code does; --skip leaves existing codes untouched."""
    import subprocess
    import sys
    from deepsdf_b200 import reconstruct as R
    from deepsdf_b200 import train as T
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    subprocess.check_call([sys.executable, os.path.join(root, "tools", "make_tiny_experiment.py"), str(tmp_path), "6"])
    exp = str(tmp_path / "exp")
    with open(os.path.join(exp, "specs.json")) as f:
        specs = json.load(f)
    specs.update(NumEpochs=30, SnapshotFrequency=30, LogFrequency=30, ScenesPerBatch=6, SamplesPerScene=2048)
    specs["LearningRateSchedule"][0]["Initial"] = 5e-4
    with open(os.path.join(exp, "specs.json"), "w") as f:
        json.dump(specs, f)
    for k in ("WORLD_SIZE", "RANK", "LOCAL_RANK"):
        monkeypatch.delenv(k, raising=False)
    T.main_function(exp)
    short = R.reconstruct_split(exp, "latest", iters=1, num_samples=1000, seed=3)
    res = R.reconstruct_split(exp, "latest", iters=200, num_samples=1000, resolution=32, seed=3)
    assert res["epoch"] == 30 and len(res["loss"]) == 6
    codes = tmp_path / "exp" / "Reconstructions" / "30" / "Codes" / "micro" / "cs"
    assert sorted(os.listdir(codes)) == [f"cs_{i}.pth" for i in range(6)]
    z = torch.load(codes / "cs_0.pth", weights_only=False)
    assert z.shape == (1, 1, 16) and torch.isfinite(z).all()
    ply = open(tmp_path / "exp" / "Reconstructions" / "30" / "Meshes" / "micro" / "cs" / "cs_0.ply", "rb").read()
    assert ply.startswith(b"ply\nformat binary_little_endian 1.0\nelement vertex ")        # workspace.py:122-135 file name
    l0, l1 = np.mean(list(short["loss"].values())), np.mean(list(res["loss"].values()))
    print(f"reconstruction loss: first iteration {l0:.5f} -> after 200 iterations {l1:.5f}")
    assert l1 < l0
    stamp = os.path.getmtime(codes / "cs_0.pth")
    again = R.reconstruct_split(exp, "latest", iters=5, num_samples=1000, skip=True)
    assert again["loss"] == {} and os.path.getmtime(codes / "cs_0.pth") == stamp


def test_sampler_properties_at_config5_size():
    """BASELINE.json config #5: 4096 scenes x 100 000 stored rows (6.5 GB resident), SamplesPerScene 16000, 64 scenes per
    batch.  Size-independent properties of one sampled batch: every row comes from its scene and from the right sign
    half, no stored row is drawn twice (the keyed bijection), the draw is a function of the seed only."""
    n_sc, n_rows, sps, spb = 4096, 100000, 16000, 64
    n_pos = 45000                                                         # 45 000 positives, 55 000 negatives per scene
    ds = object.__new__(DeviceSDFSamples)                                 # the store is built on the device: rows encode their origin
    ds.subsample, ds.geom_dimension, ds.device = sps, 3, torch.device(DEV)
    ds.npyfiles = [f"<scene {i}>" for i in range(n_sc)]
    idx = torch.arange(n_rows, device=DEV, dtype=torch.float32)
    one = torch.stack([torch.zeros_like(idx), idx, torch.zeros_like(idx), torch.where(idx < n_pos, 0.05, -0.05)], 1)
    ds.store = one.repeat(n_sc, 1)                                        # [409.6 M, 4] fp32
    ds.store[:, 0] = torch.arange(n_sc, device=DEV, dtype=torch.float32).repeat_interleave(n_rows)
    off = torch.arange(n_sc, dtype=torch.int64) * n_rows
    ds.meta_host = torch.stack([off, torch.full_like(off, n_pos), off + n_pos, torch.full_like(off, n_rows - n_pos)], 1).numpy()
    ds.meta = torch.from_numpy(ds.meta_host).to(DEV)
    ds.total_rows = n_sc * n_rows
    assert ds.store.shape == (n_sc * n_rows, 4) and ds.store.numel() * 4 > 6.5e9
    ids = torch.sort(torch.randperm(n_sc, generator=torch.Generator().manual_seed(1))[:spb]).values.to(DEV)
    b = ds.sample_batch(ids, seed=1234)
    assert b.shape == (spb * sps, 4)
    v = b.view(spb, sps, 4)
    assert torch.equal(v[:, :, 0], ids.float()[:, None].expand(-1, sps))                     # right scene, scene-major
    half = sps // 2
    assert bool((v[:, :half, 3] > 0).all()) and bool((v[:, half:, 3] < 0).all())             # pos half then neg half
    assert bool((v[:, :half, 1] < n_pos).all()) and bool((v[:, half:, 1] >= n_pos).all())
    srt = torch.sort(v[:, :, 1], dim=1).values
    assert bool((srt[:, 1:] != srt[:, :-1]).all())                                           # without replacement
    assert torch.equal(ds.sample_batch(ids, seed=1234), b) and not torch.equal(ds.sample_batch(ids, seed=1235), b)
    # the draw is spread over the whole stored range, not a prefix
    assert float(v[:, :half, 1].max()) > 0.99 * n_pos and float(v[:, half:, 1].max()) > 0.99 * n_rows
    del ds, b, v, one
    torch.cuda.empty_cache()
