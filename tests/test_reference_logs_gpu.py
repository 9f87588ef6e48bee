"""The reference-held training log as a fixture: ``python -m deepsdf_b200.train`` on the regenerated
``experiments/corner_spheres_only`` experiment (Decoder 8x512, CodeLength 16, 20 scenes, ScenesPerBatch 10, SamplesPerScene
16000) against the loss curve the reference's own ``train_deep_sdf.py`` logged for it (``Logs.pth`` ->
``tests/golden/logs_corner_spheres_only.json``, ``oracle/make_golden_logs.py``).

The data set is regenerated exactly as ``evaluation_scripts/04_simple_geom_creation.py:24-29`` builds it
(``CornerSpheresSDF(r, limit=0.9)``, r = linspace(0.4, 1, 20), 1e5 uniform samples in [-1, 1]^3 per scene, fp64 ``pos`` /
``neg`` npz arrays, sdf_sampler/sdf_sampler.py:132-146).  NumPy's generator is never seeded upstream and the reference
never seeds torch either, so the comparison is between two draws of the same experiment: an envelope, not a bitwise check.
"""
import json
import os
import subprocess
import sys

import numpy as np
import pytest
import torch

from oracle import decoder_oracle as O

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLDEN = os.path.join(ROOT, "tests", "golden", "logs_corner_spheres_only.json")


def write_corner_spheres_experiment(root: str, fixture: dict, epochs: int, seed: int = 2024) -> str:
    rng = np.random.RandomState(seed)
    names = []
    ddir = os.path.join(root, "data", "SdfSamples", "microstructure", "corner_spheres")
    os.makedirs(ddir, exist_ok=True)
    for i, r in enumerate(np.linspace(0.4, 1, 20)):
        xyz = rng.uniform(-1, 1, (100000, 3))
        sdf = O.corner_spheres_sdf(xyz, r, limit=0.9).reshape(-1, 1)
        rows = np.hstack([xyz, sdf])                                   # fp64 like the reference's files
        name = f"corner_spheres_{10000 + i}"
        np.savez(os.path.join(ddir, name + ".npz"), neg=rows[rows[:, 3] < 0.0], pos=rows[rows[:, 3] >= 0.0])
        names.append(name)
    split = os.path.join(root, "data", "splits", "corner_spheres.json")
    os.makedirs(os.path.dirname(split), exist_ok=True)
    with open(split, "w") as f:
        json.dump({"microstructure": {"corner_spheres": names}}, f)
    exp = os.path.join(root, "experiments", "corner_spheres_only")
    os.makedirs(exp, exist_ok=True)
    specs = dict(fixture["specs"])
    specs.update({"DataSource": os.path.join(root, "data"), "TrainSplit": split, "TestSplit": split, "NumEpochs": epochs,
                  "SnapshotFrequency": epochs, "AdditionalSnapshots": [], "LogFrequency": epochs})
    with open(os.path.join(exp, "specs.json"), "w") as f:
        json.dump(specs, f, indent=1)
    return exp


def test_corner_spheres_only_follows_the_reference_log(tmp_path):
    with open(GOLDEN) as f:
        fx = json.load(f)
    epochs = 60
    exp = write_corner_spheres_experiment(str(tmp_path), fx, epochs)
    env = dict(os.environ, PYTHONPATH=ROOT + os.pathsep + os.environ.get("PYTHONPATH", ""))
    for k in ("RANK", "WORLD_SIZE", "LOCAL_RANK"):
        env.pop(k, None)
    res = subprocess.run([sys.executable, "-m", "deepsdf_b200.train", "-e", exp], cwd=ROOT, env=env, capture_output=True, text=True,
                         timeout=900)
    assert res.returncode == 0, res.stderr[-2000:]
    logs = torch.load(os.path.join(exp, "Logs.pth"), map_location="cpu", weights_only=False)
    per = fx["iterations_per_epoch"]
    assert logs["epoch"] == epochs and len(logs["loss"]) == per * epochs                 # 20 scenes / 10 per batch = 2 iterations
    got = np.asarray(logs["loss"]).reshape(epochs, per).mean(1)
    ref = np.asarray(fx["loss_per_epoch_first_epochs"][:epochs])
    print("epoch   deepsdf_b200   reference Logs.pth   ratio")
    for e in (1, 2, 3, 5, 8, 10, 15, 20, 30, 40, 50, 60):
        print(f"{e:5d}   {got[e - 1]:.5f}        {ref[e - 1]:.5f}              {got[e - 1] / ref[e - 1]:.3f}")
    # same starting point: predictions ~0, loss = mean |clamp(gt)| of the same distribution
    assert abs(got[0] / ref[0] - 1.0) <= 0.08
    # the fast drop happens in the same epochs (the reference falls below half its initial loss at epoch 10)
    first_below = lambda c: int(np.argmax(c < 0.5 * c[0])) + 1
    print(f"first epoch below half the initial loss: {first_below(got)} (reference {first_below(ref)})")
    assert abs(first_below(got) - first_below(ref)) <= 6
    # plateau level over epochs 20..60 (two independent draws of a stochastic run: 25 % band on the window mean)
    w_got, w_ref = got[19:].mean(), ref[19:].mean()
    print(f"mean loss over epochs 20..{epochs}: {w_got:.5f} (reference {w_ref:.5f}, ratio {w_got / w_ref:.3f})")
    assert 0.75 <= w_got / w_ref <= 1.25
    assert got[-1] < 0.5 * got[0]
    # the reference's checkpoint files exist and carry its key names (workspace.py:9-16, train_deep_sdf.py:96-143)
    sd = torch.load(os.path.join(exp, "ModelParameters", "latest.pth"), map_location="cpu", weights_only=False)
    assert sd["epoch"] == epochs and "module.lin0.parametrizations.weight.original0" in sd["model_state_dict"]
    lat = torch.load(os.path.join(exp, "LatentCodes", "latest.pth"), map_location="cpu", weights_only=False)
    assert lat["latent_codes"]["weight"].shape == (20, 16)
    lm = float(lat["latent_codes"]["weight"].norm(dim=1).mean())
    ref_lm = fx["latent_magnitude_first_epochs"][epochs - 1]
    print(f"mean latent magnitude after {epochs} epochs: {lm:.4f} (reference {ref_lm:.4f})")
    assert 0.5 <= lm / ref_lm <= 2.0
