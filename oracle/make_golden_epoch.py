#!/usr/bin/env python3
"""Generate tests/golden/train_epochs.npz by running the REAL, unmodified reference training entry point
``train_deep_sdf.main_function`` (train_deep_sdf.py:255-600) on CPU - BASELINE.json config #1 scaled down so that the
fixture stays small.  Runs only in the build container (/root/reference mounted); the fixture it writes is committed.

  python oracle/make_golden_epoch.py [--reference /root/reference]

What is run: the network of experiments/corner_spheres_only_small_network/specs.json (4x64, CodeLength 2,
latent_in=[1], use_tanh, weight_norm on every layer) with dropout switched off (torch's dropout RNG cannot be shared
with the Philox stream), 6 scenes of the analytic CornerSpheres family written in the reference's npz format
(sdf_sampler/sdf_sampler.py:146: arrays ``pos`` / ``neg`` of [n,4] xyz+sdf), SamplesPerScene 256, ScenesPerBatch 3,
2 epochs = 4 optimiser iterations, DataLoaderThreads 0.

Nothing of the reference is edited.  Python-level observers record what the loop consumed and produced:
  * SDFSamples.__getitem__ is wrapped to record every (samples, scene index) pair the DataLoader pulled, in order;
  * torch.optim.Adam.__init__ is wrapped to snapshot the decoder / lat_vecs parameters at optimiser construction;
  * Logs.pth, ModelParameters/2.pth, LatentCodes/2.pth written by the reference give the per-iteration loss and the
    final parameters.
Modules the training loop imports but never calls on this path (plyfile, skimage, splinepy, matplotlib, trimesh) are
satisfied with empty stub modules."""
import argparse
import json
import os
import sys
import tempfile
import types

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
from oracle.decoder_oracle import corner_spheres_sdf  # noqa: E402

SPECS = {
    "Description": "golden fixture: corner_spheres_only_small_network, 2 epochs, dropout off",
    "DataSource": "data", "NetworkArch": "deep_sdf_decoder",
    "TrainSplit": "data/splits/cs.json", "TestSplit": "data/splits/cs.json", "ReconstructionSplit": "",
    "NetworkSpecs": {"dims": [64, 64, 64, 64], "dropout": [], "dropout_prob": 0.0, "norm_layers": [0, 1, 2, 3, 4, 5, 6, 7],
                     "latent_in": [1], "xyz_in_all": False, "use_tanh": True, "latent_dropout": False, "weight_norm": True,
                     "geom_dimension": 3},
    "CodeLength": 2, "NumEpochs": 2, "SnapshotFrequency": 2, "AdditionalSnapshots": [],
    "LearningRateSchedule": [{"Type": "Step", "Initial": 0.0005, "Interval": 500, "Factor": 0.5},
                             {"Type": "Step", "Initial": 0.001, "Interval": 500, "Factor": 0.5}],
    "SamplesPerScene": 256, "ScenesPerBatch": 3, "DataLoaderThreads": 0, "ClampingDistance": 0.1,
    "CodeRegularization": True, "CodeRegularizationLambda": 1e-4, "CodeBound": 1.0,
    "LogFrequency": 1,               # Logs.pth is only written every LogFrequency epochs (train_deep_sdf.py:569)
}
N_SCENES, N_STORED = 6, 400


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--reference", default="/root/reference")
    ap.add_argument("--out", default=os.path.join(os.path.dirname(HERE), "tests", "golden", "train_epochs.npz"))
    args = ap.parse_args()

    for name in ("plyfile", "skimage", "skimage.measure", "splinepy", "matplotlib", "matplotlib.pyplot", "trimesh"):
        sys.modules.setdefault(name, types.ModuleType(name))
    sys.path.insert(0, args.reference)

    work = tempfile.mkdtemp(prefix="dsdf_golden_")
    exp = os.path.join(work, "exp")
    os.makedirs(exp)
    rng = np.random.RandomState(42)
    names = []
    for s in range(N_SCENES):
        d = os.path.join(work, "data", "SdfSamples", "microstructure", "corner_spheres")
        os.makedirs(d, exist_ok=True)
        xyz = rng.uniform(-1, 1, (4 * N_STORED, 3)).astype(np.float32)
        sdf = corner_spheres_sdf(xyz, 0.4 + 0.1 * s, limit=0.9).astype(np.float32)
        rows = np.concatenate([xyz, sdf], 1)
        pos, neg = rows[rows[:, 3] > 0][:N_STORED], rows[rows[:, 3] <= 0][:N_STORED]
        assert len(pos) >= 128 and len(neg) >= 128
        nm = f"cs_{s:03d}"
        np.savez(os.path.join(d, nm + ".npz"), pos=pos, neg=neg)
        names.append(nm)
    os.makedirs(os.path.join(work, "data", "splits"))
    with open(os.path.join(work, "data", "splits", "cs.json"), "w") as f:
        json.dump({"microstructure": {"corner_spheres": names}}, f)
    with open(os.path.join(exp, "specs.json"), "w") as f:
        json.dump(SPECS, f)

    os.chdir(work)
    import deep_sdf.data
    import train_deep_sdf

    pulled = []
    orig_getitem = deep_sdf.data.SDFSamples.__getitem__

    def recording_getitem(self, idx):
        smp, i = orig_getitem(self, idx)
        pulled.append((smp.clone(), int(i)))
        return smp, i

    deep_sdf.data.SDFSamples.__getitem__ = recording_getitem
    snap = {}
    orig_adam = torch.optim.Adam.__init__

    # decoder.parameters() is a generator: materialise it before both the snapshot and the real constructor see it
    def adam_init(self, params, *a, **k):
        groups = [dict(g, params=list(g["params"])) for g in params]
        snap["decoder"] = [p.detach().clone() for p in groups[0]["params"]]
        snap["lat"] = groups[1]["params"][0].detach().clone()
        return orig_adam(self, groups, *a, **k)

    torch.optim.Adam.__init__ = adam_init
    torch.manual_seed(20240607)
    torch.set_num_threads(1)
    try:
        train_deep_sdf.main_function("exp", None, 1)
    finally:
        torch.optim.Adam.__init__ = orig_adam
        deep_sdf.data.SDFSamples.__getitem__ = orig_getitem

    logs = torch.load(os.path.join(exp, "Logs.pth"), weights_only=False)
    model = torch.load(os.path.join(exp, "ModelParameters", "2.pth"), weights_only=False)
    lat = torch.load(os.path.join(exp, "LatentCodes", "2.pth"), weights_only=False)
    final = {k[len("module."):] if k.startswith("module.") else k: v for k, v in model["model_state_dict"].items()}
    pnames = list(final.keys())
    assert len(pnames) == len(snap["decoder"])
    spb, sps = SPECS["ScenesPerBatch"], SPECS["SamplesPerScene"]
    n_it = len(pulled) // spb
    assert n_it == len(logs["loss"]) == 4
    out = {"loss": np.asarray(logs["loss"], dtype=np.float64),
           "batches": np.stack([torch.cat([pulled[i * spb + j][0] for j in range(spb)], 0).numpy() for i in range(n_it)]),
           "indices": np.asarray([[pulled[i * spb + j][1] for j in range(spb)] for i in range(n_it)], dtype=np.int64),
           "lat_init": snap["lat"].numpy(), "lat_final": lat["latent_codes"]["weight"].numpy(),
           "its_per_epoch": np.asarray(n_it // SPECS["NumEpochs"]), "specs_json": np.asarray(json.dumps(SPECS))}
    for nm, p0 in zip(pnames, snap["decoder"]):
        assert tuple(p0.shape) == tuple(final[nm].shape), nm
        out["init::" + nm] = p0.numpy()
        out["final::" + nm] = final[nm].numpy()
    assert out["batches"].shape == (n_it, spb * sps, 4)
    np.savez_compressed(args.out, **out)
    print("wrote", args.out, os.path.getsize(args.out), "bytes; losses", out["loss"])


if __name__ == "__main__":
    main()
