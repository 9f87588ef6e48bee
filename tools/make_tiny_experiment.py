#!/usr/bin/env python3
"""Write a small synthetic experiment (reference layout: data/SdfSamples/..npz, data/splits/*.json, exp/specs.json) for
smoke-testing ``python -m deepsdf_b200.train -e <dir>/exp`` on 1..N GPUs."""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import decoder_oracle as O  # noqa: E402  (test tooling: the analytic SDF family)

out = sys.argv[1]
n_scenes = int(sys.argv[2]) if len(sys.argv) > 2 else 16
rng = np.random.RandomState(0)
d = os.path.join(out, "data", "SdfSamples", "micro", "cs")
os.makedirs(d, exist_ok=True)
names = []
for s in range(n_scenes):
    xyz = rng.uniform(-1, 1, (6000, 3)).astype(np.float32)
    sdf = 0.25 * O.corner_spheres_sdf(xyz, 0.4 + 0.5 * s / n_scenes, limit=0.9).astype(np.float32)
    rows = np.concatenate([xyz, sdf], 1)
    np.savez(os.path.join(d, f"cs_{s}.npz"), pos=rows[rows[:, 3] > 0], neg=rows[rows[:, 3] <= 0])
    names.append(f"cs_{s}")
os.makedirs(os.path.join(out, "data", "splits"), exist_ok=True)
with open(os.path.join(out, "data", "splits", "cs.json"), "w") as f:
    json.dump({"micro": {"cs": names}}, f)
os.makedirs(os.path.join(out, "exp"), exist_ok=True)
net = dict(dims=[512] * 8, dropout=list(range(8)), dropout_prob=0.2, norm_layers=list(range(8)), latent_in=[4],
           weight_norm=True, xyz_in_all=False, use_tanh=False, latent_dropout=False, geom_dimension=3)
specs = {"Description": "tiny synthetic experiment", "DataSource": os.path.join(out, "data"), "NetworkArch": "deep_sdf_decoder",
         "TrainSplit": os.path.join(out, "data", "splits", "cs.json"), "NetworkSpecs": net, "CodeLength": 16, "NumEpochs": 10,
         "SnapshotFrequency": 5, "AdditionalSnapshots": [], "LogFrequency": 5,
         "LearningRateSchedule": [{"Type": "Step", "Initial": 1e-4, "Interval": 500, "Factor": 0.5},
                                  {"Type": "Step", "Initial": 1e-3, "Interval": 500, "Factor": 0.5}],
         "SamplesPerScene": 2048, "ScenesPerBatch": 8, "ClampingDistance": 0.1, "CodeRegularization": True,
         "CodeRegularizationLambda": 1e-4, "CodeBound": 1.0}
with open(os.path.join(out, "exp", "specs.json"), "w") as f:
    json.dump(specs, f)
print("wrote", out)
