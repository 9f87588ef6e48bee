#!/usr/bin/env python3
"""Extract the loss envelope of a reference-held training log into a small committed fixture.

TEST INFRASTRUCTURE (oracle/): runs in the build container only, where /root/reference exists.

  python oracle/make_golden_logs.py   ->   tests/golden/logs_corner_spheres_only.json

Source: /root/reference/experiments/corner_spheres_only/Logs.pth, written by the reference's own
train_deep_sdf.py (save_logs, train_deep_sdf.py:179-200) for experiments/corner_spheres_only/specs.json: Decoder
8x512, CodeLength 16, latent_in [4], weight_norm, dropout 0.2, 20 scenes (CornerSpheresSDF(r, limit=0.9),
r = linspace(0.4, 1, 20), 1e5 uniform samples each: evaluation_scripts/04_simple_geom_creation.py:24-29),
ScenesPerBatch 10 -> 2 iterations per epoch, SamplesPerScene 16000, 1000 epochs.  The data set is regenerable from
sdf_sampler/microstructures.py (NumPy's RNG is never seeded there, so a rerun sees the same distribution, not the
same samples): the fixture is a statistical envelope for tests/test_reference_logs_gpu.py, not a bitwise target."""
import json
import os

import numpy as np
import torch

REF = "/root/reference/experiments/corner_spheres_only"
OUT = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden", "logs_corner_spheres_only.json")


def main():
    logs = torch.load(os.path.join(REF, "Logs.pth"), map_location="cpu", weights_only=False)
    with open(os.path.join(REF, "specs.json")) as f:
        specs = json.load(f)
    loss = np.asarray(logs["loss"], dtype=np.float64)
    epochs = int(logs["epoch"])
    per = len(loss) // epochs
    per_epoch = loss[:per * epochs].reshape(epochs, per).mean(1)
    keep = 200
    out = {
        "source": "experiments/corner_spheres_only/Logs.pth (reference, train_deep_sdf.py:179-200)",
        "specs": specs,
        "epochs_total": epochs,
        "iterations_per_epoch": per,
        "loss_per_iteration_first_epochs": [float(x) for x in loss[:per * keep]],
        "loss_per_epoch_first_epochs": [float(x) for x in per_epoch[:keep]],
        "loss_per_epoch_last": float(per_epoch[-1]),
        "latent_magnitude_first_epochs": [float(x) for x in logs["latent_magnitude"][:keep]],
        "learning_rate_first_epoch": [float(x) for x in logs["learning_rate"][0]],
    }
    with open(OUT, "w") as f:
        json.dump(out, f, indent=1)
    print("wrote", OUT, "epoch 1/10/50/100 loss:", per_epoch[0], per_epoch[9], per_epoch[49], per_epoch[99])


if __name__ == "__main__":
    main()
