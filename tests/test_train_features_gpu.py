"""Training-loop features around the fused step (VERDICT r01 missing #7, ADVICE r01 train.py):
  * --batch_split (train_deep_sdf.py:495-535): chunked gradient accumulation == the unchunked step up to summation order;
  * OptimizerParameters/*.pth is a real torch.optim.Adam state_dict of the reference's two parameter groups, so that the
    reference's load_optimizer (train_deep_sdf.py:116-131) accepts it;
  * --continue (train_deep_sdf.py:420-453): an interrupted + resumed run reproduces the uninterrupted one bit for bit."""
import json
import os
import subprocess
import sys

import numpy as np
import pytest
import torch

from tests.util import build_module

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
NET = dict(dims=[512] * 8, dropout=list(range(8)), dropout_prob=0.2, norm_layers=list(range(8)), latent_in=[4],
           weight_norm=True, xyz_in_all=False, use_tanh=False, latent_dropout=False, geom_dimension=3)


def _specs(L=16, sps=1024, spb=4):
    return {"NetworkSpecs": NET, "CodeLength": L,
            "LearningRateSchedule": [{"Type": "Step", "Initial": 5e-4, "Interval": 500, "Factor": 0.5},
                                     {"Type": "Step", "Initial": 1e-3, "Interval": 500, "Factor": 0.5}],
            "SamplesPerScene": sps, "ScenesPerBatch": spb, "ClampingDistance": 0.1, "CodeRegularization": True,
            "CodeRegularizationLambda": 1e-4, "CodeBound": 1.0}


@pytest.mark.parametrize("split", [2, 3, 7])
def test_batch_split_accumulates_the_same_gradients(split):
    from deepsdf_b200.trainer import SceneTrainer
    from oracle import decoder_oracle as O
    S, sps = 4, 1024                                     # 32 tiles: 3 chunks = 11 + 11 + 10 tiles, scenes span chunk borders
    rng = np.random.RandomState(2)
    xyz = rng.uniform(-1, 1, (S, sps, 3)).astype(np.float32)
    sdf = np.stack([0.25 * O.corner_spheres_sdf(xyz[s], 0.4 + 0.1 * s, limit=0.9) for s in range(S)]).astype(np.float32)
    batch = torch.from_numpy(np.concatenate([xyz, sdf.reshape(S, sps, 1)], -1)).reshape(-1, 4).contiguous().to(DEV)
    ids = torch.tensor([0, 2, 3, 5], device=DEV)
    out = []
    for k in (1, split):
        tr = SceneTrainer(_specs(), 6, DEV, seed=4)
        tr.batch_split = k
        loss = tr.step(batch, ids, epoch=50)
        torch.cuda.synchronize()
        out.append((loss.cpu().clone(), tr.flat_g.cpu().clone(), tr.lat_grad.cpu().clone(), tr.flat_p.cpu().clone()))
    (l1, g1, z1, p1), (l2, g2, z2, p2) = out
    rel_g = float((g1 - g2).norm() / g1.norm())
    rel_z = float((z1 - z2).norm() / z1.norm())
    print(f"batch_split {split}: loss {float(l1[0]):.7f} vs {float(l2[0]):.7f}; decoder grad rel diff {rel_g:.2e}, latent grad rel diff {rel_z:.2e}")
    assert abs(float(l1[0]) - float(l2[0])) <= 1e-6 * abs(float(l1[0])) and abs(float(l1[1]) - float(l2[1])) <= 1e-6 * abs(float(l1[1]))
    assert rel_g <= 1e-5 and rel_z <= 1e-5               # same tiles, same masks, same per-tile arithmetic: only the order of the sums differs
    assert float((p1 - p2).abs().max()) <= 2.1 * 5e-4    # one Adam step moves a parameter by at most lr (sign flips of ~0 gradients)


def _write_experiment(tmp_path, epochs):
    subprocess.check_call([sys.executable, os.path.join(ROOT, "tools", "make_tiny_experiment.py"), str(tmp_path), "8"])
    exp = os.path.join(str(tmp_path), "exp")
    with open(os.path.join(exp, "specs.json")) as f:
        specs = json.load(f)
    specs.update(NumEpochs=epochs, SnapshotFrequency=2, LogFrequency=1, ScenesPerBatch=4, SamplesPerScene=1024)
    with open(os.path.join(exp, "specs.json"), "w") as f:
        json.dump(specs, f)
    return exp, specs


def test_optimizer_checkpoint_is_a_reference_adam_state_dict_and_continue_is_exact(tmp_path, monkeypatch):
    from deepsdf_b200 import train as T
    for k in ("WORLD_SIZE", "RANK", "LOCAL_RANK"):
        monkeypatch.delenv(k, raising=False)
    # ---- an uninterrupted 4-epoch run
    exp_a, specs = _write_experiment(tmp_path / "a", 4)
    res_a = T.main_function(exp_a)
    # ---- the same run stopped after epoch 2 (a snapshot epoch) and continued from that snapshot
    exp_b, _ = _write_experiment(tmp_path / "b", 4)
    T.main_function(exp_b, max_epochs=2)
    res_b = T.main_function(exp_b, continue_from="2")
    assert res_b["epochs"] == 4 and len(res_b["loss"]) == len(res_a["loss"])
    assert res_b["loss"] == res_a["loss"]                                         # bit for bit: deterministic kernels, per-epoch shuffles
    ma = torch.load(os.path.join(exp_a, "ModelParameters", "latest.pth"), weights_only=False)["model_state_dict"]
    mb = torch.load(os.path.join(exp_b, "ModelParameters", "latest.pth"), weights_only=False)["model_state_dict"]
    assert all(torch.equal(ma[k], mb[k]) for k in ma)
    la = torch.load(os.path.join(exp_a, "LatentCodes", "latest.pth"), weights_only=False)["latent_codes"]["weight"]
    lb = torch.load(os.path.join(exp_b, "LatentCodes", "latest.pth"), weights_only=False)["latent_codes"]["weight"]
    assert torch.equal(la, lb)
    # ---- the optimiser file is what the reference's save_optimizer writes: load it the way load_optimizer does (:116-131) into
    #      the optimiser train_deep_sdf.py:400-411 builds
    ck = torch.load(os.path.join(exp_a, "OptimizerParameters", "latest.pth"), weights_only=False)
    assert ck["epoch"] == 4
    decoder = torch.nn.DataParallel(build_module(dict(latent_size=specs["CodeLength"], **specs["NetworkSpecs"])))
    lat_vecs = torch.nn.Embedding(8, specs["CodeLength"], max_norm=specs["CodeBound"])
    opt = torch.optim.Adam([{"params": decoder.parameters(), "lr": 1e-4}, {"params": lat_vecs.parameters(), "lr": 1e-3}])
    opt.load_state_dict(ck["optimizer_state_dict"])
    sd = opt.state_dict()
    assert len(sd["param_groups"]) == 2 and len(sd["param_groups"][1]["params"]) == 1
    n_steps = 4 * (8 // 4)
    for p in list(decoder.parameters()) + list(lat_vecs.parameters()):
        st = opt.state[p]
        assert st["exp_avg"].shape == p.shape and st["exp_avg_sq"].shape == p.shape and int(st["step"]) == n_steps
        assert float(st["exp_avg_sq"].min()) >= 0.0
    assert float(opt.state[next(iter(lat_vecs.parameters()))]["exp_avg"].abs().sum()) > 0.0
    # learning rates of the saved epoch (train_deep_sdf.py:315-318)
    assert abs(sd["param_groups"][0]["lr"] - specs["LearningRateSchedule"][0]["Initial"]) < 1e-12


@pytest.mark.parametrize("clip", [None, 0.05])
def test_graphed_step_replays_the_eager_step_bit_for_bit(clip):
    """SceneTrainer.step_graphed: the whole optimiser step (scene prepare, three chains, finish, both Adam updates, re-pack,
    the side-stream fork/join) captured once and replayed.  Dropout stream and Adam bias corrections come from the device
    step state, so replay k must equal eager step k exactly - fresh masks each step, new batches through the static input
    buffers, a re-capture when an epoch-dependent scalar (regulariser ramp) changes."""
    from deepsdf_b200.trainer import SceneTrainer
    from oracle import decoder_oracle as O
    S, sps = 4, 1024
    rng = np.random.RandomState(11)
    batches = []
    for b in range(3):
        xyz = rng.uniform(-1, 1, (S, sps, 3)).astype(np.float32)
        sdf = np.stack([0.25 * O.corner_spheres_sdf(xyz[s], 0.35 + 0.1 * s + 0.02 * b, limit=0.9) for s in range(S)]).astype(np.float32)
        batches.append((torch.from_numpy(np.concatenate([xyz, sdf.reshape(S, sps, 1)], -1)).reshape(-1, 4).contiguous().to(DEV),
                        torch.tensor([[0, 2, 3, 5], [1, 2, 4, 5], [5, 0, 1, 3]][b], device=DEV)))
    specs = _specs()
    if clip is not None:
        specs["GradientClipNorm"] = clip
    epochs = [50, 50, 50, 51, 51, 51]
    out = []
    for graphed in (False, True):
        tr = SceneTrainer(specs, 6, DEV, seed=9)
        losses = []
        for k, ep in enumerate(epochs):
            smp, ids = batches[k % 3]
            loss = (tr.step_graphed if graphed else tr.step)(smp, ids, epoch=ep)
            losses.append(loss.cpu().clone())
        torch.cuda.synchronize()
        assert tr.step_count == len(epochs) and int(tr.step_state[0].item()) == len(epochs)
        out.append((torch.stack(losses), tr.flat_p.cpu().clone(), tr.lat_vecs.cpu().clone(), tr.exp_avg_sq.cpu().clone()))
        if graphed:
            assert tr.graph_launches >= 8, tr.graph_launches
            print(f"graphed step: {tr.graph_launches} kernels of this library per replay; losses {[round(float(l[0]), 6) for l in losses]}")
    (l0, p0, z0, v0), (l1, p1, z1, v1) = out
    assert torch.equal(l0, l1), (l0, l1)
    assert torch.equal(p0, p1) and torch.equal(z0, z1) and torch.equal(v0, v1)
    # the masks did change from step to step: same batch at step 1 and step 4 gives different losses only through masks + weights,
    # so check the mask words directly
    from deepsdf_b200 import _lib as L
    tr = SceneTrainer(specs, 6, DEV, seed=9)
    m = [torch.zeros(128, 512, dtype=torch.uint8, device=DEV) for _ in range(2)]
    for s, buf in enumerate(m):
        L.check(L.lib().dsdf_dropout_mask(tr.dropout_key(), 0, (s + 1) << 40, 128, 512, 0.2, buf.data_ptr(), None))
    torch.cuda.synchronize()
    assert not torch.equal(m[0], m[1])


@pytest.mark.parametrize("graphed", [False, True])
def test_host_feeder_overlaps_copies_without_changing_results(graphed):
    """HostFeeder (copy of batch i+1 on a copy stream under the step of batch i) == copying and stepping one after the other."""
    from deepsdf_b200.trainer import HostFeeder, SceneTrainer
    S, sps = 4, 1024
    rng = np.random.RandomState(5)
    host = []
    for b in range(5):
        xyz = rng.uniform(-1, 1, (S * sps, 3)).astype(np.float32)
        sdf = (0.2 * np.sin(3 * xyz).sum(1, keepdims=True)).astype(np.float32)
        ids = np.sort(rng.choice(6, S, replace=False)).astype(np.int64)
        host.append((torch.from_numpy(np.concatenate([xyz, sdf], 1)).pin_memory(), torch.from_numpy(ids).pin_memory()))
    out = []
    for piped in (False, True):
        tr = SceneTrainer(_specs(), 6, DEV, seed=3)
        losses = []
        if piped:
            fd = HostFeeder(tr, S * sps, S, graphed=graphed)
            fd.submit(*host[0])
            pending = None
            for b in range(5):
                if b + 1 < 5:
                    fd.submit(*host[b + 1])
                if b < 2:
                    losses.append(fd.step(epoch=7).cpu().clone())
                else:                                            # asynchronous read-back, one step behind
                    h = fd.step_async(epoch=7)
                    if pending is not None:
                        losses.append(pending.wait())
                    pending = h
            losses.append(pending.wait())
            with pytest.raises(Exception):
                fd.step(epoch=7)                                     # nothing submitted
        else:
            staging = {"samples": torch.empty(S * sps, 4, device=DEV), "scene_ids": torch.zeros(S, dtype=torch.int64, device=DEV)}
            for b in range(5):
                losses.append(tr.step_from_host(host[b][0], host[b][1], 7, staging).cpu().clone())
        torch.cuda.synchronize()
        out.append((torch.stack(losses), tr.flat_p.cpu().clone(), tr.lat_vecs.cpu().clone()))
    assert torch.equal(out[0][0], out[1][0]) and torch.equal(out[0][1], out[1][1]) and torch.equal(out[0][2], out[1][2])
