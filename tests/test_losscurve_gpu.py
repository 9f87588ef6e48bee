"""North-star criterion "the loss curve over the first 50 epochs of a fixed-seed run stays within 1 %": the fused
SceneTrainer (bf16 tcgen05 kernels, Philox dropout, fused Adam) against the fp32 oracle's training loop
(train_deep_sdf.py:472-545 restated: Embedding max_norm renorm, clamped L1 sum / N, code regulariser with the epoch
ramp, Adam on decoder and on ALL latent rows, Step LR schedules) on the same data, the same initial weights and the same
dropout masks (the trainer's per-step Philox streams exported through dsdf_dropout_mask)."""
import numpy as np
import pytest
import torch

from deepsdf_b200 import _lib as L
from oracle import decoder_oracle as O
from tests.util import build_module, sdf_scale_params

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def test_loss_curve_50_epochs_within_one_percent():
    from deepsdf_b200.trainer import SceneTrainer
    kw = dict(latent_size=16, dims=[512] * 8, geom_dimension=3, dropout=list(range(8)), dropout_prob=0.2,
              norm_layers=list(range(8)), latent_in=[4], weight_norm=True, xyz_in_all=False, use_tanh=False,
              latent_dropout=False)
    specs = {"NetworkSpecs": {k: v for k, v in kw.items() if k != "latent_size"}, "CodeLength": 16,
             "LearningRateSchedule": [{"Type": "Step", "Initial": 1e-4, "Interval": 20, "Factor": 0.5},
                                      {"Type": "Step", "Initial": 1e-3, "Interval": 20, "Factor": 0.5}],
             "SamplesPerScene": 1024, "ScenesPerBatch": 4, "ClampingDistance": 0.1, "CodeRegularization": True,
             "CodeRegularizationLambda": 1e-4, "CodeBound": 1.0}
    ospec = O.OracleSpec(**kw)
    ospec.dropout_scale = O.kernel_dropout_scale(0.2)          # shared masks: same survivor scale as the kernels
    params = sdf_scale_params(ospec, 7, out_scale=0.025)
    S_total, S, sps, epochs = 6, 4, 1024, 50
    rng = np.random.RandomState(1)
    # data: the analytic corner-spheres family of the reference's small experiments (sdf_sampler/microstructures.py:23-36)
    xyz = rng.uniform(-1, 1, (S_total, sps, 3)).astype(np.float32)
    # scaled by 0.25 so that most targets lie inside the +-0.1 clamp band (the reference samples near the surface)
    sdf = 0.25 * np.stack([O.corner_spheres_sdf(xyz[s], 0.4 + 0.1 * s, limit=0.9) for s in range(S_total)]).astype(np.float32)
    data = torch.from_numpy(np.concatenate([xyz, sdf.reshape(S_total, sps, 1)], -1))   # [S_total, sps, 4]
    mod = build_module(kw, params, DEV)
    tr = SceneTrainer(specs, S_total, DEV, seed=3, decoder=mod)
    lat0 = tr.lat_vecs.cpu().clone()
    drop_p = 0.2

    # ---- oracle state
    P = {k: v.clone() for k, v in params.items()}
    m1 = {k: torch.zeros_like(v) for k, v in P.items()}
    m2 = {k: torch.zeros_like(v) for k, v in P.items()}
    Z, zm, zv = lat0.clone(), torch.zeros_like(lat0), torch.zeros_like(lat0)
    got, ref, inside = [], [], []
    order = np.random.RandomState(5)
    for epoch in range(1, epochs + 1):
        ids = torch.from_numpy(np.sort(order.permutation(S_total)[:S])).long()
        batch = data[ids].reshape(-1, 4).contiguous()
        n = batch.shape[0]
        loss = tr.step(batch.to(DEV), ids.to(DEV), epoch)
        got.append(float(loss[0].cpu()))
        seed, r0 = tr.dropout_key(), tr.step_count << 40      # the stream the trainer just used: key + (step << 40) row offset
        masks = {}
        for l in range(ospec.num_linear - 1):
            buf = torch.empty(n, ospec.out_dims[l], dtype=torch.uint8, device=DEV)
            L.check(tr.engine.lib.dsdf_dropout_mask(seed, l, r0, n, ospec.out_dims[l], drop_p, buf.data_ptr(), None))
            masks[l] = buf.cpu().float()
        row_scene = ids.repeat_interleave(sps)
        l_ref, grads, zgrad, pred_raw = O.train_loss_and_grads(ospec, P, Z, row_scene, batch[:, :3], batch[:, 3], 0.1, 1e-4, epoch,
                                                        masks=masks, code_bound=1.0)
        ref.append(l_ref)
        inside.append(float((pred_raw.abs() < 0.1).float().mean()))
        lr_d, lr_z = O.step_lr(1e-4, 20, 0.5, epoch), O.step_lr(1e-3, 20, 0.5, epoch)
        for k in P:
            O.adam_step_(P[k], grads[k], m1[k], m2[k], epoch, lr_d)
        O.adam_step_(Z, zgrad, zm, zv, epoch, lr_z)
    got, ref = np.array(got), np.array(ref)
    rel = np.abs(got - ref) / np.abs(ref)
    print("epoch  fused      oracle     rel")
    for e in (0, 1, 4, 9, 19, 29, 39, 49):
        print(f"{e + 1:4d}  {got[e]:.6f}  {ref[e]:.6f}  {rel[e]:.2e}")
    print(f"max rel diff over 50 epochs {rel.max():.2e}; loss {ref[0]:.5f} -> {ref[-1]:.5f}")
    print(f"fraction of predictions inside the clamp band: first epoch {inside[0]:.2f}, last {inside[-1]:.2f}, min {min(inside):.2f}")
    assert min(inside) >= 0.5                      # non-degenerate: the clamped loss still sees the network
    assert ref[-1] < 0.9 * ref[0]                  # the run actually trains
    # one "epoch" here is a single 4096-row batch, so a curve point carries the sampling noise of one small batch on top of
    # the slow divergence of two Adam trajectories on a sign-gradient loss: the 1 % criterion is applied to the curve
    # smoothed over 5 epochs (what a Logs.pth loss curve averages over is far more) AND pointwise
    k = np.ones(5) / 5
    rel_s = np.abs(np.convolve(got, k, "valid") - np.convolve(ref, k, "valid")) / np.convolve(ref, k, "valid")
    print(f"max rel diff of the 5-epoch moving average {rel_s.max():.2e}")
    assert rel_s.max() <= 1e-2
    assert rel.max() <= 1e-2
    # latent codes and decoder weights stay close to the fp32 trajectory
    zrel = float((tr.lat_vecs.cpu() - Z).norm() / Z.norm())
    names = mod.flat_param_names()
    wrel = max(float((p.detach().cpu() - P[nm]).norm() / P[nm].norm()) for nm, p in zip(names, mod.flat_param_list()))
    print(f"after 50 epochs: latent rel diff {zrel:.2e}, worst parameter tensor rel diff {wrel:.2e}")
    assert zrel <= 5e-2 and wrel <= 5e-2
