"""World-size-2 (gloo, CPU) checks of the multi-GPU host logic: scene partition + global-N loss normaliser + SUM
all-reduce of the flat decoder gradient reproduces the single-process gradient of train_deep_sdf.py:483-533 (computed
by the oracle); latent gradients need no exchange; z-slabs tile the grid."""
import os
import socket

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from deepsdf_b200 import sharding
from oracle import decoder_oracle as O

KW = dict(latent_size=4, dims=[24, 24, 24], geom_dimension=3, dropout=None, dropout_prob=0.0, norm_layers=[0, 1, 2],
          latent_in=[2], weight_norm=True, xyz_in_all=False, use_tanh=False, latent_dropout=False)
S_TOTAL, SPS = 6, 64


def _data():
    rng = np.random.RandomState(0)
    table = torch.from_numpy(rng.normal(0, 0.5, (S_TOTAL, 4)).astype(np.float32))
    xyz = torch.from_numpy(rng.uniform(-1, 1, (S_TOTAL * SPS, 3)).astype(np.float32))
    gt = torch.from_numpy(rng.normal(0, 0.05, (S_TOTAL * SPS,)).astype(np.float32))
    return table, xyz, gt


def _flat(grads, names):
    return torch.cat([grads[n].reshape(-1) for n in names])


def _worker(rank, world, port, out):
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        spec = O.OracleSpec(**KW)
        params = O.make_params(spec, 1)
        table, xyz, gt = _data()
        scenes = sharding.partition_scenes(S_TOTAL, rank, world)
        rows = slice(scenes.start * SPS, scenes.stop * SPS)
        row_scene = torch.arange(scenes.start, scenes.stop).repeat_interleave(SPS)
        n_global = S_TOTAL * SPS
        assert sharding.row_offset(rank, len(scenes) * SPS) == scenes.start * SPS
        _, grads, lat_grad, _ = O.train_loss_and_grads(spec, params, table.clone(), row_scene, xyz[rows], gt[rows], 0.1, 1e-4,
                                                       100, n_total=n_global)
        flat = _flat(grads, list(params))
        sharding.allreduce_sum_(flat)
        t = sharding.max_over_ranks(float(rank + 1), "cpu")
        if rank == 0:
            torch.save({"flat": flat, "max": t}, out)
        torch.save({"lat": lat_grad, "scenes": (scenes.start, scenes.stop)}, out + f".lat{rank}")
    finally:
        dist.destroy_process_group()


def test_scene_sharded_gradients_sum_to_the_single_process_gradient(tmp_path):
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    out = str(tmp_path / "res.pt")
    mp.spawn(_worker, args=(2, port, out), nprocs=2, join=True)
    spec = O.OracleSpec(**KW)
    params = O.make_params(spec, 1)
    table, xyz, gt = _data()
    row_scene = torch.arange(S_TOTAL).repeat_interleave(SPS)
    _, grads, lat_grad, _ = O.train_loss_and_grads(spec, params, table.clone(), row_scene, xyz, gt, 0.1, 1e-4, 100)
    res = torch.load(out)
    ref = _flat(grads, list(params))
    assert torch.allclose(res["flat"], ref, rtol=1e-4, atol=1e-7)
    assert res["max"] == 2.0
    for r in range(2):
        part = torch.load(out + f".lat{r}")
        a, b = part["scenes"]
        assert torch.allclose(part["lat"][a:b], lat_grad[a:b], rtol=1e-4, atol=1e-8)      # owned rows: complete without exchange
        other = torch.ones(S_TOTAL, dtype=torch.bool)
        other[a:b] = False
        assert torch.count_nonzero(part["lat"][other]) == 0


def test_partitions_tile_their_ranges():
    for n, w in [(512, 8), (24, 5), (7, 8), (1024, 3)]:
        edges = [sharding.block_partition(n, r, w) for r in range(w)]
        assert edges[0][0] == 0 and edges[-1][1] == n
        assert all(edges[i][1] == edges[i + 1][0] for i in range(w - 1))
        sizes = [b - a for a, b in edges]
        assert max(sizes) - min(sizes) <= 1
    assert sharding.slab_for_rank(512, 3, 8) == (192, 256)
