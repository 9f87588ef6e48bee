"""GPU-resident counterpart of ``deep_sdf/data.py``: the scene list of a split file, the SDF-sample store and the
per-batch subsampler.

The reference's ``SDFSamples.__getitem__`` (data.py:142-194) re-reads a scene's npz from disk for every batch, drops NaN
rows, draws ``randperm`` subsamples on the CPU and lets the DataLoader collate them (``unpack_sdf_samples``,
data.py:75-110).  ``DeviceSDFSamples`` reads every npz ONCE (same on-disk format: arrays ``pos`` / ``neg`` of [n,4]
xyz+sdf, fp32 or fp64 - sdf_sampler/sdf_sampler.py:146, src/PreprocessMesh.cpp:196-221), keeps all rows in HBM, and
``sample_batch`` writes a scene-major [S*SamplesPerScene, 4] batch with one kernel (``dsdf_sample_batch``): half positive
then half negative rows per scene, drawn without replacement.  The batch is exactly what ``SceneTrainer.step`` consumes.
"""
from __future__ import annotations

import logging
import os
from typing import Dict, List, Optional, Sequence

import numpy as np
import torch

from . import _lib as L

sdf_samples_subdir = "SdfSamples"        # deep_sdf/workspace.py:19


def get_instance_filenames(data_source: str, split: Dict) -> List[str]:
    """deep_sdf/data.py:15-33: ``<dataset>/<class>/<instance>.npz`` for every instance of the split; a missing file is
    only a warning there, and so it is here."""
    npzfiles = []
    for dataset in split:
        for class_name in split[dataset]:
            for instance_name in split[dataset][class_name]:
                instance_filename = os.path.join(dataset, class_name, instance_name + ".npz")
                if not os.path.isfile(os.path.join(data_source, sdf_samples_subdir, instance_filename)):
                    logging.warning("Requested non-existent file '{}'".format(instance_filename))
                npzfiles += [instance_filename]
    return npzfiles


def remove_nans_np(rows: np.ndarray, geom_dimension: int) -> np.ndarray:
    """deep_sdf/data.py:61-63: drop rows whose SDF column is NaN, cast to fp32."""
    return rows[~np.isnan(rows[:, geom_dimension])].astype(np.float32)


def subsample_counts(n_pos: int, n_neg: int, subsample: int):
    """Rows taken from the positive / negative set of one scene, deep_sdf/data.py:88-99."""
    half = int(subsample / 2)
    if n_pos < half:
        return n_pos, 2 * half - n_pos
    if n_neg < half:
        return 2 * half - n_neg, n_neg
    return half, half


class DeviceSDFSamples:
    def __init__(self, data_source: Optional[str], split: Optional[Dict], subsample: int, geom_dimension: int = 3, device="cuda",
                 arrays: Optional[Sequence] = None):
        """Same constructor meaning as ``SDFSamples(data_source, split, subsample, geom_dimension)``; ``arrays`` (an iterable
        of (pos[n,4], neg[m,4]) pairs, numpy or torch - device tensors stay on the device) replaces the files for synthetic
        data."""
        if geom_dimension != 3:
            raise L.DsdfError("DeviceSDFSamples stores [n,4] rows (geom_dimension 3)")
        self.subsample = subsample
        self.geom_dimension = geom_dimension
        self.device = torch.device(device)
        if arrays is None:
            self.npyfiles = get_instance_filenames(data_source, split)

            def scenes():
                for f in self.npyfiles:
                    npz = np.load(os.path.join(data_source, sdf_samples_subdir, f))
                    yield npz["pos"], npz["neg"]
            it = scenes()
        else:
            self.npyfiles = [f"<array {i}>" for i in range(len(arrays))]
            it = iter(arrays)
        # Streamed load: one scene at a time goes host -> device (the host never holds more than one scene; a scene handed
        # over as a pair of DEVICE tensors is taken as it is), then ONE device-side concatenation builds the store.
        parts, meta, off = [], [], 0
        for pos, neg in it:
            pos, neg = self._to_device_rows(pos), self._to_device_rows(neg)
            meta.append([off, pos.shape[0], off + pos.shape[0], neg.shape[0]])
            off += pos.shape[0] + neg.shape[0]
            parts += [pos, neg]
        take = [subsample_counts(m[1], m[3], subsample) for m in meta]
        for f, m, (tp, tn) in zip(self.npyfiles, meta, take):
            if tp > m[1] or tn > m[3]:
                raise L.DsdfError(f"{f}: {m[1]} positive + {m[3]} negative rows cannot supply {subsample} samples per scene")
        self.meta_host = np.asarray(meta, dtype=np.int64).reshape(-1, 4)
        self.meta = torch.from_numpy(self.meta_host).to(self.device)
        self.store = torch.cat(parts, 0) if parts else torch.zeros(0, 4, device=self.device)
        del parts
        self.total_rows = off

    def _to_device_rows(self, rows) -> torch.Tensor:
        """One sign's rows of one scene as an fp32 [n,4] device tensor, NaN rows dropped (``remove_nans``, data.py:61-63)."""
        if isinstance(rows, torch.Tensor):
            t = rows.to(self.device, torch.float32)
            return t[~torch.isnan(t[:, 3])].contiguous()
        return torch.from_numpy(remove_nans_np(np.asarray(rows), 3)).to(self.device)

    def __len__(self) -> int:
        return len(self.npyfiles)

    def sample_batch(self, scene_ids: torch.Tensor, seed: int, out: Optional[torch.Tensor] = None) -> torch.Tensor:
        """[S * subsample, 4] scene-major batch of the scenes ``scene_ids`` (int64 device tensor): what the reference's
        DataLoader yields after ``sdf_data.reshape(-1, 4)`` (train_deep_sdf.py:481-483)."""
        L.require_cuda(scene_ids)
        S = scene_ids.numel()
        half2 = 2 * int(self.subsample / 2)
        if half2 != self.subsample:
            raise L.DsdfError("SamplesPerScene must be even (the reference silently yields subsample-1 rows otherwise)")
        out = out if out is not None else torch.empty(S * self.subsample, 4, device=self.device)
        with torch.cuda.device(self.device):
            L.check(L.lib().dsdf_sample_batch(self.store.data_ptr(), self.meta.data_ptr(), scene_ids.data_ptr(), S, self.subsample,
                                              int(seed) & 0xFFFFFFFFFFFFFFFF, out.data_ptr(), L.stream_ptr(self.device)))
        return out


def epoch_batches(num_scenes: int, scenes_per_batch: int, generator: Optional[torch.Generator] = None) -> List[torch.Tensor]:
    """Scene-id batches of one epoch like ``DataLoader(shuffle=True, drop_last=True)`` (train_deep_sdf.py:369-375)."""
    perm = torch.randperm(num_scenes, generator=generator)
    n = num_scenes // scenes_per_batch
    return [perm[i * scenes_per_batch:(i + 1) * scenes_per_batch].sort().values for i in range(n)]
