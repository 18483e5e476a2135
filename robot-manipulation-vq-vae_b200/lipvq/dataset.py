"""Action-chunk input side: SequenceDataset's window / pad / normalise semantics over demonstrations
held in HBM.

The reference samples one sequence at a time on the host (robomimic/utils/dataset.py:499-632: ``get_item`` ->
``get_sequence_from_demo`` -> ``TensorUtils.pad_sequence(pad_same=True)`` -> ``ObsUtils.normalize_dict``) and a DataLoader
collates them.  Here the demos' action arrays live on the device as one ``[L_total, A]`` tensor, the index map of
dataset.py:218-260 is three small device arrays, and a whole ``[B, T, A]`` batch (+ pad mask) is one kernel launch
(``lipvq_window_gather_f32``).  Host-side by design: the index map construction and the sampling of sequence indices.
HDF5 reading is out of scope (no h5py / data here): ``from_arrays`` takes the per-demo ``actions [L_i, A]`` arrays the
robomimic file stores under ``data/demo_i/actions``.
"""
from __future__ import annotations

from typing import Dict, List, Optional, Sequence, Tuple

import numpy as np
import torch

from . import ops


class ActionSequenceDataset:
    def __init__(self, demos: Sequence[np.ndarray], seq_length: int = 1, frame_stack: int = 1, pad_seq_length: bool = True,
                 pad_frame_stack: bool = True, action_normalization_stats: Optional[Dict[str, np.ndarray]] = None,
                 device: str = "cuda"):
        assert frame_stack >= 1 and seq_length >= 1
        self.seq_length, self.n_frame_stack = seq_length, frame_stack
        self.pad_seq_length, self.pad_frame_stack = pad_seq_length, pad_frame_stack
        lens = [int(d.shape[0]) for d in demos]
        A = int(demos[0].shape[1])
        # index map (dataset.py:229-260): one sequence per valid start frame
        self.demo_seq0: List[int] = []
        self.index_to_demo: List[int] = []
        total = 0
        for i, L in enumerate(lens):
            n = L
            if not pad_frame_stack:
                n -= frame_stack - 1
            if not pad_seq_length:
                n -= seq_length - 1
            if pad_seq_length:
                assert L >= 1
                n = max(n, 1)
            else:
                assert n >= 1, "demo shorter than frame_stack - 1 + seq_length"
            self.demo_seq0.append(total)
            self.index_to_demo += [i] * n
            total += n
        self.total_num_sequences = total
        row0 = np.concatenate([[0], np.cumsum(lens)[:-1]]).astype(np.int64)
        dev = torch.device(device)
        self.data = torch.from_numpy(np.ascontiguousarray(np.concatenate(demos, 0), dtype=np.float32)).to(dev)
        self.demo_row0 = torch.from_numpy(row0).to(dev)
        self.demo_len = torch.tensor(lens, dtype=torch.int32, device=dev)
        self.demo_seq0_dev = torch.tensor(self.demo_seq0, dtype=torch.int64, device=dev)
        self._index_to_demo_t = torch.tensor(self.index_to_demo, dtype=torch.int32)
        self.offset = self.scale = None
        if action_normalization_stats is not None:  # {"offset": [1, A], "scale": [1, A]} (obs_utils.py:464-501)
            self.offset = torch.as_tensor(np.asarray(action_normalization_stats["offset"], dtype=np.float32).reshape(A)).to(dev)
            self.scale = torch.as_tensor(np.asarray(action_normalization_stats["scale"], dtype=np.float32).reshape(A)).to(dev)

    from_arrays = classmethod(lambda cls, demos, **kw: cls(demos, **kw))

    def __len__(self) -> int:
        return self.total_num_sequences

    def get_batch(self, indices, want_mask: bool = False):
        """indices: sequence indices (as ``__getitem__`` takes) -> actions [B, T, A] (+ pad_mask [B, T] bool)."""
        idx = torch.as_tensor(indices, dtype=torch.int64)
        demo = self._index_to_demo_t[idx]
        dev = self.data.device
        return ops.window_gather(self.data, self.demo_row0, self.demo_len, self.demo_seq0_dev, demo.to(dev), idx.to(dev),
                                 self.n_frame_stack, self.seq_length, self.pad_frame_stack, self.offset, self.scale, want_mask)
