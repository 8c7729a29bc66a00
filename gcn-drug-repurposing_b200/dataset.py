"""Drop-in mirror of method/dataset.py:5-28 — the boundary between train.py and the model.

The dataset holds the fp32 feature matrix and the adjacency and yields *indices*; the loader
yields shuffled int64 index batches.  Kept as torch.utils.data classes with the reference's
signatures so train.py:125-137 runs unchanged.  `num_workers` is accepted and forced to 0:
the reference forks 6 workers that only echo integers (pure overhead), and forked workers must
not exist next to a CUDA context anyway; the sampler permutation is drawn in the main process
from the torch default generator either way, so batch order for a given seed is identical.
"""
import torch
from torch.utils.data import DataLoader, Dataset


class DiffusionDataSet(Dataset):
    def __init__(self, features, adj):
        self.features = torch.as_tensor(features).to(torch.float32) if isinstance(features, torch.Tensor) \
            else torch.tensor(features, dtype=torch.float32)
        self.adj = adj

    def __len__(self):
        return len(self.features)

    def __getitem__(self, index):
        return index


class DiffusionDataLoader(DataLoader):
    def __init__(self, dataset, batch_size, num_workers, shuffle=False, drop_last=False):
        self.requested_num_workers = num_workers
        super().__init__(dataset, batch_size=batch_size, shuffle=shuffle, num_workers=0, drop_last=drop_last)
