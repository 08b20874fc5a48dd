"""Dataset class and HDF5 cache reader (reference src/fast/data/loaders.py:11-45)."""
import numpy as np
import torch
from torch.utils.data import Dataset


class BasicDataset(Dataset):
    """(data, label) arrays -> indexable dataset.  4-D input (subjects, trials, C, T) is flattened over subjects
    (loaders.py:14-18).  ``torch.from_numpy`` keeps dtypes: float32 trials, uint8 labels."""

    def __init__(self, data, label):
        if data.ndim == 4:
            data, label = np.concatenate(data, axis=0), np.concatenate(label, axis=0)
        self.data = torch.from_numpy(data)
        self.labels = torch.from_numpy(label)

    def __len__(self):
        return len(self.data)

    def __getitem__(self, idx):
        return self.data[idx], self.labels[idx]


def load_standardized_h5(cache_fn):
    """Per-subject-group HDF5 cache ``{SID}/X, {SID}/Y`` -> X (S, n, C, T), Y (S, n)  (loaders.py:27-45)."""
    import h5py                     # imported here so the constants / dataset class work without h5py installed
    X, Y = [], []
    with h5py.File(cache_fn, 'r') as f:
        for subject in list(f.keys()):
            X.append(f[subject]['X'][()])
            Y.append(f[subject]['Y'][()])
    X, Y = np.array(X), np.array(Y)
    print('Loaded from', cache_fn, X.shape, Y.shape)
    return X, Y
