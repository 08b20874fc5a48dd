"""BCI Competition 2020 Track #3 (imagined speech) montage tables and split loaders.

Drop-in for the reference's ``src/fast/data/preprocess.py``: same constants (:19-42), same loader names, arguments and
return types (:49-190, :197-224).  File parsing (``scipy.io.loadmat`` / ``h5py`` / ``pandas.read_excel``) stays stock
Python; the array math behind it -- (T, C, N) MATLAB layout -> (N, C, 800) float32 with the last sample repeated
(:57-62) -- runs as the ``fastb200_trials_from_mat`` kernel (include/fastb200.h) and there is no NumPy fallback for it.
"""
import os

import numpy as np
import torch

from .. import _native as N

# ---------------------------------------------------------------------------------------------- dataset constants
NAME = 'BCIC2020Track3'
SUBJECTS = [f'{i:02d}' for i in range(1, 16)]
CLASSES = ['hello', 'help-me', 'stop', 'thank-you', 'yes']
TARGET_TIMEPOINTS = 800

# 64-channel montage, in the row order of the recordings
Electrodes = (
    'Fp1 Fp2 F7 F3 Fz F4 F8 FC5 FC1 FC2 FC6 T7 C3 Cz C4 T8 TP9 CP5 CP1 CP2 CP6 TP10 P7 P3 Pz P4 P8 PO9 O1 Oz O2 PO10 '
    'AF7 AF3 AF4 AF8 F5 F1 F2 F6 FT9 FT7 FC3 FC4 FT8 FT10 C5 C1 C2 C6 TP7 CP3 CPz CP4 TP8 P5 P1 P2 P6 PO7 PO3 POz PO4 PO8'
).split()

# functional areas: an exact partition of the 64 electrodes (6+9+6+7+7+10+15+4)
Zones = {
    'Pre-frontal': 'AF7 Fp1 Fp2 AF8 AF3 AF4'.split(),
    'Frontal': 'F7 F5 F3 F1 Fz F2 F4 F6 F8'.split(),
    'Pre-central': 'FC1 FC2 FC3 FC4 FC5 FC6'.split(),
    'Central': 'C1 C2 C3 Cz C4 C5 C6'.split(),
    'Post-central': 'CP1 CP2 CP3 CPz CP4 CP5 CP6'.split(),
    'Temporal': 'T7 T8 FT7 FT8 TP7 TP8 TP9 TP10 FT9 FT10'.split(),
    'Parietal': 'P1 P2 P3 P4 Pz P5 P6 P7 P8 PO3 PO4 PO7 PO8 PO9 PO10'.split(),
    'Occipital': 'O1 O2 Oz POz'.split(),
}


# ---------------------------------------------------------------------------------------------- array math (GPU)
def _device(device=None):
    if device is None:
        if not torch.cuda.is_available():
            raise RuntimeError("fast.data: the trial re-layout runs as a CUDA kernel on a B200; no CUDA device is "
                               "available and there is no CPU fallback")
        device = torch.device('cuda', torch.cuda.current_device())
    return torch.device(device)


def trials_from_mat(x_tcn, target_len=TARGET_TIMEPOINTS, device=None):
    """(T, C, N) float64/float32 array as ``loadmat`` returns it -> (N, C, target_len) float32 CUDA tensor,
    right edge-padded: ``np.transpose(x,(2,1,0)).astype(float32)`` + ``np.pad(..., 'edge')`` (preprocess.py:60-62)."""
    device = _device(device)
    x = np.ascontiguousarray(x_tcn)
    if x.ndim != 3:
        raise ValueError(f"expected a (T, C, N) array, got shape {x.shape}")
    if x.dtype not in (np.float32, np.float64):
        x = x.astype(np.float64)
    T, C, n = x.shape
    T_pad = max(int(target_len), T)
    src = torch.from_numpy(x).to(device, non_blocking=False)
    dst = torch.empty(n, C, T_pad, device=device, dtype=torch.float32)
    with torch.cuda.device(device):
        N.check(N.lib().fastb200_trials_from_mat(N.ptr(src), x.dtype.itemsize, N.ptr(dst), T, C, n, T_pad,
                                                 N.stream_ptr(device)), "trials_from_mat")
    return dst


def _pad_edge_ncT(x_ncT, target_len, device=None):
    """(N, C, T) -> (N, C, target_len) float32, edge padded (test-set layout, preprocess.py:110-114).
    Same kernel: a (N, C, T) array is the (T', C', N') layout of its own transpose, so go through a device permute."""
    device = _device(device)
    x = torch.from_numpy(np.ascontiguousarray(x_ncT)).to(device)
    tcn = x.permute(2, 1, 0).contiguous()
    if tcn.dtype not in (torch.float32, torch.float64):
        tcn = tcn.double()
    T, C, n = tcn.shape
    T_pad = max(int(target_len), T)
    dst = torch.empty(n, C, T_pad, device=device, dtype=torch.float32)
    with torch.cuda.device(device):
        N.check(N.lib().fastb200_trials_from_mat(N.ptr(tcn), tcn.element_size(), N.ptr(dst), T, C, n, T_pad,
                                                 N.stream_ptr(device)), "trials_from_mat")
    return dst


def _labels_from_onehot(y):
    return np.asarray(y).argmax(0).astype(np.uint8)


def _split(mat_path, key):
    import scipy.io
    data = scipy.io.loadmat(mat_path)
    x = np.asarray(data[key]['x'])[0][0]
    y = np.asarray(data[key]['y'])[0][0]
    return trials_from_mat(x).cpu().numpy(), _labels_from_onehot(y)


def _load_split_all(base_folder, folder_name, key, tag):
    X_all, Y_all = [], []
    folder = os.path.join(base_folder, folder_name)
    for SID in SUBJECTS:
        filepath = os.path.join(folder, f'Data_Sample{SID}.mat')
        if os.path.exists(filepath):
            x, y = _split(filepath, key)
            X_all.append(x)
            Y_all.append(y)
            print(f"  {tag} S{SID}: {x.shape}, labels: {np.unique(y, return_counts=True)}")
    return np.concatenate(X_all, axis=0), np.concatenate(Y_all, axis=0)


# ---------------------------------------------------------------------------------------------- split loaders
def load_training_set(base_folder):
    """All subjects of ``Training set`` -> (X float32 (n,64,800), Y uint8 (n,))  (preprocess.py:49-69)."""
    return _load_split_all(base_folder, 'Training set', 'epo_train', 'Train')


def load_validation_set(base_folder):
    """All subjects of ``Validation set``  (preprocess.py:72-92)."""
    return _load_split_all(base_folder, 'Validation set', 'epo_validation', 'Valid')


def _test_subject(filepath, df_labels, i):
    import h5py
    import pandas as pd
    with h5py.File(filepath, 'r') as f:
        if 'epo_test' not in f:
            return None
        x = np.array(f['epo_test']['x'])
    if x.ndim == 3 and x.shape[2] < TARGET_TIMEPOINTS:
        x = _pad_edge_ncT(x, TARGET_TIMEPOINTS).cpu().numpy()
    x = x.astype(np.float32)
    raw = pd.to_numeric(df_labels.iloc[3:53, 2 * (i + 1)], errors='coerce').values   # answer sheet: 1..5
    return x, (raw - 1).astype(np.uint8)


def load_test_set(base_folder, excel_path):
    """All subjects of ``Test set`` (MATLAB v7.3) with labels from the answer sheet  (preprocess.py:95-129)."""
    import pandas as pd
    X_all, Y_all = [], []
    folder = os.path.join(base_folder, 'Test set')
    df_labels = pd.read_excel(excel_path, header=None)
    for i, SID in enumerate(SUBJECTS):
        filepath = os.path.join(folder, f'Data_Sample{SID}.mat')
        if os.path.exists(filepath):
            got = _test_subject(filepath, df_labels, i)
            if got is not None:
                X_all.append(got[0])
                Y_all.append(got[1])
                print(f"  Test  S{SID}: {got[0].shape}, labels: {np.unique(got[1], return_counts=True)}")
    return np.concatenate(X_all, axis=0), np.concatenate(Y_all, axis=0)


def load_test_set_per_subject(base_folder, excel_path):
    """{subject_id: (X, Y)} of the test set  (preprocess.py:132-161)."""
    import pandas as pd
    test_data = {}
    folder = os.path.join(base_folder, 'Test set')
    df_labels = pd.read_excel(excel_path, header=None)
    for i, SID in enumerate(SUBJECTS):
        filepath = os.path.join(folder, f'Data_Sample{SID}.mat')
        if os.path.exists(filepath):
            got = _test_subject(filepath, df_labels, i)
            if got is not None:
                test_data[SID] = got
    return test_data


def load_subject_train_val(base_folder, SID):
    """Train + validation trials of one subject, concatenated  (preprocess.py:164-190)."""
    X_parts, Y_parts = [], []
    for folder_name, key in (('Training set', 'epo_train'), ('Validation set', 'epo_validation')):
        path = os.path.join(base_folder, folder_name, f'Data_Sample{SID}.mat')
        if os.path.exists(path):
            x, y = _split(path, key)
            X_parts.append(x)
            Y_parts.append(y)
    return np.concatenate(X_parts, axis=0), np.concatenate(Y_parts, axis=0)


# ---------------------------------------------------------------------------------------------- HDF5 cache writer
def proc_one(SID, src_folder='./', name=NAME):
    """One subject's train + validation trials  (preprocess.py:197-211)."""
    X, Y = load_subject_train_val(f'{src_folder}/{name}', SID)
    print(SID, X.shape, Y.shape)
    return SID, X, Y


def proc_all(src_folder='./', data_folder='./Processed'):
    """Per-subject groups ``{SID}/X, {SID}/Y`` in one HDF5 file  (preprocess.py:214-224).  Subjects are processed in
    this process (the re-layout is a GPU kernel; the reference's 15-process pool only parallelised the NumPy copy)."""
    import h5py
    os.makedirs(data_folder, exist_ok=True)
    with h5py.File(f'{data_folder}/{NAME}.h5', 'w') as f:
        for SID in SUBJECTS:
            _, X, Y = proc_one(SID, src_folder=src_folder, name=NAME)
            f.create_dataset(f'{SID}/X', data=X)
            f.create_dataset(f'{SID}/Y', data=Y)
            print(SID, X.shape, Y.shape, np.unique(Y, return_counts=True))


if __name__ == '__main__':
    proc_all()
