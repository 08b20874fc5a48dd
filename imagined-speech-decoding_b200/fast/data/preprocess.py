"""BCI Competition 2020 Track #3 (imagined speech) montage tables and split loaders.

Drop-in for the reference's ``src/fast/data/preprocess.py``: same constants (:19-42), same loader names, arguments and
return types (:49-190, :197-224).  File parsing (``scipy.io.loadmat`` / ``h5py`` / ``pandas.read_excel``) stays stock
Python; the array math behind it -- (T, C, N) MATLAB layout -> (N, C, 800) float32 with the last sample repeated
(:57-62) -- runs as the ``fastb200_trials_from_mat`` kernel (include/fastb200.h) and there is no NumPy fallback for it.

GPU-direct ingestion (additive): ``subject_train_val_device`` / ``_split_device`` keep the trials on the device (file ->
pinned staging -> re-layout kernel -> optional band-pass / resample kernel -> CUDA tensor) for ``fast.train.cv``; the
reference-named ``load_*`` functions are NumPy views of them.
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


_STAGE = {}      # device -> [pinned uint8 staging buffer, event recorded after the last copy out of it]


def _to_device(x_np, device):
    """Pageable NumPy array -> device tensor through a reusable PINNED staging buffer and one asynchronous copy
    (``torch.from_numpy(x).to(device)`` on pageable memory is a synchronous, driver-staged copy).  The staging buffer is
    reused by the next call only after its copy has completed (event), so callers never wait on the H2D themselves."""
    x_np = np.ascontiguousarray(x_np)
    slot = _STAGE.get(device)
    if slot is None or slot[0].numel() < x_np.nbytes:
        slot = [torch.empty(max(x_np.nbytes, 1 << 20), dtype=torch.uint8, pin_memory=True), None]
        _STAGE[device] = slot
    if slot[1] is not None:
        slot[1].synchronize()
    view = slot[0][:x_np.nbytes].view(torch.from_numpy(x_np).dtype).view(x_np.shape)
    view.copy_(torch.from_numpy(x_np))
    with torch.cuda.device(device):
        out = view.to(device, non_blocking=True)
        slot[1] = torch.cuda.Event()
        slot[1].record(torch.cuda.current_stream(device))
    return out


def trials_from_mat(x_tcn, target_len=TARGET_TIMEPOINTS, device=None, pad=None):
    """(T, C, N) float64/float32 array as ``loadmat`` returns it -> (N, C, T_pad) float32 CUDA tensor, right edge-padded:
    ``np.transpose(x,(2,1,0)).astype(float32)`` + ``np.pad(..., 'edge')`` (preprocess.py:60-62) as ONE kernel over a
    pinned-staged copy of the file's array.  ``pad=k`` appends exactly k samples (the train / validation loaders' fixed
    ``(0, 5)``, preprocess.py:62,85,177,186,209); otherwise the trial is padded up to ``target_len`` (test-set rule,
    preprocess.py:110-113)."""
    device = _device(device)
    x = np.asarray(x_tcn)
    if x.ndim != 3:
        raise ValueError(f"expected a (T, C, N) array, got shape {x.shape}")
    if x.dtype not in (np.float32, np.float64):
        x = x.astype(np.float64)
    T, C, n = x.shape
    T_pad = T + int(pad) if pad is not None else max(int(target_len), T)
    src = _to_device(x, device)
    dst = torch.empty(n, C, T_pad, device=device, dtype=torch.float32)
    with torch.cuda.device(device):
        N.check(N.lib().fastb200_trials_from_mat(N.ptr(src), x.dtype.itemsize, N.ptr(dst), T, C, n, T_pad,
                                                 N.stream_ptr(device)), "trials_from_mat")
    return dst


def _pad_edge_ncT(x_ncT, target_len, device=None):
    """(N, C, T) -> (N, C, target_len) float32, edge padded (test-set layout, preprocess.py:110-114).
    Same kernel: a (N, C, T) array is the (T', C', N') layout of its own transpose, so go through a device permute."""
    device = _device(device)
    x = _to_device(x_ncT, device)
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


def _split_device(mat_path, key, device=None, bandpass=None):
    """One split of one subject, device resident: ``loadmat`` (stock parser) -> pinned staging -> fused transpose + cast +
    edge-pad kernel (-> optional zero-phase band-pass / resample kernel) -> (X (n, C, T') float32 CUDA, Y (n,) uint8 CUDA).
    Nothing comes back to the host.  ``bandpass``: None, or keyword arguments of ``fast.data.filters.bandpass_resample``."""
    import scipy.io
    data = scipy.io.loadmat(mat_path)
    x = np.asarray(data[key]['x'])[0][0]
    y = np.asarray(data[key]['y'])[0][0]
    X = trials_from_mat(x, device=device, pad=5)
    if bandpass is not None:
        from .filters import bandpass_resample
        X = bandpass_resample(X, **dict(bandpass))
    Y = torch.from_numpy(_labels_from_onehot(y)).to(X.device, non_blocking=True)
    return X, Y


def _split(mat_path, key):
    X, Y = _split_device(mat_path, key)
    return X.cpu().numpy(), Y.cpu().numpy()


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


def subject_train_val_device(base_folder, SID, device=None, bandpass=None):
    """GPU-direct form of ``load_subject_train_val``: the subject's train + validation trials as CUDA
    tensors (X (n, 64, 800) float32, Y (n,) uint8) that ``fast.train.cv.run_cv_jobs`` / ``train_fold`` consume in place --
    file -> pinned staging -> one re-layout kernel per split (+ the band-pass / resample kernel when ``bandpass`` is
    given) -> concatenation on the device; no host round trip of the trial data."""
    X_parts, Y_parts = [], []
    for folder_name, key in (('Training set', 'epo_train'), ('Validation set', 'epo_validation')):
        path = os.path.join(base_folder, folder_name, f'Data_Sample{SID}.mat')
        if os.path.exists(path):
            x, y = _split_device(path, key, device=device, bandpass=bandpass)
            X_parts.append(x)
            Y_parts.append(y)
    if not X_parts:
        raise ValueError("need at least one array to concatenate")          # np.concatenate's message on an empty list
    return torch.cat(X_parts, dim=0), torch.cat(Y_parts, dim=0)


def load_subject_train_val(base_folder, SID):
    """Train + validation trials of one subject, concatenated  (preprocess.py:164-190).  NumPy view of
    ``subject_train_val_device`` for the unchanged scripts."""
    X, Y = subject_train_val_device(base_folder, SID)
    return X.cpu().numpy(), Y.cpu().numpy()


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
