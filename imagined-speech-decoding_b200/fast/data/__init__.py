"""Data loading and preprocessing for FAST (drop-in for the reference's ``fast.data``, src/fast/data/__init__.py:5-30).

Every name the reference exports is kept; the band-pass / resample functions of ``fast.data.filters`` are additive.
"""
from .loaders import BasicDataset, load_standardized_h5
from .preprocess import (
    Electrodes,
    Zones,
    SUBJECTS,
    CLASSES,
    NAME,
    TARGET_TIMEPOINTS,
    load_training_set,
    load_validation_set,
    load_test_set,
    load_test_set_per_subject,
    load_subject_train_val,
    subject_train_val_device,
    trials_from_mat,
)
from .filters import bandpass_sos, sosfiltfilt, resample_poly, sosfiltfilt_resample, bandpass_resample, bandpass_resample_host

__all__ = [
    "BasicDataset",
    "load_standardized_h5",
    "Electrodes",
    "Zones",
    "SUBJECTS",
    "CLASSES",
    "load_training_set",
    "load_validation_set",
    "load_test_set",
    "load_test_set_per_subject",
    "load_subject_train_val",
    # additive (B200 path)
    "NAME",
    "TARGET_TIMEPOINTS",
    "trials_from_mat",
    "subject_train_val_device",
    "bandpass_sos",
    "sosfiltfilt",
    "resample_poly",
    "sosfiltfilt_resample",
    "bandpass_resample",
    "bandpass_resample_host",
]
