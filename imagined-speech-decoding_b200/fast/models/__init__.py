"""FAST model family (B200-native drop-in for the reference's ``fast.models``)."""
from .fast import (AttentionBlock, Conv4Layers, CVBlock, EEGNet_Encoder, FAST, Head, HeadConv_Paper_Version)

__all__ = ["FAST", "AttentionBlock", "CVBlock", "Conv4Layers", "EEGNet_Encoder", "HeadConv_Paper_Version", "Head"]
