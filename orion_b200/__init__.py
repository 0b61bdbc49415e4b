"""orion_b200: B200-native CKKS evaluation backend for Orion (ctypes mirror of the reference's backend object)."""
import os as _os

# The dataflow lanes and side streams of liborion_b200 (csrc/capi.cu, csrc/context.cu) want their own hardware queues:
# the CUDA default of 8 connections makes 8 lanes + 16 side streams share channels (measured on the ResNet-20 replay:
# 0.45 -> 0.40 s/image with 32).  Only effective if set before the process creates its CUDA context.
_os.environ.setdefault("CUDA_DEVICE_MAX_CONNECTIONS", "32")
