"""B200-native IQ / BLT-VQG training-step kernels behind the reference's `models.IQ` API."""
from ._lib import B200iqError, LIB_PATH  # noqa: F401
