"""TEST INFRASTRUCTURE -- nmigen.utils subset (see __init__.py)."""
from ._ast import bits_for  # noqa: F401

__all__ = ["log2_int", "bits_for"]


def log2_int(n, need_pow2=True):
    if n == 0:
        return 0
    r = (n - 1).bit_length()
    if need_pow2 and (1 << r) != n:
        raise ValueError("{} is not a power of 2".format(n))
    return r
