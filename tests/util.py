import numpy as np


def bits(a):
    """uint32 view with every NaN mapped to one canonical pattern: the sign/payload of a NaN is
    defined neither by WGSL nor by IEEE-754 arithmetic (x86 SSE produces 0xFFC00000, CUDA
    0x7FFFFFFF), everything else must match bit for bit."""
    a = np.ascontiguousarray(a, dtype=np.float32)
    b = a.view(np.uint32).copy()
    b[np.isnan(a)] = 0x7FC00000
    return b


def same_bits(a, b):
    return np.array_equal(bits(a), bits(b))


def count_diff(a, b):
    return int((bits(a) != bits(b)).sum())
