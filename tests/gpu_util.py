"""Helpers shared by the GPU parity tests."""
import numpy as np


def to_dev(x):
    import torch
    x = np.ascontiguousarray(x, np.complex64)
    return torch.from_numpy(x.view(np.float32).reshape(-1, 2)).cuda()


def err_db(got, ref):
    """20*log10(rms(got-ref)/rms(ref)); -inf when identical."""
    got = np.asarray(got).astype(np.complex128 if np.iscomplexobj(got) else np.float64)
    ref = np.asarray(ref).astype(got.dtype)
    assert got.shape == ref.shape, (got.shape, ref.shape)
    if got.size == 0:
        return -np.inf
    num = np.sqrt(np.mean(np.abs(got - ref) ** 2))
    den = np.sqrt(np.mean(np.abs(ref) ** 2))
    if num == 0:
        return -np.inf
    if den == 0:
        return np.inf
    return 20 * np.log10(num / den)


def split(x, sizes):
    out, i, k = [], 0, 0
    while i < len(x):
        n = sizes[k % len(sizes)]
        out.append(x[i:i + n])
        i += n
        k += 1
    return out
