"""Input normalisation constants for raw uint8 images.

The reference normalises on the host (``transforms.ToTensor()`` + ``transforms.Normalize(mean, std)`` in
its dataset pipeline, /root/reference/distributed.py:112-131) and ships fp32 tensors to the GPU.  Here the
loader may ship the raw bytes instead (``data.build_datasets(..., uint8=True)``): a quarter of the PCIe
traffic and no float conversion on the host cores; the same arithmetic, ``(x / 255 - mean) / std``, then
runs on the GPU, fused into the stem's layout kernel on the native path.
"""
import ctypes

import torch

IMAGENET_MEAN = (0.485, 0.456, 0.406)
IMAGENET_STD = (0.229, 0.224, 0.225)

_mean, _std = IMAGENET_MEAN, IMAGENET_STD
_c_cache = None


def set_input_norm(mean, std):
    """Per-channel mean / std applied to uint8 inputs (defaults: the reference's ImageNet constants)."""
    global _mean, _std, _c_cache
    _mean, _std = tuple(float(m) for m in mean), tuple(float(s) for s in std)
    _c_cache = None


def get_input_norm():
    return _mean, _std


def scale_bias():
    """x_uint8 * scale + bias == (x / 255 - mean) / std."""
    return tuple(1.0 / (255.0 * s) for s in _std), tuple(-m / s for m, s in zip(_mean, _std))


def scale_bias_ctypes():
    global _c_cache
    if _c_cache is None:
        sc, bs = scale_bias()
        _c_cache = ((ctypes.c_float * 3)(*sc), (ctypes.c_float * 3)(*bs))
    return _c_cache


_t_cache = {}


def normalize_u8(x):
    """uint8 [N,3,H,W] -> fp32, the torch formulation (CPU / non-native paths and the test oracle).
    The constants are cached per device: creating them is a host-to-device copy, which a graph capture forbids."""
    sc, bs = scale_bias()
    key = (x.device, sc, bs)
    t = _t_cache.get(key)
    if t is None:
        t = _t_cache[key] = (torch.tensor(sc, dtype=torch.float32, device=x.device).view(1, 3, 1, 1),
                             torch.tensor(bs, dtype=torch.float32, device=x.device).view(1, 3, 1, 1))
    return x.float() * t[0] + t[1]
