"""numpy twin of the input generator of oracle/hwprobe/gen_amx_bf16_golden.cpp (counter-based hash), so that
tests/golden/amx_bf16_hw.npz only has to store what the AMX hardware produced.  Test infrastructure."""
import numpy as np

from oracle import oracle as O

N, P, H, J, V = 32, 320, 1024, 512, 32


def _mix(x):
    x = x.astype(np.uint64)
    m = np.uint64(0xFFFFFFFF)
    x ^= x >> np.uint64(16)
    x = (x * np.uint64(0x85EBCA6B)) & m
    x ^= x >> np.uint64(13)
    x = (x * np.uint64(0xC2B2AE35)) & m
    x ^= x >> np.uint64(16)
    return x


def uni(seed, n, bound):
    i = np.arange(n, dtype=np.uint64)
    h = _mix((i * np.uint64(2654435761) + np.uint64(seed) * np.uint64(0x9E3779B9) + np.uint64(12345)) & np.uint64(0xFFFFFFFF))
    u = (h >> np.uint64(8)).astype(np.float32) * np.float32(1.0 / 16777216.0)
    return ((np.float32(2.0) * u - np.float32(1.0)) * np.float32(bound)).astype(np.float32)


def bf16(seed, shape, bound):
    return O.f32_to_bf16_bits(uni(seed, int(np.prod(shape)), bound)).reshape(shape)


def f32(seed, shape, bound):
    return uni(seed, int(np.prod(shape)), bound).reshape(shape)


def inputs():
    return dict(
        x=bf16(1, (N, P), 2.0), h=bf16(2, (N, P), 1.0), c=f32(3, (N, P), 2.0), w_ih=bf16(4, (4 * P, P), 0.12),
        w_hh=bf16(5, (4 * P, P), 0.12), bias=f32(6, (4 * P,), 0.5),
        f=bf16(11, (N, H), 1.0), g=bf16(12, (N, P), 1.0), w_trans=bf16(13, (J, H), 0.1), w_pred=bf16(14, (J, P), 0.1),
        b1=f32(15, (J,), 0.5), w2=bf16(21, (V, J), 0.2), b2=f32(22, (V,), 0.5))
