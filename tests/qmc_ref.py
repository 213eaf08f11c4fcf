"""TEST INFRASTRUCTURE: numpy/pure-Python restatement of the reference's QMC helpers, used to pin the oracle KAT dump
(/root/reference/include/utility/util_mcqmc.h:93-162) independently of the C++ build."""
import numpy as np

M = 0xFFFFFFFF


def fnv32a(v):
    h = 0x811c9dc5
    for i in range(4):
        h ^= (v >> (8 * i)) & 0xff
        h = (h * 0x01000193) & M
    return h


def ri_vdc(bits, r=0):
    bits = ((bits << 16) | (bits >> 16)) & M
    bits = (((bits & 0x00ff00ff) << 8) | ((bits & 0xff00ff00) >> 8)) & M
    bits = (((bits & 0x0f0f0f0f) << 4) | ((bits & 0xf0f0f0f0) >> 4)) & M
    bits = (((bits & 0x33333333) << 2) | ((bits & 0xcccccccc) >> 2)) & M
    bits = (((bits & 0x55555555) << 1) | ((bits & 0xaaaaaaaa) >> 1)) & M
    return np.float32(min(1.0, max(0.0, float(np.float32((bits ^ r) * 2.0 ** -32)))))


def ri_s(i, r=0):
    v = 0x80000000
    while i:
        if i & 1:
            r ^= v
        i >>= 1
        v ^= v >> 1
    return np.float32(min(1.0, max(0.0, float(np.float32(r * 2.0 ** -32)))))


def ri_lp(i, r=0):
    v = 0x80000000
    while i:
        if i & 1:
            r ^= v
        i >>= 1
        v |= v >> 1
    return np.float32(min(1.0, max(0.0, float(np.float32(r * 2.0 ** -32)))))
