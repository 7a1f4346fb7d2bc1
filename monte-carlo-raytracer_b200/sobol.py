"""Sobol generator matrices in the layout the reference uploads to its kernels (sampling/sobol.h:31-34,
scene/RTScene.cpp:55): per dimension 52 columns, each holding the upper 32 bits of a 52-bit direction
number.  The table is an INPUT of the device path (ptb_scene_desc.sobol); the engine passes its own
1024-dimension table, the headless scenes of this repo generate the leading dimensions here from the
Joe-Kuo primitive polynomials / initial direction numbers (new-joe-kuo-6), which is what the
reference's table was generated from -- tests/test_sobol_cpu.py checks the two agree column for column.
"""
import numpy as np

MATRIX_SIZE = 52          # SOBOL_MATRIX_SIZE
RT_SAMPLER_SOBOL, RT_SAMPLER_RANDOM = 0, 1      # samplers.cl:15-17

# (s, a, m_1..m_s) for dimensions 2, 3, ... ; dimension 1 is the van der Corput sequence
JOE_KUO = [
    (1, 0, [1]),
    (2, 1, [1, 3]),
    (3, 1, [1, 3, 1]),
    (3, 2, [1, 1, 1]),
    (4, 1, [1, 1, 3, 3]),
    (4, 4, [1, 3, 5, 13]),
    (5, 2, [1, 1, 5, 5, 17]),
    (5, 4, [1, 1, 5, 5, 5]),
    (5, 7, [1, 1, 7, 11, 19]),
    (5, 11, [1, 1, 5, 1, 1]),
    (5, 13, [1, 1, 1, 3, 11]),
    (5, 14, [1, 3, 5, 5, 31]),
    (6, 1, [1, 3, 3, 9, 7, 49]),
    (6, 13, [1, 1, 1, 15, 21, 21]),
    (6, 16, [1, 3, 1, 13, 27, 49]),
    (6, 19, [1, 1, 1, 15, 7, 5]),
    (6, 22, [1, 3, 1, 15, 13, 25]),
    (6, 25, [1, 1, 5, 5, 19, 61]),
    (7, 1, [1, 3, 7, 11, 23, 15, 103]),
    (7, 4, [1, 3, 7, 13, 13, 15, 69]),
]
MAX_DIMENSIONS = 1 + len(JOE_KUO)


def direction_numbers(dim):
    """52 direction numbers (as 52-bit integers) of 0-based dimension `dim`."""
    L = MATRIX_SIZE
    if dim == 0:
        return [1 << (L - 1 - i) for i in range(L)]
    s, a, m = JOE_KUO[dim - 1]
    v = [0] * L
    for i in range(min(s, L)):
        v[i] = m[i] << (L - 1 - i)
    for i in range(s, L):
        x = v[i - s] ^ (v[i - s] >> s)
        for k in range(1, s):
            if (a >> (s - 1 - k)) & 1:
                x ^= v[i - k]
        v[i] = x
    return v


def matrices(num_dimensions=MAX_DIMENSIONS):
    """uint32[num_dimensions * 52], the reference's g_SobolMatrices32 layout."""
    if not 1 <= num_dimensions <= MAX_DIMENSIONS:
        raise ValueError("this generator covers %d dimensions" % MAX_DIMENSIONS)
    out = np.zeros(num_dimensions * MATRIX_SIZE, np.uint32)
    for d in range(num_dimensions):
        out[d * MATRIX_SIZE:(d + 1) * MATRIX_SIZE] = [x >> (MATRIX_SIZE - 32) for x in direction_numbers(d)]
    return out


def sobol_sample(idx, dimension, scramble, table):
    """samplers.cl:64-72 on the host (uint32 in, float32 out)."""
    v = int(scramble)
    i = dimension * MATRIX_SIZE
    idx = int(idx)
    while idx:
        if idx & 1:
            v ^= int(table[i])
        idx >>= 1
        i += 1
    return np.float32(np.float32(v) * np.float32(2.0 ** -32))
