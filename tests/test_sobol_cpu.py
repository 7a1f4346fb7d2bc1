"""CPU suite: the Sobol generator matrices this repo feeds its headless scenes with
(monte-carlo-raytracer_b200/sobol.py) against the reference's own table, and the sample function."""
import os
import re

import numpy as np
import pytest

from mcrt_b200 import sobol, camera

REF_TABLE = "/root/reference/source/application/PathTracer/raytracing/sampling/sobol.h"


@pytest.mark.skipif(not os.path.exists(REF_TABLE), reason="reference tree not present")
def test_generated_matrices_equal_the_reference_table():
    src = open(REF_TABLE, encoding="latin-1").read()
    body = src[src.index("g_SobolMatrices32"):]
    body = body[body.index("{") + 1:body.index("};")]
    ref = np.array([int(x, 16) for x in re.findall(r"0x[0-9a-fA-F]+", body)], dtype=np.uint64).astype(np.uint32)
    assert ref.size == 1024 * sobol.MATRIX_SIZE
    mine = sobol.matrices()
    assert np.array_equal(mine, ref[:mine.size])


def test_dimension_zero_is_the_radical_inverse():
    t = sobol.matrices(2)
    for i, want in [(0, 0.0), (1, 0.5), (2, 0.25), (3, 0.75), (4, 0.125), (255, 255 / 256.0)]:
        assert sobol.sobol_sample(i, 0, 0, t) == np.float32(want)
    # scrambling is an XOR on the 32-bit integer
    assert sobol.sobol_sample(1, 0, 0x80000000, t) == np.float32(0.0)


def test_taa_jitter_uses_the_same_sequence():
    """PathTracingApp::update draws the sub-pixel offset from dimensions 0 / 1 (scramble 0)."""
    t = sobol.matrices(2)
    for f in range(8):
        u = (float(sobol.sobol_sample(f, 0, 0, t)), float(sobol.sobol_sample(f, 1, 0, t)))
        ox, oy = camera.taa_pixel_offset(f)
        assert np.isclose(ox, -2.0 + 4.0 * u[0]) and np.isclose(oy, -2.0 + 4.0 * u[1])
