"""Host-side camera: restates what the reference's host does per frame to fill RTPinholeCamera.

Follows RTPrimaryRaysPass::generatePrimaryRays (renderPasses/RTPrimaryRaysPass.cpp:81-107),
RTUtil::screenToRay (util/RTUtil.cpp:9-41), CameraComponent (engine/camera/CameraComponent.cpp:61-165;
glm::perspective is the *left-handed* variant because the vendored glm.hpp:56 forces
GLM_FORCE_LEFT_HANDED) and the TAA jitter of PathTracingApp::update (PathTracingApp.cpp:208-215,
sampling/sampling.h:7-15).

The camera struct is an INPUT of the hot path: oracle and CUDA consume the same 176 bytes, so the
closed-form unprojection used here (instead of glm's numeric 4x4 inverse) does not enter parity.
"""
import math

import numpy as np

from . import layouts as L

f4 = np.float32


def euler_yxz_matrix(euler_deg):
    """Rotation matrix (columns = right, up, forward) of math::eulerYXZQuat: R = Ry * Rx * Rz."""
    ex, ey, ez = (math.radians(float(a)) for a in euler_deg)
    cx, sx, cy, sy, cz, sz = math.cos(ex), math.sin(ex), math.cos(ey), math.sin(ey), math.cos(ez), math.sin(ez)
    rx = np.array([[1, 0, 0], [0, cx, -sx], [0, sx, cx]], np.float64)
    ry = np.array([[cy, 0, sy], [0, 1, 0], [-sy, 0, cy]], np.float64)
    rz = np.array([[cz, -sz, 0], [sz, cz, 0], [0, 0, 1]], np.float64)
    return ry @ rx @ rz


def sobol_sample_dim01(index: int, dim: int) -> float:
    """Sampler::sobolSample(frame, dim, scramble 0) for dimensions 0 and 1 (sampling.h:7-15).

    Dimension 0 of the Joe-Kuo/pbrt table is the van der Corput radical inverse; dimension 1 has
    direction numbers v_j = v_{j-1} ^ (v_{j-1} >> 1) starting from 1<<31.  Only these two
    dimensions are consumed on the host (TAA jitter x / y)."""
    v = 0
    if dim == 0:
        c = 1 << 31
        i = index
        while i:
            if i & 1:
                v ^= c
            i >>= 1
            c >>= 1
    else:
        c = 1 << 31
        i = index
        while i:
            if i & 1:
                v ^= c
            i >>= 1
            c ^= c >> 1
    return float(np.float32(v) * np.float32(2.0 ** -32))


def taa_pixel_offset(frame: int, radius=(2.0, 2.0), use_taa=True):
    """curPixelOffset = lerp(-r, r, sobol(frame, dim)) (PathTracingApp.cpp:208-215)."""
    if not use_taa:
        return (0.0, 0.0)
    ux, uy = sobol_sample_dim01(frame, 0), sobol_sample_dim01(frame, 1)
    return (-radius[0] + 2.0 * radius[0] * ux, -radius[1] + 2.0 * radius[1] * uy)


def make_camera(spec, width: int, height: int, pixel_offset=(0.0, 0.0)) -> np.ndarray:
    """Fill one RTPinholeCamera (kernel_data.h:246-264) for the given jitter."""
    R = euler_yxz_matrix(spec.euler_deg)
    right, up, fwd = R[:, 0], R[:, 1], R[:, 2]
    pos = np.asarray(spec.position, np.float64)
    aspect = float(width) / float(height)
    th = math.tan(spec.fov_y * 0.5)          # fov_y used as radians, exactly like the reference
    ox, oy = pixel_offset[0] / width, pixel_offset[1] / height   # RTUtil.cpp:26-30 (NDC translation)

    def corner(sx, sy):
        nx, ny = sx / width * 2.0 - 1.0 - ox, sy / height * 2.0 - 1.0 - oy
        d = right * (nx * aspect * th) + up * (ny * th) + fwd
        return d / np.linalg.norm(d)

    cam = np.zeros(1, dtype=L.CAMERA)
    cam["r00"][0, :3] = corner(0, 0)
    cam["r10"][0, :3] = corner(width, 0)
    cam["r11"][0, :3] = corner(width, height)
    cam["r01"][0, :3] = corner(0, height)
    cam["pos"][0, :3] = pos
    cam["direction"][0, :3] = fwd
    cam["width"] = width
    cam["height"] = height
    # BDPT extras (RTBDPTPass.cpp:158-166): image-plane area at unit distance and world->clip
    # (row-major, so that transformVector4 = M * v, cameras.cl:20).
    cam["area"] = (2.0 * th * aspect) * (2.0 * th)
    n, f = spec.near, spec.far
    proj = np.array([[1.0 / (aspect * th), 0, 0, 0], [0, 1.0 / th, 0, 0],
                     [0, 0, (f + n) / (f - n), -(2 * f * n) / (f - n)], [0, 0, 1, 0]], np.float64)
    view = np.eye(4)
    view[0, :3], view[1, :3], view[2, :3] = right, up, fwd
    view[:3, 3] = -(view[:3, :3] @ pos)
    cam["worldToClip"][0] = (proj @ view).astype(f4)
    return cam
