"""Seeded ray-set generators shared by the CPU and GPU parity tests."""
import numpy as np

from mcrt_b200 import layouts as L


def random_rays(n, lo, hi, seed, maxt=1000.0):
    """The reference conformance test's ray distribution (radeon_rays_conformance_test_cl.h:556-560):
    origins uniform in a box, directions = normalize(rand^3 - 0.5), maxt = 1000."""
    rng = np.random.RandomState(seed)
    rays = L.make_rays(n)
    rays["o"][:, :3] = (lo + (hi - lo) * rng.rand(n, 3)).astype(np.float32)
    d = rng.rand(n, 3) - 0.5
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    rays["d"][:, :3] = d.astype(np.float32)
    rays["o"][:, 3] = maxt
    return rays


def camera_rays(scene, width, height, offset=(0.0, 0.0)):
    """Primary rays exactly as GeneratePerspectiveRays makes them (PathTracing.cl:13-35), in numpy fp32."""
    from mcrt_b200 import camera
    cam = camera.make_camera(scene.camera, width, height, offset)
    f = np.float32
    xs = (np.arange(width, dtype=f) * (f(1.0) / f(width))).astype(f)
    ys = (np.arange(height, dtype=f) * (f(1.0) / f(height))).astype(f)
    u, v = np.meshgrid(xs, ys)
    u, v = u.reshape(-1, 1), v.reshape(-1, 1)
    r00, r10, r11, r01 = (cam[k][0, :3].astype(f) for k in ("r00", "r10", "r11", "r01"))
    a = r00 + (r10 - r00) * u
    b = r01 + (r11 - r01) * u
    d = a + (b - a) * v
    d = d / np.sqrt((d[:, 0:1] * d[:, 0:1] + d[:, 1:2] * d[:, 1:2]) + d[:, 2:3] * d[:, 2:3])
    rays = L.make_rays(width * height)
    rays["o"][:, :3] = cam["pos"][0, :3]
    rays["o"][:, 3] = 1000.0
    rays["d"][:, :3] = d.astype(f)
    return rays, cam
