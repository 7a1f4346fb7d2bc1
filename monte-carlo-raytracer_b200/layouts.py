"""Binary layouts of the hot path's host/device structs as numpy dtypes.

These mirror, byte for byte, the structs the reference shares between host and kernels
(assets/kernels/kernel_data.h:36-272) and Radeon Rays' query structs
(RadeonRays/include/math/ray.h:103-107, radeon_rays.h:140-154).  Sizes are asserted at import
and cross-checked against the compiled reference in tests/test_oracle_ref.py.
"""
import numpy as np

f4 = np.float32
i4 = np.int32
u4 = np.uint32

# RadeonRays::ray / RTRay -- 48 B
RAY = np.dtype([("o", f4, 4), ("d", f4, 4), ("extra", i4, 2), ("doBackfaceCulling", i4), ("padding", i4)])
# RadeonRays::Intersection / RTIntersection -- 32 B
INTERSECTION = np.dtype([("shapeid", i4), ("primid", i4), ("pad", i4, 2), ("uvwt", f4, 4)])
# RTShape -- 160 B (kernel_data.h:36-52); matrices row-major
SHAPE = np.dtype([("toWorldTransform", f4, (4, 4)), ("toWorldInverseTranspose", f4, (4, 4)),
                  ("startIdx", u4), ("startVertex", u4), ("numTriangles", u4), ("materialId", i4),
                  ("lightID", i4), ("area", f4), ("pad", i4, 2)])
# RTMaterial -- 128 B (kernel_data.h:87-113)
MATERIAL = np.dtype([("uber_kd", f4, 4), ("uber_ks", f4, 4), ("uber_kr", f4, 4), ("uber_kt", f4, 4),
                     ("uber_opacity", f4, 4), ("uber_roughness", f4, 2), ("uber_eta", f4), ("type", i4),
                     ("uber_normalMapId", i4), ("uber_diffuseTexId", i4), ("uber_glossyTexId", i4),
                     ("uber_specReflectionTexId", i4), ("uber_transmissionTexId", i4),
                     ("uber_opacityTexId", i4), ("uber_roughnessTexId", i4), ("uber_iorTexId", i4)])
# RTLight -- 80 B (kernel_data.h:137-152)
LIGHT = np.dtype([("d", f4, 4), ("p", f4, 4), ("intensity", f4, 4), ("radius", f4), ("area", f4),
                  ("choicePdf", f4), ("shapeId", i4), ("type", i4), ("flags", i4), ("pad", i4, 2)])
# RTThroughput -- 32 B (kernel_data.h:154-160)
THROUGHPUT = np.dtype([("throughput", f4, 4), ("prevBsdfFlags", i4), ("ignoreOcclusion", i4), ("pad", i4, 2)])
# RTPinholeCamera -- 176 B (kernel_data.h:246-264)
CAMERA = np.dtype([("worldToClip", f4, (4, 4)), ("r00", f4, 4), ("r10", f4, 4), ("r11", f4, 4), ("r01", f4, 4),
                   ("pos", f4, 4), ("direction", f4, 4), ("width", u4), ("height", u4), ("area", f4),
                   ("padding", i4)])
# TextureDesc2D -- 16 B (kernel_data.h:317-326)
TEXDESC = np.dtype([("width", np.uint16), ("height", np.uint16), ("numMipLevels", np.uint16),
                    ("format", np.uint16), ("wrap", np.uint16), ("pad", np.uint16), ("memOffset", u4)])
# RTRayDifferentials -- 64 B (kernel_data.h:266-272)
RAYDIFF = np.dtype([("xOrigin", f4, 4), ("yOrigin", f4, 4), ("xDirection", f4, 4), ("yDirection", f4, 4)])

assert RAY.itemsize == 48 and INTERSECTION.itemsize == 32 and SHAPE.itemsize == 160
assert MATERIAL.itemsize == 128 and LIGHT.itemsize == 80 and THROUGHPUT.itemsize == 32
assert CAMERA.itemsize == 176 and TEXDESC.itemsize == 16 and RAYDIFF.itemsize == 64

# enums (kernel_data.h:54-61,115-135)
RT_DIRECTIONAL_LIGHT, RT_POINT_LIGHT, RT_DISK_AREA_LIGHT, RT_TRIANGLE_MESH_AREA_LIGHT = 0, 1, 2, 3
RT_LIGHT_FLAG_DELTA_POSITION = 1 << 1
RT_LIGHT_FLAG_DELTA_DIRECTION = 1 << 2
RT_LIGHT_FLAG_AREA = 1 << 3
RT_LIGHT_FLAG_INFINITE = 1 << 4
RT_INVALID_ID = -1
TEX_FORMAT_RGBA8 = 3
TEX_WRAP_REPEAT = 0
RT_BOX_FILTER, RT_TRIANGLE_FILTER, RT_GAUSSIAN_FILTER, RT_MITCHELL_FILTER, RT_LANCZOS_SINC_FILTER = range(5)


def make_rays(n: int) -> np.ndarray:
    """n default-constructed RadeonRays::ray (ray.h:33-45): mask -1, active, maxt = FLT_MAX."""
    r = np.zeros(n, dtype=RAY)
    r["o"][:, 3] = np.finfo(np.float32).max
    r["extra"][:, 0] = -1
    r["extra"][:, 1] = 1
    return r


def default_material(n: int = 1) -> np.ndarray:
    """RTMaterial() defaults (kernel_data.h:90-96)."""
    m = np.zeros(n, dtype=MATERIAL)
    m["uber_kd"][:, :3] = 0.25
    m["uber_ks"][:, :3] = 0.25
    m["uber_opacity"][:, :3] = 1.0
    m["uber_roughness"] = 0.1
    m["uber_eta"] = 1.5
    for k in ("uber_normalMapId", "uber_diffuseTexId", "uber_glossyTexId", "uber_specReflectionTexId",
              "uber_transmissionTexId", "uber_opacityTexId", "uber_roughnessTexId", "uber_iorTexId"):
        m[k] = RT_INVALID_ID
    return m


def default_shape(n: int = 1) -> np.ndarray:
    s = np.zeros(n, dtype=SHAPE)
    s["toWorldTransform"] = np.eye(4, dtype=f4)
    s["toWorldInverseTranspose"] = np.eye(4, dtype=f4)
    s["materialId"] = RT_INVALID_ID
    s["lightID"] = RT_INVALID_ID
    s["area"] = 1.0
    return s
