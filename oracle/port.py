"""oracle/port.py -- TEST INFRASTRUCTURE.  Python driver of the CPU restatement
(oracle/port/*.cpp -> oracle/_build/libmcrt_oracle.so).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg may import this module.
"""
import ctypes
import os
import subprocess

import numpy as np

import mcrt_b200  # noqa: F401
from mcrt_b200 import layouts as L

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB = os.path.join(_HERE, "_build", "libmcrt_oracle.so")

TRI = np.dtype([("v0", np.float32, 3), ("shape", np.int32), ("v1", np.float32, 3), ("prim", np.int32),
                ("v2", np.float32, 3), ("pad", np.int32)])
assert TRI.itemsize == 48

_lib = None


def build():
    subprocess.check_call(["make", "-C", _HERE, "port"], stdout=subprocess.DEVNULL)


def load():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB):
            build()
        lib = ctypes.CDLL(LIB)
        vp, ll, i = ctypes.c_void_p, ctypes.c_longlong, ctypes.c_int
        lib.oracle_make_triangles.argtypes = [vp, i, vp, i, vp, i, vp]
        lib.oracle_intersect.argtypes = [vp, ll, vp, ll, vp, vp]
        lib.oracle_occluded.argtypes = [vp, ll, vp, ll, vp]
        lib.oracle_hit_in_tied_set.argtypes = [vp, ll, vp, ctypes.c_float, i, i]
        lib.oracle_hit_in_tied_set.restype = i
        lib.oracle_wang_hash.argtypes = [ctypes.c_uint32]
        lib.oracle_wang_hash.restype = ctypes.c_uint32
        lib.oracle_sampler_seed.argtypes = [ctypes.c_uint32, i, i, i, i]
        lib.oracle_sampler_seed.restype = ctypes.c_uint32
        lib.oracle_rand_float.argtypes = [ctypes.POINTER(ctypes.c_uint32)]
        lib.oracle_rand_float.restype = ctypes.c_float
        _lib = lib
    return _lib


def _p(a):
    return None if a is None else a.ctypes.data_as(ctypes.c_void_p)


def scene_triangles(scene) -> np.ndarray:
    """World-space triangle records of every shape (id = index into the RTShape array)."""
    return mesh_triangles([(m.positions, m.indices, i, m.transform) for i, m in enumerate(scene.meshes)])


def mesh_triangles(meshes) -> np.ndarray:
    lib = load()
    chunks = []
    for pos, idx, sid, xf in meshes:
        pos = np.ascontiguousarray(pos, np.float32)
        idx = np.ascontiguousarray(idx, np.int32)
        m = np.eye(4, dtype=np.float32) if xf is None else np.ascontiguousarray(xf, np.float32)
        out = np.zeros(idx.shape[0], dtype=TRI)
        lib.oracle_make_triangles(_p(pos), 12, _p(idx), idx.shape[0], _p(m), int(sid), _p(out))
        chunks.append(out)
    return np.concatenate(chunks)


def intersect(tris, rays, hits=None, want_ties=False):
    lib = load()
    if hits is None:
        hits = np.zeros(rays.shape[0], dtype=L.INTERSECTION)
        hits["shapeid"] = -1
        hits["primid"] = -1
    ties = np.zeros(rays.shape[0], np.int32)
    lib.oracle_intersect(_p(tris), tris.shape[0], _p(rays), rays.shape[0], _p(hits), _p(ties))
    return (hits, ties) if want_ties else hits


def occluded(tris, rays, out=None):
    lib = load()
    if out is None:
        out = np.zeros(rays.shape[0], np.int32)
    lib.oracle_occluded(_p(tris), tris.shape[0], _p(rays), rays.shape[0], _p(out))
    return out


SHAPE2L = np.dtype([("minv", np.float32, 12), ("first_tri", np.int32), ("num_tris", np.int32), ("id", np.int32),
                    ("pad", np.int32)])


def two_level_world(shapes):
    """shapes: iterable of (mesh_key, positions, indices, shape_id, minv 4x4 or None) -- one entry per ATTACHED shape;
    entries with the same mesh_key share one set of object-space triangles (instances).  Returns (tris, table)."""
    chunks, first, table, at = [], {}, [], 0
    for key, pos, idx, sid, minv in shapes:
        if key not in first:
            t = mesh_triangles([(pos, idx, 0, None)])
            first[key] = (at, t.shape[0])
            chunks.append(t)
            at += t.shape[0]
        rec = np.zeros(1, dtype=SHAPE2L)
        m = np.eye(4, dtype=np.float32) if minv is None else np.ascontiguousarray(minv, np.float32)
        rec["minv"][0] = m[:3].reshape(-1)
        rec["first_tri"], rec["num_tris"] = first[key]
        rec["id"] = sid
        table.append(rec)
    return np.concatenate(chunks), np.concatenate(table)


def intersect_2level(tris, table, rays, hits=None):
    """Brute-force restatement of the reference's two-level kernel (intersect_bvh2level_skiplinks.cl:135-300)."""
    lib = load()
    lib.oracle_intersect_2level.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p,
                                            ctypes.c_longlong, ctypes.c_void_p]
    if hits is None:
        hits = np.zeros(rays.shape[0], dtype=L.INTERSECTION)
        hits["shapeid"] = -1
        hits["primid"] = -1
    lib.oracle_intersect_2level(_p(tris), _p(table), table.shape[0], _p(rays), rays.shape[0], _p(hits))
    return hits


def occluded_2level(tris, table, rays, out=None):
    lib = load()
    lib.oracle_occluded_2level.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p,
                                           ctypes.c_longlong, ctypes.c_void_p]
    if out is None:
        out = np.zeros(rays.shape[0], np.int32)
    lib.oracle_occluded_2level(_p(tris), _p(table), table.shape[0], _p(rays), rays.shape[0], _p(out))
    return out


def hit_in_tied_set(tris, ray, t, shape, prim) -> bool:
    lib = load()
    r = np.ascontiguousarray(ray)
    return bool(lib.oracle_hit_in_tied_set(_p(tris), tris.shape[0], _p(r), float(t), int(shape), int(prim)))


def compare_hits(test_hits, oracle_hits, tris, rays, ties=None):
    """Classify a CUDA (or reference-kernel) closest-hit result against the brute-force oracle.

    Returns a dict: exact (ids and t bit-equal), tie_ok (t bit-equal, ids differ, test winner is in
    the tied set -- SURVEY Appendix A, E1), b_farther (test found a hit that `oracle_hits` missed or
    placed farther, E4), bad (everything else), and the indices of each class."""
    a, b = test_hits, oracle_hits
    active = rays["extra"][:, 1] != 0
    same_id = (a["shapeid"] == b["shapeid"]) & (a["primid"] == b["primid"])
    hit = b["shapeid"] >= 0
    t_equal = a["uvwt"][:, 3].view(np.uint32) == b["uvwt"][:, 3].view(np.uint32)
    uv_equal = (a["uvwt"][:, :3].view(np.uint32) == b["uvwt"][:, :3].view(np.uint32)).all(axis=1)
    exact = active & same_id & (~hit | (t_equal & uv_equal))
    cand = np.nonzero(active & ~exact)[0]
    tie_ok, bad, b_farther = [], [], []
    for i in cand:
        if hit[i] and a["shapeid"][i] >= 0 and t_equal[i] and hit_in_tied_set(
                tris, rays[i:i + 1], a["uvwt"][i, 3], a["shapeid"][i], a["primid"][i]):
            tie_ok.append(int(i))
        elif a["shapeid"][i] >= 0 and (not hit[i] or b["uvwt"][i, 3] > a["uvwt"][i, 3]):
            # `oracle_hits` missed, or found only a farther hit than `test_hits`: when the second
            # argument is the reference's BVH kernel this is its slab-test culling a leaf whose
            # triangle test would have accepted -- never the brute force.
            b_farther.append(int(i))
        else:
            bad.append(int(i))
    return {"active": int(active.sum()), "exact": int(exact.sum()), "tie_ok": len(tie_ok), "bad": len(bad),
            "b_farther": len(b_farther), "bad_idx": bad, "tie_idx": tie_ok, "b_farther_idx": b_farther}
