"""Scalar Python emulation of the wide-BVH walk (csrc/traverse.cuh) used by the CPU test-suite to
validate the HOST builder's output (node encoding, slot order, quantised boxes) without a GPU.
It restates the kernel's control flow on the host arrays; the triangle test is delegated to the
brute-force oracle's formulation through `tri_test`."""
import numpy as np

f32 = np.float32


def _safe_inv(d):
    d = f32(d)
    if abs(d) <= f32(1e-20):
        d = f32(np.copysign(f32(1e-20), d))
    return f32(1.0) / d


def decode_child_boxes(node):
    """(lo[8,3], hi[8,3], meta[8], imask, child_base, tri_base) of one 20-word node."""
    p = node[0:3].view(np.float32)
    e = [(int(node[3]) >> (8 * a)) & 0xFF for a in range(3)]
    imask = (int(node[3]) >> 24) & 0xFF
    scale = [np.array([ex << 23], np.uint32).view(np.float32)[0] for ex in e]
    b = node[6:20].view(np.uint8)
    meta = b[0:8]
    q = b[8:].reshape(6, 8)
    lo = np.zeros((8, 3), np.float64)
    hi = np.zeros((8, 3), np.float64)
    for a in range(3):
        lo[:, a] = float(p[a]) + q[a].astype(np.float64) * float(scale[a])
        hi[:, a] = float(p[a]) + q[3 + a].astype(np.float64) * float(scale[a])
    return lo, hi, meta, imask, int(node[4]), int(node[5])


def walk_tree(nodes, tris):
    """Yields (node_index, slot, is_leaf, lo, hi, tri_indices or child_index) for every child."""
    stack = [0]
    while stack:
        ni = stack.pop()
        lo, hi, meta, imask, cbase, tbase = decode_child_boxes(nodes[ni])
        for s in range(8):
            m = int(meta[s])
            if m == 0:
                continue
            if (imask >> s) & 1:
                assert m == ((1 << 5) | (24 + s)), "internal child meta"
                ci = cbase + bin(imask & ((1 << s) - 1)).count("1")
                yield ni, s, False, lo[s], hi[s], ci
                stack.append(ci)
            else:
                cnt = {1: 1, 3: 2, 7: 3}[m >> 5]
                off = m & 31
                assert off + cnt <= 24
                yield ni, s, True, lo[s], hi[s], list(range(tbase + off, tbase + off + cnt))


def subtree_bounds(nodes, tris):
    """Exact bounds of every node's subtree from the triangle records: dict node -> (lo, hi)."""
    children = {}
    for ni, s, leaf, lo, hi, ref in walk_tree(nodes, tris):
        children.setdefault(ni, []).append((s, leaf, lo, hi, ref))
    bounds = {}

    def rec(ni):
        blo, bhi = np.full(3, np.inf), np.full(3, -np.inf)
        for s, leaf, lo, hi, ref in children.get(ni, []):
            if leaf:
                v = tris[ref][:, [0, 1, 2, 4, 5, 6, 8, 9, 10]].reshape(-1, 3).astype(np.float64)
                clo, chi = v.min(0), v.max(0)
            else:
                clo, chi = rec(ref)
            assert (lo <= clo + 0).all() and (hi >= chi - 0).all(), "quantised child box does not enclose its subtree"
            blo, bhi = np.minimum(blo, clo), np.maximum(bhi, chi)
        bounds[ni] = (blo, bhi)
        return blo, bhi

    rec(0)
    return bounds


def closest_hit(nodes, tris, ray_o, ray_d, tmax, tri_test):
    """Emulates Traverser<false>::run for one ray; returns (t, tri_index or -1, nodes_visited)."""
    o = np.asarray(ray_o, f32)
    d = np.asarray(ray_d, f32)
    inv = [_safe_inv(d[a]) for a in range(3)]
    octant = (1 if inv[0] < 0 else 0) | (2 if inv[1] < 0 else 0) | (4 if inv[2] < 0 else 0)
    octinv = 7 - octant
    closest, best, best_key = f32(tmax), -1, None
    visited = 0
    stack = []
    ngroup = [0, 0x80000000]
    while True:
        tgroup = [0, 0]
        if ngroup[1] > 0x00FFFFFF:
            hits = ngroup[1]
            bit = hits.bit_length() - 1
            ngroup[1] &= ~(1 << bit)
            if ngroup[1] > 0x00FFFFFF:
                stack.append(list(ngroup))
            slot = (bit - 24) ^ octinv
            rel = bin(hits & 0xFF & ((1 << slot) - 1)).count("1")
            ni = ngroup[0] + rel
            visited += 1
            node = nodes[ni]
            p = node[0:3].view(np.float32)
            e = [(int(node[3]) >> (8 * a)) & 0xFF for a in range(3)]
            imask = (int(node[3]) >> 24) & 0xFF
            b = node[6:20].view(np.uint8)
            meta, q = b[0:8], b[8:].reshape(6, 8)
            hitmask = 0
            for s in range(8):
                m = int(meta[s])
                if m == 0:
                    continue
                tmin, tmx = f32(0.0), closest
                for a in range(3):
                    # traverse.cuh (MCRT_QUANT_PRMT): byte -> 1 + b * 2^-15 (+ 2 ulp for the HIGH planes), scale and
                    # offset absorb the affine map
                    sc = np.array([(e[a] + 15) << 23], np.uint32).view(np.float32)[0]
                    ai = f32(sc * inv[a])
                    org = f32(f32(f32(p[a] - o[a]) * inv[a]) - ai)
                    lo = np.array([0x3F800000 | (int(q[a][s]) << 8)], np.uint32).view(np.float32)[0]
                    hi = np.array([0x3F800002 | (int(q[3 + a][s]) << 8)], np.uint32).view(np.float32)[0]
                    qa, qb = (hi, lo) if inv[a] < 0 else (lo, hi)
                    ta = f32(np.float64(qa) * np.float64(ai) + np.float64(org))   # fma
                    tb = f32(np.float64(qb) * np.float64(ai) + np.float64(org))
                    tmin, tmx = max(tmin, ta), min(tmx, tb)
                if tmin <= f32(tmx * f32(1.0000004)):
                    inner = (imask >> s) & 1
                    bit_index = ((m & 31) ^ (octinv if inner else 0)) & 31
                    hitmask |= (m >> 5) << bit_index
            ngroup = [int(node[4]), (hitmask & 0xFF000000) | imask]
            tgroup = [int(node[5]), hitmask & 0x00FFFFFF]
        while tgroup[1]:
            ti = tgroup[1].bit_length() - 1
            tgroup[1] &= ~(1 << ti)
            ta = tgroup[0] + ti
            t = tri_test(ta, closest)
            if t is not None:
                key = (int(tris[ta, 3:4].view(np.int32)[0]), int(tris[ta, 7:8].view(np.int32)[0]))
                if t < closest or (t == closest and best >= 0 and key < best_key):
                    closest, best, best_key = f32(t), ta, key
        if ngroup[1] <= 0x00FFFFFF:
            if not stack:
                break
            ngroup = stack.pop()
    return closest, best, visited
