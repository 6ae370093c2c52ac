"""Shared helpers for the parity tests."""
import numpy as np

MISS = 0xFFFFFFFF


def srgb_oetf(x):
    x = np.clip(x, 0.0, 1.0)
    return np.where(x <= 0.0031308, 12.92 * x, 1.055 * np.power(x, 1 / 2.4) - 0.055)


def quantise8(y):
    """8-bit quantiser of the reference's PNG encoder: floor(256 y) clamped (pinned in test_oracle_golden)."""
    return np.minimum(255, np.floor(np.clip(y, 0.0, 1.0) * 256.0)).astype(np.uint8)


def srgb_eotf(c):
    return np.where(c <= 0.04045, c / 12.92, ((c + 0.055) / 1.055) ** 2.4)


def block_mean(a, b):
    h, w = a.shape[0] // b * b, a.shape[1] // b * b
    a = a[:h, :w]
    return a.reshape(h // b, b, w // b, b, -1).mean(axis=(1, 3))


def luminance(img):
    return 0.2126 * img[..., 0] + 0.7152 * img[..., 1] + 0.0722 * img[..., 2]


def z_scores(mean_a, sumsq_a, n_a, mean_b, sumsq_b, n_b):
    """Per pixel/channel z of the difference of two Monte Carlo means from Σx² buffers."""
    var_a = np.maximum(sumsq_a / n_a - mean_a ** 2, 0.0) / max(n_a - 1, 1) * n_a / n_a
    var_b = np.maximum(sumsq_b / n_b - mean_b ** 2, 0.0) / max(n_b - 1, 1) * n_b / n_b
    floor = (1e-3 * np.maximum(np.abs(mean_b), 1e-2)) ** 2      # fp32-vs-fp64 floor for zero-variance pixels
    return (mean_a - mean_b) / np.sqrt(var_a + var_b + floor)


def check_device_tree(nodes, info):
    """Structural invariants of a flattened tree (rt_device.h): every inner node's box contains its children's boxes, leaf
    references address valid primitive ranges, pair indices are even, and every primitive hangs on some leaf (except the
    records of a medium boundary that walks the surface tree).  `nodes`: (n, 8) float32 view of the DNode array."""
    assert info["n_bvh_nodes"] == len(nodes) and info["max_depth"] < 32
    lo, hi = nodes[:, 0:3], nodes[:, 4:7]
    ref, cnt = nodes[:, 3].view(np.uint32), nodes[:, 7].view(np.uint32)
    LEAF = 0x80000000
    seen = np.zeros(info["n_prims"], bool)
    for k in range(len(nodes)):
        if ref[k] & LEAF:
            first, c = int(ref[k] & 0x7FFFFFFF) >> 5, int(ref[k] & 31) + 1
            assert c == cnt[k] and 1 <= c <= 32 and first + c <= info["n_prims"]
            assert not seen[first:first + c].any()                  # leaves are disjoint
            seen[first:first + c] = True
        else:
            c = int(ref[k])
            assert c % 2 == 0 and cnt[k] == 0 and c + 1 < len(nodes)
            for ch in (c, c + 1):
                assert (lo[k] <= lo[ch] + 1e-3).all() and (hi[k] >= hi[ch] - 1e-3).all()
    if len(nodes):
        # a leaf root is referenced from its DMedium record, not a node; a medium whose boundary IS the surface geometry walks
        # the surface tree, so its own copies of the records (half of the primitives: pawn, bunny) hang on no leaf
        spare = info["n_prims"] // 2 if info["n_shared_trees"] else 0
        assert (~seen).sum() <= 32 * max(1, info["n_media"]) + spare
        if info["n_shared_trees"]:
            assert (~seen).sum() == spare and not seen[spare:].any()      # exactly the boundary's records, stored behind the tree's
    return seen
