"""Synthetic inputs of SURVEY.md section 8(d): seeded occupancy patches and pose samples."""
import numpy as np


def gen_map(patch_dim, seed):
    """u8 patch_dim^2 sensor patch: background 0 (free), 2-cell border wall 255,
    patch_dim^2/16384 occupied rectangles (sides 8..96), patch_dim^2/65536 unknown rectangles
    (127, sides 16..128); numpy PCG64, values drawn in the fixed order x, y, w, h."""
    rng = np.random.default_rng(seed)
    m = np.zeros((patch_dim, patch_dim), np.uint8)
    n_occ = max(1, patch_dim * patch_dim // 16384)
    n_unk = max(1, patch_dim * patch_dim // 65536)
    for _ in range(n_occ):
        x, y = int(rng.integers(0, patch_dim)), int(rng.integers(0, patch_dim))
        w, h = int(rng.integers(8, 97)), int(rng.integers(8, 97))
        m[y:y + h, x:x + w] = 255
    for _ in range(n_unk):
        x, y = int(rng.integers(0, patch_dim)), int(rng.integers(0, patch_dim))
        w, h = int(rng.integers(16, 129)), int(rng.integers(16, 129))
        blk = m[y:y + h, x:x + w]
        blk[blk == 0] = 127
    m[:2, :] = 255
    m[-2:, :] = 255
    m[:, :2] = 255
    m[:, -2:] = 255
    return m


def sample_free_cells(astar_grid, n, seed):
    """n astar cells (x, y) that are not OCC, uniformly, seeded."""
    rng = np.random.default_rng(seed)
    ys, xs = np.nonzero(astar_grid != 2)
    pick = rng.integers(0, len(xs), n)
    return np.stack([xs[pick], ys[pick]], 1).astype(np.int32)
