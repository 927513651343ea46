"""Seeded synthetic inputs for the ORB front end.

Every generator draws from splitmix64(seed) so the same frames can be produced by any
language; the stream element i is mix(seed + (i+1)*GOLDEN), which vectorises in numpy.
"""
import numpy as np

_GOLDEN = np.uint64(0x9E3779B97F4A7C15)
_M1 = np.uint64(0xBF58476D1CE4E5B9)
_M2 = np.uint64(0x94D049BB133111EB)


def splitmix64(seed, n, offset=0):
    """n consecutive outputs of splitmix64 seeded with `seed`, starting at stream position `offset`."""
    with np.errstate(over="ignore"):
        i = np.arange(offset + 1, offset + n + 1, dtype=np.uint64)
        z = np.uint64(seed) + i * _GOLDEN
        z = (z ^ (z >> np.uint64(30))) * _M1
        z = (z ^ (z >> np.uint64(27))) * _M2
        return z ^ (z >> np.uint64(31))


def _uniform_int(u64, lo, hi):
    """Map raw 64-bit draws to integers in [lo, hi] (inclusive)."""
    span = np.uint64(hi - lo + 1)
    return (u64 % span).astype(np.int64) + lo


def _rect_params(W, H, seed):
    K = int(round(W * H / 768.0))
    r = splitmix64(seed, 5 * K)
    cx = _uniform_int(r[0::5], 0, W - 1)
    cy = _uniform_int(r[1::5], 0, H - 1)
    rw = _uniform_int(r[2::5], 4, 96)
    rh = _uniform_int(r[3::5], 4, 96)
    g = _uniform_int(r[4::5], 0, 255)
    return K, cx, cy, rw, rh, g


def _background(W, H):
    x = np.arange(W, dtype=np.float64)[None, :]
    y = np.arange(H, dtype=np.float64)[:, None]
    return np.floor(96 + 64 * x / W + 32 * y / H).astype(np.int64)


def _noise(W, H, seed, amp):
    r = splitmix64(seed ^ 0x5EED5EED, W * H)
    return _uniform_int(r, -amp, amp).reshape(H, W)


def rects(W, H, seed):
    """Background ramp + K=round(W*H/768) grey rectangles + uniform noise in [-8, 8]."""
    img = _background(W, H)
    K, cx, cy, rw, rh, g = _rect_params(W, H, seed)
    for k in range(K):
        x0 = max(int(cx[k] - rw[k] // 2), 0); x1 = min(int(cx[k] + (rw[k] + 1) // 2), W)
        y0 = max(int(cy[k] - rh[k] // 2), 0); y1 = min(int(cy[k] + (rh[k] + 1) // 2), H)
        img[y0:y1, x0:x1] = g[k]
    img = img + _noise(W, H, seed, 8)
    return np.clip(img, 0, 255).astype(np.uint8)


def stereo_rects(W, H, seed):
    """Left/right pair: the right image paints the same rectangles shifted left by a per-rectangle
    disparity in [1, 64]; rectangles are painted in order of increasing disparity (nearer = later)."""
    K, cx, cy, rw, rh, g = _rect_params(W, H, seed)
    disp = _uniform_int(splitmix64(seed ^ 0xD15FA417, K), 1, 64)
    order = np.argsort(disp, kind="stable")
    left = _background(W, H)
    right = _background(W, H)
    for k in order:
        x0 = int(cx[k] - rw[k] // 2); x1 = int(cx[k] + (rw[k] + 1) // 2)
        y0 = max(int(cy[k] - rh[k] // 2), 0); y1 = min(int(cy[k] + (rh[k] + 1) // 2), H)
        left[y0:y1, max(x0, 0):min(x1, W)] = g[k]
        d = int(disp[k])
        if x1 - d > 0:
            right[y0:y1, max(x0 - d, 0):min(x1 - d, W)] = g[k]
    left = np.clip(left + _noise(W, H, seed, 8), 0, 255).astype(np.uint8)
    right = np.clip(right + _noise(W, H, seed + 77777, 8), 0, 255).astype(np.uint8)
    return left, right


def texture(W, H, seed):
    """Uniform u8 noise, 5x5 box filtered (many weak corners)."""
    r = (splitmix64(seed, (W + 4) * (H + 4)) & np.uint64(255)).astype(np.int64).reshape(H + 4, W + 4)
    c = np.cumsum(np.cumsum(np.pad(r, ((1, 0), (1, 0))), axis=0), axis=1)
    s = c[5:, 5:] - c[:-5, 5:] - c[5:, :-5] + c[:-5, :-5]
    return (s // 25).astype(np.uint8)


def lowcontrast(W, H, seed):
    """128 +- uniform[-10, 10]: the ini threshold fails in most cells."""
    return (128 + _noise(W, H, seed, 10)).astype(np.uint8)


def flat(W, H, seed=0):
    return np.full((H, W), 128, dtype=np.uint8)


KINDS = {"rects": rects, "texture": texture, "lowcontrast": lowcontrast, "flat": flat}


def frame(kind, W, H, seed):
    return KINDS[kind](W, H, seed)


def descriptor_sets(N, seed):
    """Config 4: A = N random 256-bit rows; B[pi(i)] = A[i] with k~U[0,60] bits flipped for 70% of the rows,
    the rest random.  Returns (A, B) as uint8 [N, 32]."""
    A = splitmix64(seed, 4 * N).view(np.uint8).reshape(N, 32).copy()
    rnd = splitmix64(seed ^ 0xB0B, 4 * N).view(np.uint8).reshape(N, 32)
    ctl = splitmix64(seed ^ 0xC0FFEE, 3 * N)
    keep = (ctl[0::3] % np.uint64(10)) < np.uint64(7)
    k = (ctl[1::3] % np.uint64(61)).astype(np.int64)
    perm = np.argsort(ctl[2::3], kind="stable")
    flips = splitmix64(seed ^ 0xF11B, 60 * N).reshape(N, 60) % np.uint64(256)
    B = np.empty_like(A)
    Abits = np.unpackbits(A, axis=1, bitorder="little")
    for i in range(N):
        if keep[i]:
            row = Abits[i].copy()
            pos = flips[i, :k[i]].astype(np.int64)
            row[pos] ^= 1          # repeated positions flip once (fancy-index semantics): deterministic
            B[perm[i]] = np.packbits(row, bitorder="little")
        else:
            B[perm[i]] = rnd[i]
    return A, B
