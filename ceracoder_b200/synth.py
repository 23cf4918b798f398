"""Deterministic synthetic I420 sources: S1 testsrc, S2 pan-sprites, S3 ball.

Pure numpy, no reference code involved; used by tests and bench.py for inputs only.
S1 mirrors the shape of `videotestsrc` as used by the reference's test-pattern pipelines
(/root/reference/pipeline/jetson/h265_test_pattern:1).
"""
import numpy as np


def _value_noise(h, w, rng, periods=(256, 64, 16, 4), amps=(64, 32, 16, 8)):
    out = np.full((h, w), 128.0, dtype=np.float32)
    for p, a in zip(periods, amps):
        gh, gw = h // p + 2, w // p + 2
        g = rng.uniform(-1, 1, size=(gh, gw)).astype(np.float32)
        ys = np.arange(h, dtype=np.float32) / p
        xs = np.arange(w, dtype=np.float32) / p
        y0 = ys.astype(np.int32); x0 = xs.astype(np.int32)
        fy = (ys - y0)[:, None]; fx = (xs - x0)[None, :]
        a00 = g[y0][:, x0]; a01 = g[y0][:, x0 + 1]; a10 = g[y0 + 1][:, x0]; a11 = g[y0 + 1][:, x0 + 1]
        out += a * ((a00 * (1 - fx) + a01 * fx) * (1 - fy) + (a10 * (1 - fx) + a11 * fx) * fy)
    return out


class PanSprites:
    """S2: panning value-noise background + bouncing textured sprites + grain; scene cut every 300 frames."""

    def __init__(self, width, height, seed=1234, n_sprites=8, cut_every=300, grain=2.0):
        self.w, self.h, self.seed, self.n, self.cut, self.grain = width, height, seed, n_sprites, cut_every, grain
        self._scene = -1

    def _setup(self, scene):
        rng = np.random.default_rng(self.seed + 7919 * scene)
        self.bg = _value_noise(self.h + 512, self.w + 1024, rng)
        self.bgc = [_value_noise(self.h // 2 + 256, self.w // 2 + 512, rng, (128, 32), (24, 12)) for _ in range(2)]
        s = min(128, self.h // 4 * 2, self.w // 4 * 2)
        self.ss = s
        self.spr = [np.clip(_value_noise(s, s, rng, (32, 8, 2), (60, 40, 20)) + rng.uniform(-40, 40), 0, 255) for _ in range(self.n)]
        self.pos = rng.uniform(0, 1, size=(self.n, 2)) * [self.w - s, self.h - s]
        self.vel = np.stack([rng.integers(5, 12, self.n) * rng.choice([-1, 1], self.n),
                             rng.integers(3, 8, self.n) * rng.choice([-1, 1], self.n)], 1).astype(np.float64)
        self._scene = scene

    def frame(self, i):
        scene, k = divmod(i, self.cut)
        if scene != self._scene:
            self._setup(scene)
        ox, oy = (3 * k) % 1024, (1 * k) % 512
        y = self.bg[oy:oy + self.h, ox:ox + self.w].copy()
        s = self.ss
        for j in range(self.n):
            p = self.pos[j] + self.vel[j] * k
            span = np.array([self.w - s, self.h - s], dtype=np.float64)
            p = np.abs((p % (2 * span)) - span)  # triangle-wave bounce
            px, py = int(p[0]) & ~1, int(p[1]) & ~1
            y[py:py + s, px:px + s] = self.spr[j]
        g = np.random.default_rng(2000 + i)
        y = y + g.normal(0, self.grain, size=y.shape).astype(np.float32)
        u = self.bgc[0][oy // 2:oy // 2 + self.h // 2, ox // 2:ox // 2 + self.w // 2]
        v = self.bgc[1][oy // 2:oy // 2 + self.h // 2, ox // 2:ox // 2 + self.w // 2]
        f = lambda a: np.clip(np.rint(a), 16, 235).astype(np.uint8)
        return f(y), f(u), f(v)


class TestSrc:
    """S1: static colour bars + a white-noise patch that changes every frame."""
    __test__ = False

    def __init__(self, width, height):
        self.w, self.h = width, height
        bars = [(180, 128, 128), (162, 44, 142), (131, 156, 44), (112, 72, 58), (84, 184, 198), (65, 100, 212), (35, 212, 114)]
        y = np.zeros((height, width), np.uint8); u = np.zeros((height // 2, width // 2), np.uint8); v = u.copy()
        for k, (a, b, c) in enumerate(bars):
            x0, x1 = width * k // 7, width * (k + 1) // 7
            y[:, x0:x1] = a; u[:, x0 // 2:x1 // 2] = b; v[:, x0 // 2:x1 // 2] = c
        self.base = (y, u, v)

    def frame(self, i):
        y, u, v = (p.copy() for p in self.base)
        g = np.random.default_rng(1000 + i)
        ph, pw = self.h // 4 & ~1, self.w // 6 & ~1
        y[self.h - ph:, self.w - pw:] = g.integers(16, 236, size=(ph, pw), dtype=np.uint8)
        return y, u, v


class Ball:
    """S3: black field + moving disc."""

    def __init__(self, width, height):
        self.w, self.h = width, height
        self.yy, self.xx = np.mgrid[0:height, 0:width]

    def frame(self, i):
        r = self.h // 10
        cx = r + (i * 7) % (self.w - 2 * r); cy = r + (i * 3) % (self.h - 2 * r)
        y = np.where((self.xx - cx) ** 2 + (self.yy - cy) ** 2 < r * r, 235, 16).astype(np.uint8)
        u = np.full((self.h // 2, self.w // 2), 128, np.uint8)
        return y, u, u.copy()


def make_source(name, width, height):
    return {"pan": PanSprites, "testsrc": TestSrc, "ball": Ball}[name](width, height)
