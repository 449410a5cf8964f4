"""Seeded procedural frame pairs / sequences with a known flow.

Sintel is not available offline, so inputs are multi-octave band-limited noise textures warped
by an analytic flow.  Pure numpy/scipy host code: it only fabricates INPUTS (u8 grey frames and
the ground-truth flow); no part of the optical-flow path lives here.
"""
from __future__ import annotations

import numpy as np
from scipy import ndimage

_SIGMAS = (1.0, 2.5, 6.0, 16.0, 40.0)
_AMPS = (0.6, 0.8, 1.0, 1.0, 0.8)


def texture(h: int, w: int, rng: np.random.Generator) -> np.ndarray:
    """Multi-octave noise canvas scaled to 0..255 (float64)."""
    acc = np.zeros((h, w), np.float64)
    for s, a in zip(_SIGMAS, _AMPS):
        f = ndimage.gaussian_filter(rng.standard_normal((h, w)), s, mode="wrap")
        acc += a * f / (f.std() + 1e-12)
    acc -= acc.min()
    acc *= 255.0 / acc.max()
    return acc


def gt_flow(h: int, w: int, amp: float) -> np.ndarray:
    """Smooth analytic flow (h, w, 2): u = A sin(3pi x/W + .3) cos(pi y/H), v = .6A cos(2pi x/W + 1) sin(2pi y/H)."""
    y, x = np.mgrid[0:h, 0:w].astype(np.float64)
    u = amp * np.sin(3 * np.pi * x / w + 0.3) * np.cos(np.pi * y / h)
    v = 0.6 * amp * np.cos(2 * np.pi * x / w + 1.0) * np.sin(2 * np.pi * y / h)
    return np.stack([u, v], -1)


def _to_u8(a: np.ndarray) -> np.ndarray:
    return np.clip(np.rint(a), 0, 255).astype(np.uint8)


def make_pair(w: int = 1024, h: int = 436, seed: int = 0, amp: float | None = None, margin: int = 32, canvas_fn=None):
    """Frame A, frame B (u8, h x w) and the flow A->B (float32 h x w x 2).

    B(x + flow(x)) ~= A(x): B is the canvas sampled (bicubic) at x - flow_b where flow_b is the
    backward flow obtained by one fixed-point step; for the smooth few-pixel flows used here the
    forward/backward difference is far below the matcher's accuracy.
    """
    rng = np.random.default_rng(seed)
    if amp is None:
        amp = 3.0 * w / 1024.0
    canvas = texture(h + 2 * margin, w + 2 * margin, rng)
    if canvas_fn is not None:
        canvas = canvas_fn(canvas)
    f = gt_flow(h, w, amp)
    y, x = np.mgrid[0:h, 0:w].astype(np.float64)
    a = canvas[margin:margin + h, margin:margin + w]
    # backward warp: B(p) = A(p - f(p - f(p)))  (one fixed-point refinement of the inverse)
    fx = ndimage.map_coordinates(f[..., 0], [y - f[..., 1], x - f[..., 0]], order=1, mode="nearest")
    fy = ndimage.map_coordinates(f[..., 1], [y - f[..., 1], x - f[..., 0]], order=1, mode="nearest")
    b = ndimage.map_coordinates(canvas, [y - fy + margin, x - fx + margin], order=3, mode="nearest")
    return _to_u8(a), _to_u8(b), f.astype(np.float32)


def make_lowtex_pair(w: int = 1024, h: int = 436, seed: int = 0, amp: float | None = None):
    """make_pair on a canvas whose texture fades out from the middle to the right: the right 40 % of both
    frames is exactly flat (zero gradients => singular Hessians), the band before it
    has weak texture (long, badly conditioned iterations)."""
    def fade(canvas):
        ww = canvas.shape[1]
        x = np.arange(ww, dtype=np.float64) / ww
        m = np.clip((0.6 - x) / 0.2, 0.0, 1.0)          # 1 left of 40 %, 0 right of 60 %
        return 128.0 + (canvas - 128.0) * m[None, :]
    return make_pair(w, h, seed=seed, amp=amp, canvas_fn=fade)


def make_sequence(n_frames: int = 50, w: int = 1024, h: int = 436, seed: int = 0, margin: int = 176):
    """n_frames u8 frames sharing one canvas; frame k samples the canvas at S^-(k)(x) where
    S(x) = c + 1.002 R(0.15deg)(x - c) + (0.8, -0.4).  Every consecutive pair has GT flow S(x) - x.
    Returns (frames [n,h,w] u8, flow [h,w,2] float32)."""
    rng = np.random.default_rng(seed)
    canvas = texture(h + 2 * margin, w + 2 * margin, rng)
    c = np.array([(w - 1) / 2.0, (h - 1) / 2.0])
    th = np.deg2rad(0.15)
    A = 1.002 * np.array([[np.cos(th), -np.sin(th)], [np.sin(th), np.cos(th)]])
    t = np.array([0.8, -0.4])
    Ainv = np.linalg.inv(A)
    y, x = np.mgrid[0:h, 0:w].astype(np.float64)
    pts = np.stack([x - c[0], y - c[1]], 0).reshape(2, -1)
    frames = np.empty((n_frames, h, w), np.uint8)
    M, tt = np.eye(2), np.zeros(2)      # S^-k as x -> M (x - c) + tt + c
    for k in range(n_frames):
        q = M @ pts + tt[:, None]
        img = ndimage.map_coordinates(canvas, [q[1] + c[1] + margin, q[0] + c[0] + margin], order=3,
                                      mode="nearest").reshape(h, w)
        frames[k] = _to_u8(img)
        M, tt = Ainv @ M, Ainv @ (tt - t)   # compose S^-1 after: S^-1(x) = Ainv (x - c - t) + c
    fwd = A @ pts + t[:, None] - pts
    flow = np.stack([fwd[0].reshape(h, w), fwd[1].reshape(h, w)], -1).astype(np.float32)
    return frames, flow
