"""Deterministic synthetic inputs shared by bench.py, tools/ and the tests (same integers from numpy and torch)."""
import numpy as np


def gen_rows(xp, y0, y1, w, device=None):
    """Rows [y0, y1) x w x 4 (uint8, alpha 255) of the giant-image workload (BASELINE configs[4]): an integer hash
    blended with a coarse gradient, computable row by row on any rank without the rest of the image."""
    if xp is np:
        y = np.arange(y0, y1, dtype=np.int64)[:, None, None]
        x = np.arange(w, dtype=np.int64)[None, :, None]
        c = np.arange(4, dtype=np.int64)[None, None, :]
    else:
        y = xp.arange(y0, y1, dtype=xp.int64, device=device)[:, None, None]
        x = xp.arange(w, dtype=xp.int64, device=device)[None, :, None]
        c = xp.arange(4, dtype=xp.int64, device=device)[None, None, :]
    h = (x * 73856093) ^ (y * 19349663) ^ (c * 83492791)
    h = (h ^ (h >> 13)) * 1274126177
    noise = (h >> 7) & 127
    smooth = ((x >> 3) * 5 + (y >> 2) * 3 + c * 40) & 127
    v = (noise + smooth) & 255
    if xp is np:
        v = v.astype(np.uint8)
        v[..., 3] = 255
    else:
        v = v.to(xp.uint8)
        v[..., 3] = 255
    return v


def make_frames(torch, frame0, n, h, w, device, seed=0x6E6F726D):
    """Frames [frame0, frame0 + n) of the batch workload (BASELINE configs[3]) as an (n, h, w, 4) uint8 tensor on
    `device`: PRNG noise blended with a smooth gradient whose phase depends on the GLOBAL frame index, so a frame has
    the same content whichever rank holds it."""
    out = torch.empty((n, h, w, 4), dtype=torch.uint8, device=device)
    ys = torch.linspace(0, 1, h, device=device).view(1, h, 1, 1)
    xs = torch.linspace(0, 1, w, device=device).view(1, 1, w, 1)
    ch = torch.arange(4, device=device, dtype=torch.float32).view(1, 1, 1, 4)
    step = 32
    for i in range(0, n, step):
        m = min(step, n - i)
        g = torch.Generator(device=device)
        g.manual_seed(seed + (frame0 + i) * 7919)
        noise = torch.randint(0, 256, (m, h, w, 4), dtype=torch.uint8, device=device, generator=g)
        ph = (torch.arange(frame0 + i, frame0 + i + m, device=device, dtype=torch.float32) * 0.37).view(m, 1, 1, 1)
        grad = 127.5 + 127.5 * torch.sin(6.0 * xs * (1 + 0.1 * ch) + 4.0 * ys + ph)
        out[i:i + m] = ((noise.to(torch.float32) + grad) * 0.5).to(torch.uint8)
        out[i:i + m, :, :, 3] = 255
    return out
