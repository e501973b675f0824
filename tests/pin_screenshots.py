#!/usr/bin/env python
"""Pins the oracle's driver-defined conventions against the only effect OUTPUTS the reference repository holds:
the window screenshots at its root (`obama-luma-gaussian-blur.png`, `obama-line-detection.png`).

    python tests/pin_screenshots.py            # prints the tables of DESIGN.md section 2, writes the JSON fixture

What can be decided from them, and how:
  * The screenshots are 1:1 window grabs of tests/obama.png (1024 x 683) under a macOS title bar.  The content
    origin is read off the window chrome (first / last smooth row and column), NOT fitted, so a one-pixel kernel
    shift cannot hide in the registration.
  * They went through the display's colour management, so absolute values are not comparable.  Each hypothesis
    is scored by the residual after the best monotone value mapping (mean screenshot value per predicted level).
  * Blur: hypotheses = radius x {floor, ceil tie-break of the half-texel NEAREST taps} x {reference colour
    space (sRGB decode at load, linear intermediates, sRGB encode at the surface), raw}.
  * Sobel: hypotheses = {`-left` as written at image_shader.wgsl:151, textbook `-top_left`} x colour space.

Test infrastructure: uses the CPU oracle; nothing in norm_b200/ imports this.
"""
import json
import os
import sys

import numpy as np
from PIL import Image

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
import oracle_lib as orc  # noqa: E402

GOLD = os.path.join(HERE, "golden")
BLUR_SHOT = os.path.join(GOLD, "images", "ref_obama_luma_gaussian_blur.png")   # copy of /root/reference/obama-luma-gaussian-blur.png
EDGE_SHOT_REF = "/root/reference/obama-line-detection.png"                      # read only when the reference tree is present
FIXTURE = os.path.join(GOLD, "screenshot_pin.json")
# content origin (row, col) inside each screenshot, from the window chrome: title bar + 1-px frame lines
BLUR_ORIGIN, EDGE_ORIGIN = (60, 2), (58, 0)


def obama_rgba():
    rgb = np.array(Image.open(os.path.join(GOLD, "images", "obama.png")).convert("RGB"))
    out = np.zeros(rgb.shape[:2] + (4,), np.uint8)  # the reference's expansion: alpha 0 for RGB (grammar.rs:130)
    out[..., :3] = rgb
    return out


def srgb_encode(lin_u8):
    """What the sRGB surface does to a linear unorm value on store (gpu_state.rs:94-99), to 8 bits."""
    x = lin_u8.astype(np.float64) / 255.0
    y = np.where(x <= 0.0031308, 12.92 * x, 1.055 * np.power(x, 1 / 2.4) - 0.055)
    return np.clip(np.rint(y * 255), 0, 255)


def predict_blur(src, radius, tie, cs):
    """grayscale + blur through the oracle; tie = 'floor' (the oracle's convention) or 'ceil' (its mirror image:
    flipping the image in both axes turns floor into ceil)."""
    h, w = src.shape[:2]
    f = orc.make_uniform(w, h, grayscale=1, blur=1, blur_radius=radius)
    img = src if tie == "floor" else np.ascontiguousarray(src[::-1, ::-1])
    out = orc.effects(img, f, cs)
    if tie != "floor":
        out = np.ascontiguousarray(out[::-1, ::-1])
    g = out[..., 0]
    return srgb_encode(g) if cs == orc.CS_REFERENCE else g.astype(np.float64)


def residual(pred_levels, shot_gray, origin):
    """rms / mean |.| of screenshot - best monotone mapping of the prediction (per predicted 8-bit level)."""
    h, w = pred_levels.shape
    crop = shot_gray[origin[0]:origin[0] + h, origin[1]:origin[1] + w]
    q = np.clip(np.rint(pred_levels), 0, 255).astype(np.int64).ravel()
    c = crop.ravel()
    cnt = np.bincount(q, minlength=256)
    lut = np.bincount(q, weights=c, minlength=256) / np.maximum(cnt, 1)
    res = c - lut[q]
    return float(np.sqrt((res ** 2).mean())), float(np.abs(res).mean())


def blur_table(radii=(21, 37, 39)):
    src = obama_rgba()
    shot = np.array(Image.open(BLUR_SHOT).convert("RGB")).astype(np.float64).mean(axis=2)
    rows = []
    for r in radii:
        for cs, csname in ((orc.CS_REFERENCE, "reference"), (orc.CS_RAW, "raw")):
            for tie in ("floor", "ceil"):
                rms, mae = residual(predict_blur(src, r, tie, cs), shot, BLUR_ORIGIN)
                rows.append({"radius": r, "colour_space": csname, "tie_break": tie, "rms": rms, "mae": mae})
    return rows


def sobel_np(rgb, quirk, cs_ref):
    """image_shader.wgsl:132-156 in numpy (f64), with either gy formula; texture A holds linear unorm8."""
    lin = np.array([(c / 255) / 12.92 if c / 255 <= 0.04045 else (((c / 255) + 0.055) / 1.055) ** 2.4 for c in range(256)])
    t = (np.rint(lin * 255) / 255.0) if cs_ref else (np.arange(256) / 255.0)
    inten = np.sqrt((t[rgb] ** 2).sum(axis=2))
    p = np.pad(inten, 1, mode="edge")
    tl, tp, tr = p[2:, :-2], p[2:, 1:-1], p[2:, 2:]   # "top" is row + 1: texture v grows downward
    lf, rt = p[1:-1, :-2], p[1:-1, 2:]
    bl, bt, br = p[:-2, :-2], p[:-2, 1:-1], p[:-2, 2:]
    gx = tl + 2 * lf + bl - tr - 2 * rt - br
    gy = (-lf if quirk else -tl) - 2 * tp - tr + bl + 2 * bt + br
    return np.clip(np.sqrt(gx * gx + gy * gy), 0, 1) * 255.0


def edge_table():
    if not os.path.exists(EDGE_SHOT_REF):
        return None
    rgb = obama_rgba()[..., :3]
    shot = np.array(Image.open(EDGE_SHOT_REF).convert("RGB")).astype(np.float64).mean(axis=2)
    rows = []
    for cs_ref in (True, False):
        for quirk in (True, False):
            pred = sobel_np(rgb, quirk, cs_ref)
            rms, mae = residual(pred, shot, EDGE_ORIGIN)
            h, w = pred.shape
            crop = shot[EDGE_ORIGIN[0]:EDGE_ORIGIN[0] + h, EDGE_ORIGIN[1]:EDGE_ORIGIN[1] + w]
            rows.append({"colour_space": "reference" if cs_ref else "raw", "gy": "-left (as written)" if quirk else "-top_left (textbook)",
                         "rms": rms, "mae": mae, "corr": float(np.corrcoef(pred.ravel(), crop.ravel())[0, 1])})
    return rows


def main():
    out = {"blur": blur_table((17, 21, 25, 31, 35, 37, 39)), "edge": edge_table(),
           "blur_origin": BLUR_ORIGIN, "edge_origin": EDGE_ORIGIN}
    print("blur screenshot (grayscale + blur), residual after the best monotone value mapping:")
    for r in out["blur"]:
        print("  r=%2d  %-9s %-5s  rms %.3f  mae %.3f" % (r["radius"], r["colour_space"], r["tie_break"], r["rms"], r["mae"]))
    if out["edge"]:
        print("line-detection screenshot:")
        for r in out["edge"]:
            print("  %-9s %-22s rms %.3f  mae %.3f  corr %.4f" % (r["colour_space"], r["gy"], r["rms"], r["mae"], r["corr"]))
    with open(FIXTURE, "w") as f:
        json.dump(out, f, indent=1)
    print("wrote", FIXTURE)


if __name__ == "__main__":
    main()
