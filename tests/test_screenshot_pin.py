"""The reference repository's own output images against the oracle's driver-defined conventions
(tests/pin_screenshots.py).  CPU only.

The blur screenshot at the reference's root is the one effect OUTPUT the reference holds.  It is a colour-managed
window grab, so it cannot pin values to the LSB; it can rank hypotheses.  What it decides, and what it does not,
is in DESIGN.md section 2."""
import json
import os

import pin_screenshots as pin
import oracle_lib as orc


def _by(rows, **kw):
    return [r for r in rows if all(r[k] == v for k, v in kw.items())]


def test_blur_screenshot_prefers_floor_and_reference_colour_space():
    rows = pin.blur_table(radii=(21, 39))
    for r in (21, 39):
        ref_floor = _by(rows, radius=r, colour_space="reference", tie_break="floor")[0]
        ref_ceil = _by(rows, radius=r, colour_space="reference", tie_break="ceil")[0]
        raw_floor = _by(rows, radius=r, colour_space="raw", tie_break="floor")[0]
        # the oracle's floor tie-break of the half-texel taps fits the reference's output better than ceil ...
        assert ref_floor["rms"] < ref_ceil["rms"] and ref_floor["mae"] < ref_ceil["mae"]
        # ... and sRGB-decode-at-load / linear intermediates / sRGB surface far better than bytes/255 throughout
        assert ref_floor["rms"] < 0.9 * raw_floor["rms"]
    # the screenshot was taken at the UI's largest radius: r = 39 fits better than the default 21
    assert _by(rows, radius=39, colour_space="reference", tie_break="floor")[0]["rms"] < \
        0.5 * _by(rows, radius=21, colour_space="reference", tie_break="floor")[0]["rms"]


def test_committed_pin_table_is_current():
    """tests/golden/screenshot_pin.json (the table DESIGN.md quotes) is what the script produces today."""
    fixture = json.load(open(pin.FIXTURE))
    now = pin.blur_table(radii=(39,))
    want = [r for r in fixture["blur"] if r["radius"] == 39]
    for a, b in zip(now, want):
        assert (a["colour_space"], a["tie_break"]) == (b["colour_space"], b["tie_break"])
        assert abs(a["rms"] - b["rms"]) < 1e-6
    # line detection: recorded as inconclusive (the screenshot predates the current shader: dark lines on white)
    edge = fixture["edge"]
    assert edge is not None
    q = [r for r in edge if r["colour_space"] == "reference"]
    assert abs(q[0]["rms"] - q[1]["rms"]) < 0.05 * q[0]["rms"]      # -left vs -top_left: no measurable preference
    assert all(r["corr"] < -0.7 for r in edge)                       # inverted polarity relative to image_shader.wgsl:155
    assert orc.CS_REFERENCE == 0
