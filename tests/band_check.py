"""Checks one rank's band result against the CPU oracle (test infrastructure: bench.py's band key, tools/band_run.py)."""
import numpy as np

import oracle_lib as oracle
from synth import gen_rows


def check_band(job, f_kw, width, height, full=True, chunk_rows=2048):
    """-> dict(halo_rows_ok, rows_bit_exact, seams_bit_exact, oracle_moments).

    The rank's extended source band (owned rows + the halo rows the peer copies delivered) is downloaded.  Its halo
    rows and a sample of its owned rows must equal the generator (that checks the exchange itself); then the oracle
    chain runs on row chunks of it, each with `reach` rows of margin, and every owned output row (`full`) or the 40
    rows next to each seam must match the GPU result byte for byte.  A margin of the stencil reach is exact: an
    artificial chunk edge only disturbs output rows within the reach of it, and those are never compared."""
    b = job.band
    ext0 = b.row0 - b.halo_up
    src = b.src.download()[0]
    got = b.own_rows_of(b.dst)
    halo_ok = True
    for (lo, hi) in ((0, b.halo_up), (b.halo_up + b.rows, b.halo_up + b.rows + b.halo_down)):
        if hi > lo:
            halo_ok = halo_ok and bool(np.array_equal(src[lo:hi], gen_rows(np, ext0 + lo, ext0 + hi, width)))
    for y in (0, b.rows // 2, b.rows - 1):  # owned rows: a sample (they were generated on the device)
        halo_ok = halo_ok and bool(np.array_equal(src[b.halo_up + y], gen_rows(np, b.row0 + y, b.row0 + y + 1, width)[0]))
    reach = max(b.halo_up, b.halo_down, 1)
    if full:
        todo = [(y, min(y + chunk_rows, b.rows)) for y in range(0, b.rows, chunk_rows)]
    else:
        todo = [(max(0, s - 40), min(b.rows, s + 40)) for s in ({0} if b.halo_up else set()) | ({b.rows} if b.halo_down else set())]
    same = np.ones(b.rows, bool)
    mom = np.zeros(5)
    for (y0, y1) in todo:  # owned-row range [y0, y1) -> extended rows with margin
        e0, e1 = max(0, b.halo_up + y0 - reach), min(src.shape[0], b.halo_up + y1 + reach)
        part = np.ascontiguousarray(src[e0:e1])
        eo, _ = oracle.chain(part, oracle.make_uniform(width, e1 - e0, **f_kw), oracle.CS_RAW)
        exp = eo[b.halo_up + y0 - e0:b.halo_up + y1 - e0]
        same[y0:y1] = (got[y0:y1] == exp).all(axis=(1, 2))
        if full:
            mom += oracle.ssim_moments(exp, src[b.halo_up + y0:b.halo_up + y1])
    seams = True
    if b.halo_up:
        seams = seams and bool(same[:40].all())
    if b.halo_down:
        seams = seams and bool(same[-40:].all())
    return {"halo_rows_ok": halo_ok, "rows_bit_exact": bool(same.all()) if full else None, "seams_bit_exact": seams,
            "oracle_moments": mom}
