"""Multi-GPU scheduling of the path, host side.

Frame batches shard with no communication (plan_frames).  One giant image shards into row
bands: `band_source_rows` says which global rows a rank must hold (owned rows + stencil
halo), `halo_requests` which of them come from which neighbour - the halo rows are the
neighbours' OWNED rows next to the seam, fetched with peer copies over NVLink.
"""
from .api import plan_band, plan_frames  # noqa: F401


def band_layout(f, height, nranks, rank):
    """-> dict(row0, rows, up, down, ext_row0, ext_rows, own_off): the rank's extended band is global
    rows [ext_row0, ext_row0 + ext_rows); its owned rows start at local row own_off."""
    row0, rows, up, down = plan_band(f, height, nranks, rank)
    return {"row0": row0, "rows": rows, "up": up, "down": down, "ext_row0": row0 - up, "ext_rows": up + rows + down,
            "own_off": up}


def halo_requests(f, height, nranks, rank):
    """Peer copies this rank needs: list of (peer_rank, peer_local_row, my_local_row, nrows).

    Halo rows may span more than one neighbour when bands are shorter than the halo."""
    me = band_layout(f, height, nranks, rank)
    out = []
    for lo, hi in ((me["ext_row0"], me["row0"]), (me["row0"] + me["rows"], me["ext_row0"] + me["ext_rows"])):
        g = lo
        while g < hi:
            # owner of global row g
            for pr in range(nranks):
                pl = band_layout(f, height, nranks, pr)
                if pl["row0"] <= g < pl["row0"] + pl["rows"]:
                    n = min(hi, pl["row0"] + pl["rows"]) - g
                    out.append((pr, pl["own_off"] + (g - pl["row0"]), g - me["ext_row0"], n))
                    g += n
                    break
            else:
                raise AssertionError("row without owner")
    return out
