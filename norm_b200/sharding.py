"""Multi-GPU scheduling of the path, host side.

Frame batches shard with no communication (plan_frames).  One giant image shards into row
bands: `band_source_rows` says which global rows a rank must hold (owned rows + stencil
halo), `halo_requests` which of them come from which neighbour - the halo rows are the
neighbours' OWNED rows next to the seam, fetched with peer copies over NVLink.
"""
import ctypes as _C

import numpy as _np

from .api import Band, plan_band, plan_frames, ssim_from_moments  # noqa: F401


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


class BandJob:
    """One rank of a row-band sharded image (BASELINE configs[4]) ready to step: a Band (norm_cuda_band_*), the
    neighbours connected through `allgather` (any callable that returns every rank's bytes object, ordered by rank:
    torch.distributed.all_gather_object, mpi4py, a list for ranks inside one process), two device buffers for the
    SSIM moments and, when `comm` (norm_b200.nccl.Comm) is given, the all-reduce of the moments on the band's side
    stream.  The host blocks nowhere inside step()."""

    def __init__(self, ctx, f, width, height, nranks, rank, allgather, comm=None):
        self.ctx, self.comm, self.k = ctx, comm, 0
        self.band = Band(ctx, f, width, height, nranks, rank)
        self.band.connect(allgather(self.band.export()))
        self.mom = [ctx.image_create(64, 1, 1) for _ in range(2)]  # 256 bytes of device memory each
        self.side = self.band.side_stream()
        self.halo_bytes = (self.band.halo_up + self.band.halo_down) * width * 4

    def fill_rows(self, device_ptr, pitch, first_row, nrows):
        """Copies rows [first_row, first_row + nrows) of the rank's OWNED rows from a device buffer into the band."""
        self.ctx.copy_rows(self.band.src, self.band.halo_up + first_row, device_ptr, pitch, 0, nrows)

    def publish(self):
        self.band.publish()

    def step(self, flags):
        m = self.mom[self.k & 1]
        self.band.step_async(flags, m.device_ptr)
        if self.comm is not None:
            self.comm.allreduce_device(m.device_ptr, 5, self.side)
        self.k += 1
        return m

    def sync(self):
        self.band.sync()

    def moments(self):
        """The last step's moments (summed over ranks when a communicator was given); waits for the band."""
        self.sync()
        m = self.mom[(self.k - 1) & 1]
        return _np.frombuffer(m.download()[0].tobytes()[:40], _np.float64).copy()

    def close(self):
        self.sync()
        for m in self.mom:
            m.destroy()
        self.band.destroy()
