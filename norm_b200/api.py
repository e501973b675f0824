"""Host-side mirror of the reference's effect API over the C ABI.

The reference is compiled Rust; no Rust toolchain exists in this image, so the host layer a
norm user would touch is mirrored here with the reference's names, argument meaning and
error behaviour (INTEGRATION.md holds the Rust binding a maintainer would add):

  FeatureUniform   src/renderer/feature_uniform.rs
  Png / ImageExt   src/png/grammar.rs:65-198, src/image/grammar.rs:48-68
  Texture          src/renderer/texture.rs:22-80
  EffectPipeline   src/renderer/effect_pipeline.rs:8-132
  compute_sim      src/png/ssim.rs:91-106
  Circle / ShapeUniform / CircleData / DrawUniform
                   src/renderer/shape.rs:188-243, shape_uniform.rs:3-60, draw_uniform.rs:3-56

Everything that computes goes through libnorm_cuda.so; numpy is used only to hold host bytes.
"""
import ctypes as C

import numpy as np

from . import _lib
from ._lib import (NORM_CS_RAW, NORM_CS_REFERENCE, NORM_EXPAND_COMPAT, NORM_EXPAND_SPEC, NORM_OK,
                   NORM_MAX_CIRCLES, NormCircleData, NormDrawUniform, NormFeatureUniform, NormInflatedDesc,
                   NormPngDesc, NormShapeUniform)


class NormError(RuntimeError):
    def __init__(self, status, message):
        super().__init__(f"[{_lib.load().norm_cuda_status_string(status).decode()}] {message}")
        self.status = status


class FeatureUniform:
    """feature_uniform.rs: same fields, defaults, toggles and clamps."""
    DEFAULT_BLUR_RADIUS, MAX_BLUR_RADIUS, MIN_BLUR_RADIUS = 21, 39, 3
    DEFAULT_SHARPEN_FACTOR, MAX_SHARPEN_FACTOR, MIN_SHARPEN_FACTOR = 16, 40, 1

    def __init__(self, width, height, gamma=0):  # feature_uniform.rs:20-36 `new`
        r = NormFeatureUniform()
        r.width, r.height, r.gamma = width, height, gamma
        r.blur_radius, r.sharpen_factor = self.DEFAULT_BLUR_RADIUS, self.DEFAULT_SHARPEN_FACTOR
        for i in range(4):
            r.transform[i][i] = 1.0
        self.raw = r

    # predicates compare with 1 exactly, as the reference does (feature_uniform.rs:40-42)
    def grayscale(self): return self.raw.grayscale == 1
    def invert(self): return self.raw.invert == 1
    def blur(self): return self.raw.blur == 1
    def sharpen(self): return self.raw.sharpen == 1
    def edge_detect(self): return self.raw.edge_detect == 1
    def gamma(self): return self.raw.gamma

    def toggle_grayscale(self): self.raw.grayscale = int(not self.grayscale())
    def toggle_invert(self): self.raw.invert = int(not self.invert())
    def toggle_blur(self): self.raw.blur = int(not self.blur())
    def toggle_sharpen(self): self.raw.sharpen = int(not self.sharpen())
    def toggle_edge_detect(self): self.raw.edge_detect = int(not self.edge_detect())

    def increase_blur_radius(self):  # feature_uniform.rs:76-78
        self.raw.blur_radius = min(self.raw.blur_radius + 2, self.MAX_BLUR_RADIUS)

    def decrease_blur_radius(self):  # feature_uniform.rs:80-82
        self.raw.blur_radius = max(self.raw.blur_radius - 2, self.MIN_BLUR_RADIUS)

    def increase_sharpen_factor(self):
        self.raw.sharpen_factor = min(self.raw.sharpen_factor + 1, self.MAX_SHARPEN_FACTOR)

    def decrease_sharpen_factor(self):
        self.raw.sharpen_factor = max(self.raw.sharpen_factor - 1, self.MIN_SHARPEN_FACTOR)

    def update_window_dimensions(self, width, height):
        self.raw.width, self.raw.height = width, height

    def apply_transform(self, action):  # feature_uniform.rs:161-175; action in {"FlipX", "FlipY"}
        d = {"FlipX": (0, 0), "FlipY": (1, 1)}[action]
        self.raw.transform[d[0]][d[1]] *= -1.0

    def copy(self):
        f = FeatureUniform(self.raw.width, self.raw.height, self.raw.gamma)
        C.memmove(C.byref(f.raw), C.byref(self.raw), C.sizeof(NormFeatureUniform))
        return f


class Circle:
    """src/renderer/shape.rs:188-243: centre + radius, either in pixel or in normalised window space."""

    def __init__(self, center, radius):
        self.center, self.radius = (np.float32(center[0]), np.float32(center[1])), np.float32(radius)

    @classmethod
    def from_pixel_space(cls, center, radius, window_dimension):
        w, h = np.float32(window_dimension[0]), np.float32(window_dimension[1])
        return cls((np.float32(center[0]) / w, np.float32(center[1]) / h), np.float32(radius) / min(w, h))

    def in_normalized_space(self, window_dimension):
        return Circle.from_pixel_space(self.center, self.radius, window_dimension)

    def in_pixel_space(self, window_dimension):
        w, h = np.float32(window_dimension[0]), np.float32(window_dimension[1])
        return Circle((self.center[0] * w, self.center[1] * h), self.radius * min(w, h))

    def translate(self, offset):
        return Circle((self.center[0] + np.float32(offset[0]), self.center[1] + np.float32(offset[1])), self.radius)

    def data(self):
        """impl From<&Circle> for CircleData (shape_uniform.rs:47-58)"""
        return NormCircleData(float(self.center[0]), float(self.center[1]), float(self.radius), 0.0)


class ShapeUniform:
    """src/renderer/shape_uniform.rs:3-34"""

    def __init__(self, width, height):
        self.raw = NormShapeUniform(width, height, 0, 0xFFFFFFFF)

    def update_dimensions(self, width, height): self.raw.width, self.raw.height = width, height
    def set_num_circles(self, count): self.raw.num_circles = count
    def set_selected_circle(self, selected): self.raw.selected_circle = 0xFFFFFFFF if selected is None else selected


class DrawUniform:
    """src/renderer/draw_uniform.rs:3-56"""

    def __init__(self):
        self.raw = NormDrawUniform()
        for i in range(4):
            self.raw.camera_view_proj[i][i] = 1.0

    def crosshair(self): return self.raw.crosshair == 1
    def toggle_crosshair(self): self.raw.crosshair = int(not self.crosshair())
    def set_circle_center(self, x, y): self.raw.circle_center_x, self.raw.circle_center_y = x, y
    def set_circle_radius(self, radius): self.raw.circle_radius = radius


class Png:
    """Decoded image as the reference holds it (grammar.rs:65-70) + PLTE/tRNS bodies."""

    def __init__(self, width, height, bit_depth, color_type, pixel_buffer, gamma=0, plte=None, trns=None):
        self.width, self.height = int(width), int(height)
        self.bit_depth, self.color_type, self._gamma = int(bit_depth), int(color_type), int(gamma)
        self.pixel_buffer = np.ascontiguousarray(np.frombuffer(bytes(pixel_buffer), np.uint8)
                                                 if not isinstance(pixel_buffer, np.ndarray) else pixel_buffer,
                                                 dtype=np.uint8).reshape(-1)
        self.plte = None if plte is None or len(plte) == 0 else np.ascontiguousarray(plte, np.uint8).reshape(-1)
        self.trns = None if trns is None or len(trns) == 0 else np.ascontiguousarray(trns, np.uint8).reshape(-1)

    def dimensions(self): return (self.width, self.height)
    def gamma(self): return self._gamma

    def desc(self):
        d = NormPngDesc()
        d.width, d.height, d.bit_depth, d.color_type, d.gamma = (self.width, self.height, self.bit_depth,
                                                                 self.color_type, self._gamma)
        d.pixels, d.len = self.pixel_buffer.ctypes.data, self.pixel_buffer.size
        if self.plte is not None:
            d.plte, d.n_plte = self.plte.ctypes.data, self.plte.size // 3
        if self.trns is not None:
            d.trns, d.n_trns = self.trns.ctypes.data, self.trns.size
        return d

    def rgba8(self, ctx, mode=NORM_EXPAND_COMPAT):
        """ImageExt::rgba8 (grammar.rs:123-155), computed on the GPU."""
        return ctx.expand_rgba8_host(self, mode)

    def luma_mode(self):
        """The branch of Png::luma_buffer (ssim.rs:18-87) this image's colour type takes."""
        if self.color_type in (2, 6):
            return _lib.NORM_LUMA_RGB
        if self.color_type == 0:
            return _lib.NORM_LUMA_GRAY
        if self.color_type == 4:
            return _lib.NORM_LUMA_GRAY_UNIT
        raise NormError(_lib.NORM_E_UNSUPPORTED, "luma buffer of a palette image: unimplemented in the reference (ssim.rs:83-86)")

    def compute_sim(self, reference_image, ctx, mode=NORM_EXPAND_COMPAT):
        """Png::compute_sim (ssim.rs:91-106): `self` is the test image; each image's luma follows its colour type."""
        if self.dimensions() != reference_image.dimensions():
            raise NormError(_lib.NORM_E_DIM_MISMATCH,
                            "Expect reference and test images to have identical dimensions.")
        a, b = Texture.from_image(ctx, self, mode), Texture.from_image(ctx, reference_image, mode)
        try:
            return float(ctx.ssim_luma(a.image, self.luma_mode(), b.image, reference_image.luma_mode())[0])
        finally:
            a.destroy(), b.destroy()


class DeviceImage:
    """`frames` pitched RGBA8 frames in HBM (norm_img)."""

    def __init__(self, ctx, handle):
        self.ctx, self.handle = ctx, handle
        w, h, n, pitch, fs, ptr = C.c_uint32(), C.c_uint32(), C.c_uint32(), C.c_size_t(), C.c_size_t(), C.c_void_p()
        ctx._ok(ctx.lib.norm_cuda_image_info(handle, w, h, n, pitch, fs, C.byref(ptr)))
        self.width, self.height, self.frames = w.value, h.value, n.value
        self.pitch, self.frame_stride, self.device_ptr = pitch.value, fs.value, ptr.value

    @property
    def nbytes(self): return self.frame_stride * self.frames

    def upload(self, host, frame0=0):
        host = np.ascontiguousarray(host, np.uint8)
        n = host.size // (self.width * self.height * 4)
        assert host.size == n * self.width * self.height * 4
        self.ctx._ok(self.ctx.lib.norm_cuda_upload(self.ctx.h, self.handle, frame0, n, host.ctypes.data, 0))
        return self

    def download(self, frame0=0, nframes=None):
        n = self.frames - frame0 if nframes is None else nframes
        out = np.empty((n, self.height, self.width, 4), np.uint8)
        self.ctx._ok(self.ctx.lib.norm_cuda_download(self.ctx.h, self.handle, frame0, n, out.ctypes.data, 0))
        return out

    def destroy(self):
        if self.handle:
            self.ctx.lib.norm_cuda_image_destroy(self.ctx.h, self.handle)
            self.handle = None


class Context:
    """One CUDA device + streams (stands where GpuResourceAllocator does, gpu_state.rs:49-132)."""

    def __init__(self, device=0, stream=None):
        self.lib = _lib.load()
        h = C.c_void_p()
        if stream is None:
            st = self.lib.norm_cuda_create(device, C.byref(h))
        else:
            st = self.lib.norm_cuda_create_on_stream(device, C.c_void_p(stream), C.byref(h))
        if st != NORM_OK:
            raise NormError(st, self.lib.norm_cuda_last_error(None).decode())
        self.h, self.device = h, device

    def _ok(self, st):
        if st != NORM_OK:
            raise NormError(st, self.lib.norm_cuda_last_error(self.h).decode())

    def close(self):
        if self.h:
            self.lib.norm_cuda_destroy(self.h)
            self.h = None

    def __enter__(self): return self
    def __exit__(self, *a): self.close()

    # -- facts / timing
    def device_info(self):
        sm, mem, name = C.c_int32(), C.c_size_t(), C.create_string_buffer(128)
        self._ok(self.lib.norm_cuda_device_info(self.h, sm, mem, name, 128))
        return {"sm_count": sm.value, "total_mem": mem.value, "name": name.value.decode()}

    def launch_count(self): return int(self.lib.norm_cuda_launch_count(self.h))
    def sync(self): self._ok(self.lib.norm_cuda_sync(self.h))
    def timer_start(self): self._ok(self.lib.norm_cuda_timer_start(self.h))
    def timer_stop(self): self._ok(self.lib.norm_cuda_timer_stop(self.h))

    def timer_elapsed_ms(self):
        ms = C.c_float()
        self._ok(self.lib.norm_cuda_timer_elapsed_ms(self.h, ms))
        return ms.value

    # -- buffers
    def image_create(self, width, height, frames=1):
        h = C.c_void_p()
        self._ok(self.lib.norm_cuda_image_create(self.h, width, height, frames, C.byref(h)))
        return DeviceImage(self, h)

    def image_wrap(self, device_ptr, width, height, frames, pitch, frame_stride):
        h = C.c_void_p()
        self._ok(self.lib.norm_cuda_image_wrap(self.h, C.c_void_p(device_ptr), width, height, frames, pitch,
                                               frame_stride, C.byref(h)))
        return DeviceImage(self, h)

    def image_from_host(self, host):
        host = np.ascontiguousarray(host, np.uint8)
        if host.ndim == 3:
            host = host[None]
        n, hgt, wid, c = host.shape
        assert c == 4
        return self.image_create(wid, hgt, n).upload(host)

    def host_alloc(self, nbytes):
        """Pinned host buffer as a numpy uint8 array (free with host_free)."""
        p = C.c_void_p()
        self._ok(self.lib.norm_cuda_host_alloc(self.h, nbytes, C.byref(p)))
        arr = np.ctypeslib.as_array(C.cast(p, C.POINTER(C.c_uint8)), shape=(nbytes,))
        return arr, p

    def host_free(self, p): self.lib.norm_cuda_host_free(self.h, p)

    # -- expansion
    def expand_rgba8(self, png, mode, out, frame=0):
        d = png.desc()
        self._ok(self.lib.norm_cuda_expand_rgba8(self.h, C.byref(d), mode, out.handle, frame))

    def expand_rgba8_device(self, png, device_ptr, row_pitch, frame_stride, frames, mode, out, frame0=0):
        """Scanlines already in device memory (png supplies geometry / PLTE / tRNS); asynchronous, one launch."""
        d = png.desc()
        d.pixels = C.c_void_p(device_ptr)
        d.len = row_pitch * png.height
        self._ok(self.lib.norm_cuda_expand_rgba8_device(self.h, C.byref(d), row_pitch, frame_stride, frames, mode, out.handle, frame0))

    def expand_rgba8_host(self, png, mode):
        out = np.empty((png.height, png.width, 4), np.uint8)
        d = png.desc()
        self._ok(self.lib.norm_cuda_expand_rgba8_host(self.h, C.byref(d), mode, out.ctypes.data))
        return out

    # -- unfilter (ScanlineReader::read_lines on the GPU)
    @staticmethod
    def _inflated(width, height, bit_depth, color_type, interlace, data):
        d = NormInflatedDesc()
        d.width, d.height, d.bit_depth, d.color_type, d.interlace = width, height, bit_depth, color_type, interlace
        buf = np.ascontiguousarray(np.frombuffer(bytes(data), np.uint8))
        d.data, d.len = buf.ctypes.data, buf.size
        return d, buf

    def unfilter(self, width, height, bit_depth, color_type, interlace, inflated):
        """inflated IDAT stream -> pixel_buffer (numpy uint8), scanline_reader.rs:30-168."""
        ch = {0: 1, 2: 3, 3: 1, 4: 2, 6: 4}[color_type]
        out = np.empty(height * ((width * ch * bit_depth + 7) // 8), np.uint8)
        d, keep = self._inflated(width, height, bit_depth, color_type, interlace, inflated)
        self._ok(self.lib.norm_cuda_unfilter(self.h, C.byref(d), out.ctypes.data))
        return out

    def unfilter_expand(self, width, height, bit_depth, color_type, interlace, inflated, mode, out, frame=0,
                        plte=None, trns=None):
        d, keep = self._inflated(width, height, bit_depth, color_type, interlace, inflated)
        pl = None if plte is None or len(plte) == 0 else np.ascontiguousarray(plte, np.uint8)
        tr = None if trns is None or len(trns) == 0 else np.ascontiguousarray(trns, np.uint8)
        self._ok(self.lib.norm_cuda_unfilter_expand_rgba8(
            self.h, C.byref(d), pl.ctypes.data if pl is not None else None, pl.size // 3 if pl is not None else 0,
            tr.ctypes.data if tr is not None else None, tr.size if tr is not None else 0, mode, out.handle, frame))

    def unfilter_expand_batch(self, width, height, bit_depth, color_type, interlace, streams, mode, out, frame0=0,
                              plte=None, trns=None):
        """streams: inflated IDAT streams of identical geometry; stream i -> frame frame0 + i of out."""
        pairs = [self._inflated(width, height, bit_depth, color_type, interlace, st) for st in streams]
        arr = (NormInflatedDesc * max(len(pairs), 1))(*[d for d, _ in pairs])
        pl = None if plte is None or len(plte) == 0 else np.ascontiguousarray(plte, np.uint8)
        tr = None if trns is None or len(trns) == 0 else np.ascontiguousarray(trns, np.uint8)
        self._ok(self.lib.norm_cuda_unfilter_expand_rgba8_batch(
            self.h, arr, len(pairs), pl.ctypes.data if pl is not None else None, pl.size // 3 if pl is not None else 0,
            tr.ctypes.data if tr is not None else None, tr.size if tr is not None else 0, mode, out.handle, frame0))

    # -- encoder-side scanline filtering (ScanlineWriter::write on the GPU)
    def filter_scanlines(self, width, height, bit_depth, color_type, pixel_buffer, paeth_compat=0):
        """pixel_buffer -> (stream of `filter byte + filtered row`, chosen filter per row)."""
        ch = {0: 1, 2: 3, 3: 1, 4: 2, 6: 4}[color_type]
        rowb = (width * ch * bit_depth + 7) // 8
        px = np.ascontiguousarray(pixel_buffer, np.uint8).reshape(-1)
        assert px.size == rowb * height
        stream, ft = np.empty(height * (rowb + 1), np.uint8), np.empty(height, np.uint8)
        self._ok(self.lib.norm_cuda_filter_scanlines(self.h, width, height, bit_depth, color_type, px.ctypes.data,
                                                     paeth_compat, stream.ctypes.data, ft.ctypes.data))
        return stream, ft

    def filter_image(self, img, frame=0, paeth_compat=0):
        stream, ft = np.empty(img.height * (img.width * 4 + 1), np.uint8), np.empty(img.height, np.uint8)
        self._ok(self.lib.norm_cuda_filter_image(self.h, img.handle, frame, paeth_compat, stream.ctypes.data, ft.ctypes.data))
        return stream, ft

    # -- effects
    @staticmethod
    def _raw(f): return C.byref(f.raw if isinstance(f, FeatureUniform) else f)

    def effects(self, f, src, dst, flags=NORM_CS_REFERENCE):
        self._ok(self.lib.norm_cuda_effects(self.h, self._raw(f), flags, src.handle, dst.handle))

    def chain(self, f, src, dst, flags=NORM_CS_REFERENCE, want_ssim=True):
        s = np.zeros(src.frames, np.float64) if want_ssim else None
        self._ok(self.lib.norm_cuda_chain(self.h, self._raw(f), flags, src.handle, dst.handle,
                                          s.ctypes.data_as(C.POINTER(C.c_double)) if want_ssim else None))
        return s

    def chain_async(self, f, src, dst, flags=NORM_CS_REFERENCE, ssim_dev=None):
        self._ok(self.lib.norm_cuda_chain_async(self.h, self._raw(f), flags, src.handle, dst.handle,
                                                C.c_void_p(ssim_dev) if ssim_dev else None))

    def chain_host(self, f, host_in, host_out, frames, flags=NORM_CS_REFERENCE, want_ssim=True):
        s = np.zeros(frames, np.float64) if want_ssim else None
        self._ok(self.lib.norm_cuda_chain_host(self.h, self._raw(f), flags, host_in.ctypes.data, host_out.ctypes.data,
                                               frames, s.ctypes.data_as(C.POINTER(C.c_double)) if want_ssim else None))
        return s

    def chain_band(self, f, src, dst, own_row0, own_rows, flags=NORM_CS_REFERENCE, want_moments=True):
        """Chain on band + halo; returns the SSIM moments of the owned rows (frames x 5)."""
        m = np.zeros((src.frames, 5), np.float64) if want_moments else None
        self._ok(self.lib.norm_cuda_chain_band(self.h, self._raw(f), flags, src.handle, dst.handle, own_row0, own_rows,
                                               m.ctypes.data_as(C.POINTER(C.c_double)) if want_moments else None))
        return m

    def chain_band_async(self, f, src, dst, own_row0, own_rows, flags=NORM_CS_REFERENCE, moments_dev=None):
        self._ok(self.lib.norm_cuda_chain_band_async(self.h, self._raw(f), flags, src.handle, dst.handle, own_row0,
                                                     own_rows, C.c_void_p(moments_dev) if moments_dev else None))

    def stream(self):
        p = C.c_void_p()
        self._ok(self.lib.norm_cuda_stream(self.h, C.byref(p)))
        return p.value or 0

    def ipc_export(self, img):
        h = (C.c_uint8 * 64)()
        self._ok(self.lib.norm_cuda_ipc_export(self.h, img.handle, h))
        return bytes(h)

    def ipc_open(self, handle):
        p = C.c_void_p()
        buf = (C.c_uint8 * 64).from_buffer_copy(handle)
        self._ok(self.lib.norm_cuda_ipc_open(self.h, buf, C.byref(p)))
        return p.value

    def ipc_close(self, ptr): self._ok(self.lib.norm_cuda_ipc_close(self.h, C.c_void_p(ptr)))

    def copy_rows(self, dst, dst_row, src_ptr, src_pitch, src_row, nrows):
        self._ok(self.lib.norm_cuda_copy_rows(self.h, dst.handle, dst_row, C.c_void_p(src_ptr), src_pitch, src_row, nrows))

    def front(self, f, src, dst, flags=NORM_CS_REFERENCE):
        self._ok(self.lib.norm_cuda_front(self.h, self._raw(f), flags, src.handle, dst.handle))

    def blur(self, radius, src, dst): self._ok(self.lib.norm_cuda_blur(self.h, radius, src.handle, dst.handle))
    def sobel(self, src, dst): self._ok(self.lib.norm_cuda_sobel(self.h, src.handle, dst.handle))
    def sharpen(self, factor, src, dst): self._ok(self.lib.norm_cuda_sharpen(self.h, factor, src.handle, dst.handle))

    def box_blur(self, src, dst): self._ok(self.lib.norm_cuda_box_blur(self.h, src.handle, dst.handle))

    def overlay(self, src, dst, circles=(), selected=None, draw=None):
        """Shape pass + composite + crosshair (shape_shader.wgsl:41-89, image_shader.wgsl:176-193).
        circles: Circle objects in normalised space (as EditorState holds them); selected: index
        or None (ShapeUniform::set_selected_circle); draw: DrawUniform or None."""
        su = ShapeUniform(src.width, src.height)
        su.set_num_circles(len(circles))
        su.set_selected_circle(selected)
        arr = (NormCircleData * max(len(circles), 1))(*[c.data() for c in circles])
        self._ok(self.lib.norm_cuda_overlay(self.h, C.byref(su.raw), arr, C.byref(draw.raw) if draw else None,
                                            src.handle, dst.handle))

    def ssim(self, a, b):
        s = np.zeros(a.frames, np.float64)
        self._ok(self.lib.norm_cuda_ssim(self.h, a.handle, b.handle, s.ctypes.data_as(C.POINTER(C.c_double))))
        return s

    def ssim_luma(self, a, luma_a, b, luma_b):
        s = np.zeros(a.frames, np.float64)
        self._ok(self.lib.norm_cuda_ssim_luma(self.h, a.handle, luma_a, b.handle, luma_b, s.ctypes.data_as(C.POINTER(C.c_double))))
        return s

    def ssim_moments(self, a, b):
        m = np.zeros((a.frames, 5), np.float64)
        self._ok(self.lib.norm_cuda_ssim_moments(self.h, a.handle, b.handle, m.ctypes.data_as(C.POINTER(C.c_double))))
        return m


class Band:
    """One rank's share of a row-band sharded image (norm_cuda_band_*, include/norm_cuda.h): the halo exchange over
    NVLink, the overlapped interior / edge launches and the SSIM moments behind one handle."""

    def __init__(self, ctx, f, width, height, nranks, rank):
        self.ctx, self.lib = ctx, ctx.lib
        h = C.c_void_p()
        ctx._ok(self.lib.norm_cuda_band_create(ctx.h, Context._raw(f), width, height, nranks, rank, C.byref(h)))
        self.h, self.width, self.height, self.nranks, self.rank = h, width, height, nranks, rank
        a, b, c, d = C.c_uint32(), C.c_uint32(), C.c_uint32(), C.c_uint32()
        ctx._ok(self.lib.norm_cuda_band_layout(h, a, b, c, d))
        self.row0, self.rows, self.halo_up, self.halo_down = a.value, b.value, c.value, d.value
        self.src = DeviceImage(ctx, C.c_void_p(self.lib.norm_cuda_band_src(h)))
        self.dst = DeviceImage(ctx, C.c_void_p(self.lib.norm_cuda_band_dst(h)))

    def export(self):
        hd = _lib.NormBandHandle()
        self.ctx._ok(self.lib.norm_cuda_band_export(self.h, C.byref(hd)))
        return bytes(hd)

    def connect(self, handles):
        """handles: the export() of every rank, ordered by rank."""
        arr = (_lib.NormBandHandle * len(handles))(*[_lib.NormBandHandle.from_buffer_copy(x) for x in handles])
        self.ctx._ok(self.lib.norm_cuda_band_connect(self.h, arr, len(handles)))

    def begin_update(self): self.ctx._ok(self.lib.norm_cuda_band_begin_update(self.h))
    def publish(self): self.ctx._ok(self.lib.norm_cuda_band_publish(self.h))

    def step_async(self, flags=NORM_CS_REFERENCE, moments_dev=None):
        self.ctx._ok(self.lib.norm_cuda_band_step_async(self.h, flags, C.c_void_p(moments_dev) if moments_dev else None))

    def side_stream(self):
        p = C.c_void_p()
        self.ctx._ok(self.lib.norm_cuda_band_side_stream(self.h, C.byref(p)))
        return p.value or 0

    def sync(self): self.ctx._ok(self.lib.norm_cuda_band_sync(self.h))

    def own_rows_of(self, img):
        """The owned rows of an extended band image (src or dst) as a (rows, width, 4) array."""
        return img.download()[0][self.halo_up:self.halo_up + self.rows]

    def destroy(self):
        if self.h:
            self.lib.norm_cuda_band_destroy(self.h)
            self.h = None
            self.src.handle = self.dst.handle = None  # owned by the band


class Texture:
    """texture.rs:22-80: `from_image` = rgba8() + upload; here both happen on the GPU."""

    def __init__(self, image):
        self.image = image

    @classmethod
    def from_image(cls, ctx, img, mode=NORM_EXPAND_COMPAT):
        w, h = img.dimensions()
        dev = ctx.image_create(w, h, 1)
        try:
            ctx.expand_rgba8(img, mode, dev, 0)
        except Exception:
            dev.destroy()
            raise
        return cls(dev)

    from_bytes = from_image  # texture.rs:18-20

    def destroy(self): self.image.destroy()


class EffectPipeline:
    """effect_pipeline.rs: ordered (effect, predicate) list over ping-pong textures A / B.

    `add_effect(label, shader, uniform, condition)`: `shader` names one of the reference's
    compute shaders -- "gamma", "grayscale", "invert", "copy" -- instead of carrying WGSL.
    `execute(feature_uniform)` runs each enabled effect as its own pass with the reference's
    texture routing (#0: input->A, odd: A->B, even: B->A, effect_pipeline.rs:105-111);
    `render` is the whole frame (execute + the fragment effect fs_main selects) in fused form.
    """

    def __init__(self, ctx, input_texture, texture_a, texture_b, effects, flags):
        self.ctx, self.input, self._a, self._b, self.effects, self.flags = ctx, input_texture, texture_a, texture_b, effects, flags

    @staticmethod
    def builder(ctx, input_texture, width, height, flags=NORM_CS_REFERENCE):
        return EffectPipelineBuilder(ctx, input_texture, width, height, flags)

    def texture_a(self): return self._a

    def update_effect_uniform(self, index, data):  # effect_pipeline.rs:53-62
        if 0 <= index < len(self.effects):
            self.effects[index]["uniform"] = data

    def execute(self, feature_uniform):
        w, h = self._a.image.width, self._a.image.height
        for e in self.effects:
            if not e["condition"](feature_uniform):
                continue
            f = FeatureUniform(w, h, 0)
            flags = NORM_CS_RAW  # only the gamma pass reads the sRGB input texture
            if e["shader"] == "gamma":
                f.raw.gamma = int(e["uniform"] or 0)
                flags = self.flags
            elif e["shader"] == "grayscale":
                f.raw.grayscale = 1
            elif e["shader"] == "invert":
                f.raw.invert = 1
            self.ctx.front(f, e["src"].image, e["dst"].image, flags)

    def render(self, feature_uniform, out=None):
        out = out or self.ctx.image_create(self.input.image.width, self.input.image.height, self.input.image.frames)
        self.ctx.effects(feature_uniform, self.input.image, out, self.flags)
        return out


class EffectPipelineBuilder:
    SHADERS = ("gamma", "grayscale", "invert", "copy")

    def __init__(self, ctx, input_texture, width, height, flags):
        self.ctx, self.input, self.flags = ctx, input_texture, flags
        self.a = Texture(ctx.image_create(width, height, 1))
        self.b = Texture(ctx.image_create(width, height, 1))
        self.effects = []

    def add_effect(self, label, shader, uniform, condition):
        if shader not in self.SHADERS:
            raise NormError(_lib.NORM_E_UNSUPPORTED, f"unknown compute effect {shader!r}")
        n = len(self.effects)
        src, dst = (self.input, self.a) if n == 0 else ((self.a, self.b) if n % 2 == 1 else (self.b, self.a))
        self.effects.append({"label": label, "shader": shader, "uniform": uniform, "condition": condition,
                             "src": src, "dst": dst})
        return self

    def build(self):
        return EffectPipeline(self.ctx, self.input, self.a, self.b, self.effects, self.flags)


def reference_pipeline(ctx, input_texture, feature_uniform, flags=NORM_CS_REFERENCE):
    """The wiring of AppState::new (app_state.rs:89-125)."""
    w, h = input_texture.image.width, input_texture.image.height
    return (EffectPipeline.builder(ctx, input_texture, w, h, flags)
            .add_effect("gamma", "gamma", feature_uniform.gamma(), lambda f: True)
            .add_effect("grayscale", "grayscale", None, lambda f: f.grayscale())
            .add_effect("copy_after_grayscale", "copy", None, lambda f: f.grayscale())
            .add_effect("invert", "invert", None, lambda f: f.invert())
            .add_effect("copy_after_invert", "copy", None, lambda f: f.invert())
            .build())


# -- host-only helpers -------------------------------------------------------------------

def blur_taps(radius):
    taps, dmin = np.zeros(radius, np.float32), C.c_int32()
    st = _lib.load().norm_blur_taps(radius, taps.ctypes.data_as(C.POINTER(C.c_float)), dmin)
    if st != NORM_OK:
        raise NormError(st, "blur radius outside 1..63")
    return taps, dmin.value


def srgb_tables():
    eotf, thresh = np.zeros(256, np.float32), np.zeros(256, np.float32)
    _lib.load().norm_srgb_tables(eotf.ctypes.data_as(C.POINTER(C.c_float)), thresh.ctypes.data_as(C.POINTER(C.c_float)))
    return eotf, thresh


def front_lut(flags, gamma):
    lut = np.zeros(256, np.uint8)
    _lib.load().norm_front_lut(flags, gamma, lut.ctypes.data_as(C.POINTER(C.c_uint8)))
    return lut


def ssim_from_moments(m, npix):
    m = np.ascontiguousarray(m, np.float64)
    return float(_lib.load().norm_ssim_from_moments(m.ctypes.data_as(C.POINTER(C.c_double)), float(npix)))


def plan_frames(frames, nranks, rank):
    a, b = C.c_uint32(), C.c_uint32()
    st = _lib.load().norm_plan_frames(frames, nranks, rank, a, b)
    if st != NORM_OK:
        raise NormError(st, "bad frame plan arguments")
    return a.value, b.value


def plan_band(f, height, nranks, rank):
    r0, rows, up, down = C.c_uint32(), C.c_uint32(), C.c_uint32(), C.c_uint32()
    st = _lib.load().norm_plan_band(Context._raw(f), height, nranks, rank, r0, rows, up, down)
    if st != NORM_OK:
        raise NormError(st, "bad band plan arguments")
    return r0.value, rows.value, up.value, down.value
