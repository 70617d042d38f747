"""soil.tiff / soil.geotiff / soil.geotiff_meta - the raster IO either side of the flow path (SURVEY.md section 8f row 2;
reference io/tiff.hpp:69-241, io/geotiff.hpp:131-314, bindings python/source/io.cpp:20-101).

The reference goes through libtiff (an un-vendored submodule, absent here).  This is a small own codec for what its
reader and writer handle: classic (non-Big) TIFF, either byte order, ONE float32 or float64 sample per pixel, stored in
strips or in tiles, uncompressed or Deflate-compressed (zlib, with or without the horizontal / floating-point
predictor - what GDAL writes for DEMs); LZW and the rest raise.  The writer emits what tiff::write emits: uncompressed
strips, top-left orientation, min-is-black, IEEE samples.  GeoTIFF adds the tags of geotiff.hpp:12-20 (33550 pixel
scale, 33922 tie points, 34735/6/7 geo keys, 42112/3 GDAL metadata / nodata) and the nodata -> NaN substitution.

Layout: the tensor is (height, width) - dim 0 is the image row, dim 1 the column - so that
`tensor.numpy()[row, col]` is the pixel, and `shape` is that shape.  Tensors are host tensors; `.gpu()` moves them.
"""
from __future__ import annotations

import struct
import zlib
from typing import Dict, Optional, Tuple

import numpy as np

from . import silt

# tag ids
IMAGEWIDTH, IMAGELENGTH, BITSPERSAMPLE, COMPRESSION, PHOTOMETRIC = 256, 257, 258, 259, 262
STRIPOFFSETS, ORIENTATION, SAMPLESPERPIXEL, ROWSPERSTRIP, STRIPBYTECOUNTS = 273, 274, 277, 278, 279
PLANARCONFIG, PREDICTOR, TILEWIDTH, TILELENGTH, TILEOFFSETS, TILEBYTECOUNTS, SAMPLEFORMAT = 284, 317, 322, 323, 324, 325, 339
GEOPIXELSCALE, GEOTIEPOINTS, GEOKEYDIRECTORY, GEODOUBLEPARAMS, GEOASCIIPARAMS = 33550, 33922, 34735, 34736, 34737
GDAL_METADATA, GDAL_NODATA = 42112, 42113

_TYPES = {1: ("B", 1), 2: ("c", 1), 3: ("H", 2), 4: ("I", 4), 5: ("II", 8), 6: ("b", 1), 7: ("B", 1), 8: ("h", 2),
          9: ("i", 4), 10: ("ii", 8), 11: ("f", 4), 12: ("d", 8), 16: ("Q", 8)}


class missing_file(FileNotFoundError):
    """silt::error::missing_file (tiff.hpp:73)."""


def _read_ifd(buf: bytes) -> Tuple[str, Dict[int, tuple]]:
    """First image file directory of a classic TIFF: {tag: tuple of values} (ASCII tags as one str)."""
    if len(buf) < 8 or buf[:2] not in (b"II", b"MM"):
        raise ValueError("not a TIFF file")
    e = "<" if buf[:2] == b"II" else ">"
    magic, off = struct.unpack(e + "HI", buf[2:8])
    if magic == 43:
        raise ValueError("BigTIFF is not supported")
    if magic != 42:
        raise ValueError("not a TIFF file")
    (n,) = struct.unpack(e + "H", buf[off:off + 2])
    tags = {}
    for i in range(n):
        tag, typ, count, raw = struct.unpack(e + "HHI4s", buf[off + 2 + 12 * i: off + 14 + 12 * i])
        if typ not in _TYPES:
            continue
        fmt, size = _TYPES[typ]
        nbytes = size * count
        data = raw[:nbytes] if nbytes <= 4 else buf[struct.unpack(e + "I", raw)[0]:][:nbytes]
        if typ == 2:
            tags[tag] = (data.split(b"\0")[0].decode("latin-1"),)
        elif typ in (5, 10):
            v = struct.unpack(e + fmt[0] * (2 * count), data)
            tags[tag] = tuple(v[2 * j] / v[2 * j + 1] if v[2 * j + 1] else 0.0 for j in range(count))
        else:
            tags[tag] = struct.unpack(e + fmt * count, data)
    return e, tags


def _undo_predictor(a: np.ndarray, predictor: int, dtype: np.dtype, rows: int, cols: int, byteorder: str) -> np.ndarray:
    """`a`: the decompressed bytes of rows x cols samples."""
    if predictor == 1:
        return a
    size = dtype.itemsize
    if predictor == 2:      # horizontal differencing on the integer view of the samples
        v = a.view(np.dtype(("u%d" % size)).newbyteorder(byteorder)).reshape(rows, cols)
        return np.cumsum(v, axis=1, dtype=v.dtype).astype(v.dtype).view(np.uint8).reshape(-1)
    if predictor == 3:      # floating-point predictor: byte planes (most significant first), differenced per row
        b = a.reshape(rows, cols * size)
        b = np.cumsum(b, axis=1, dtype=np.uint8)
        planes = b.reshape(rows, size, cols)                 # plane p holds byte (size-1-p) of a little-endian sample
        order = range(size - 1, -1, -1) if byteorder == "<" else range(size)
        out = np.empty((rows, cols, size), np.uint8)
        for k, p in enumerate(order):
            out[:, :, k] = planes[:, p, :]
        return out.reshape(-1)
    raise ValueError(f"TIFF predictor {predictor} is not supported")


def _decode(chunk: bytes, compression: int) -> bytes:
    if compression == 1:
        return chunk
    if compression in (8, 32946):
        return zlib.decompress(chunk)
    names = {5: "LZW", 7: "JPEG", 32773: "PackBits", 50000: "ZSTD", 34925: "LZMA"}
    raise ValueError(f"TIFF compression {names.get(compression, compression)} is not supported (uncompressed and Deflate are)")


def _read_raster(buf: bytes, e: str, tags: Dict[int, tuple]) -> np.ndarray:
    width, height = int(tags[IMAGEWIDTH][0]), int(tags[IMAGELENGTH][0])
    bits = int(tags.get(BITSPERSAMPLE, (1,))[0])
    if int(tags.get(SAMPLESPERPIXEL, (1,))[0]) != 1:
        raise ValueError("only one sample per pixel is supported")
    fmt = int(tags.get(SAMPLEFORMAT, (1,))[0])
    if bits not in (32, 64) or fmt != 3:
        raise ValueError(f"only 32- and 64-bit IEEE samples are supported (got {bits} bits, sample format {fmt})")
    dt = np.dtype("f%d" % (bits // 8)).newbyteorder(e)
    comp = int(tags.get(COMPRESSION, (1,))[0])
    pred = int(tags.get(PREDICTOR, (1,))[0])
    out = np.zeros((height, width), dt.newbyteorder("="))
    if TILEWIDTH in tags:                                    # tiff.hpp:161-205
        tw, th = int(tags[TILEWIDTH][0]), int(tags[TILELENGTH][0])
        across = (width + tw - 1) // tw
        for t, (off, cnt) in enumerate(zip(tags[TILEOFFSETS], tags[TILEBYTECOUNTS])):
            raw = np.frombuffer(_decode(buf[off:off + cnt], comp), np.uint8)[: tw * th * dt.itemsize]
            tile = np.frombuffer(_undo_predictor(raw, pred, dt, th, tw, e).tobytes(), dt).reshape(th, tw)
            r0, c0 = (t // across) * th, (t % across) * tw
            out[r0:r0 + th, c0:c0 + tw] = tile[: height - r0, : width - c0]
    else:                                                    # tiff.hpp:126-159
        rps = min(int(tags.get(ROWSPERSTRIP, (height,))[0]), height)
        counts = tags.get(STRIPBYTECOUNTS)
        for s, off in enumerate(tags[STRIPOFFSETS]):
            r0 = s * rps
            rows = min(rps, height - r0)
            cnt = counts[s] if counts else rows * width * dt.itemsize
            raw = np.frombuffer(_decode(buf[off:off + cnt], comp), np.uint8)[: rows * width * dt.itemsize]
            out[r0:r0 + rows] = np.frombuffer(_undo_predictor(raw, pred, dt, rows, width, e).tobytes(), dt).reshape(rows, width)
    return out


def _write_tiff(filename: str, a: np.ndarray, extra: Dict[int, tuple]) -> None:
    """Uncompressed strips, as tiff::write / geotiff::write (tiff.hpp:213-238, geotiff.hpp:191-240).
    extra: {tag: (tiff type, values)} for the GeoTIFF tags."""
    a = np.ascontiguousarray(a)
    if a.ndim != 2 or a.dtype not in (np.float32, np.float64):
        raise ValueError("tiff.write expects a 2-D float32 or float64 tensor")
    height, width = a.shape
    bits = a.dtype.itemsize * 8
    rps = max(1, min(height, 8192 // max(1, width * a.dtype.itemsize)))      # TIFFDefaultStripSize: ~8 KiB strips
    nstrips = (height + rps - 1) // rps
    data = a.astype(a.dtype.newbyteorder("<")).tobytes()
    entries = {IMAGEWIDTH: (4, (width,)), IMAGELENGTH: (4, (height,)), BITSPERSAMPLE: (3, (bits,)), COMPRESSION: (3, (1,)),
               PHOTOMETRIC: (3, (1,)), ORIENTATION: (3, (1,)), SAMPLESPERPIXEL: (3, (1,)), ROWSPERSTRIP: (4, (rps,)),
               PLANARCONFIG: (3, (1,)), SAMPLEFORMAT: (3, (3,))}
    row_bytes = width * a.dtype.itemsize
    entries[STRIPBYTECOUNTS] = (4, tuple(min(rps, height - s * rps) * row_bytes for s in range(nstrips)))
    entries[STRIPOFFSETS] = (4, tuple(8 + s * rps * row_bytes for s in range(nstrips)))
    entries.update(extra)
    ifd_off = 8 + len(data)
    ifd_off += ifd_off & 1
    tags = sorted(entries)
    blob, table = b"", b""
    blob_off = ifd_off + 2 + 12 * len(tags) + 4
    for tag in tags:
        typ, vals = entries[tag]
        if typ == 2:
            payload = vals[0].encode("latin-1") + b"\0"
            count = len(payload)
        else:
            payload = struct.pack("<" + _TYPES[typ][0] * len(vals), *vals)
            count = len(vals)
        if len(payload) <= 4:
            table += struct.pack("<HHI4s", tag, typ, count, payload.ljust(4, b"\0"))
        else:
            table += struct.pack("<HHII", tag, typ, count, blob_off + len(blob))
            blob += payload + (b"\0" if len(payload) & 1 else b"")
    with open(filename, "wb") as f:
        f.write(b"II" + struct.pack("<HI", 42, ifd_off))
        f.write(data)
        f.write(b"\0" * (ifd_off - 8 - len(data)))
        f.write(struct.pack("<H", len(tags)) + table + struct.pack("<I", 0) + blob)


class tiff:
    """soil.tiff (io.cpp:20-36): tiff(), tiff(filename) or tiff(tensor); peek / read / write; width, height, bits,
    tensor, shape."""

    def __init__(self, arg=None):
        self.filename = ""
        self._width = self._height = 0
        self._bits = 32
        self._tensor: Optional[silt.tensor] = None
        self._tags: Dict[int, tuple] = {}
        if isinstance(arg, silt.tensor):                    # tiff.hpp:31-45
            t = arg.cpu() if arg.host is silt.gpu else arg
            if t.shape.dim() != 2:
                raise ValueError("tiff: expected a 2-D tensor")
            self._tensor = t
            self._height, self._width = int(t.shape[0]), int(t.shape[1])
            self._bits = 64 if t.type is silt.float64 else 32
        elif arg is not None:
            self.read(arg)

    def _open(self, filename) -> bytes:
        try:
            with open(filename, "rb") as f:
                return f.read()
        except OSError as exc:
            raise missing_file(str(filename)) from exc

    def peek(self, filename) -> bool:
        e, tags = _read_ifd(self._open(filename))
        self._width, self._height = int(tags[IMAGEWIDTH][0]), int(tags[IMAGELENGTH][0])
        self._bits = int(tags.get(BITSPERSAMPLE, (1,))[0])
        self._tags, self.filename = tags, str(filename)
        return True

    def read(self, filename) -> bool:
        buf = self._open(filename)
        self.peek(filename)
        e, tags = _read_ifd(buf)
        self._tensor = silt.tensor.from_numpy(_read_raster(buf, e, tags))
        return True

    def write(self, filename) -> bool:
        if self._tensor is None:
            raise ValueError("tiff.write: no tensor")
        _write_tiff(str(filename), self._tensor.numpy(), {})
        return True

    width = property(lambda self: self._width)
    height = property(lambda self: self._height)
    bits = property(lambda self: self._bits)
    tensor = property(lambda self: self._tensor)
    shape = property(lambda self: None if self._tensor is None else self._tensor.shape)


class geotiff_meta:
    """soil.geotiff_meta (geotiff.hpp:88-106, io.cpp:65-95)."""

    def __init__(self):
        self.filename = ""
        self.width = self.height = 0
        self.bits = 32
        self.gdal_nodata = ""
        self.gdal_metadata = ""
        self.gdal_ascii = ""                                   # geoasciiparams
        self.scale = [1.0, 1.0, 1.0]
        self.coords = [0.0, 0.0, 0.0, 0.0, 0.0, 0.0]
        self.params: list = []
        self.keydir: list = []

    def _corners(self):
        o = np.array([self.coords[3], self.coords[4]], np.float32)
        f = o + np.array(self.scale[:2], np.float32) * np.array([self.width, self.height], np.float32)
        return o, f

    @property
    def min(self):
        o, f = self._corners()
        return tuple(np.minimum(o, f).tolist())

    @min.setter
    def min(self, v):                                          # io.cpp:86-89
        self.coords[3], self.coords[4] = float(v[0]), float(v[1])

    @property
    def max(self):
        o, f = self._corners()
        return tuple(np.maximum(o, f).tolist())


class geotiff(tiff):
    """soil.geotiff (io.cpp:40-63): tiff + GeoTIFF / GDAL tags; nodata values become NaN on read (geotiff.hpp:242-276)."""

    def __init__(self, arg=None):
        self.meta = geotiff_meta()
        super().__init__(arg)
        if isinstance(arg, silt.tensor):                       # geotiff.hpp:71-75
            self.meta.coords[3], self.meta.coords[4] = float(self._tensor.shape[0]), float(self._tensor.shape[1])
            self.meta.width, self.meta.height, self.meta.bits = self._width, self._height, self._bits

    def peek(self, filename) -> bool:                          # geotiff.hpp:131-180
        super().peek(filename)
        m, t = self.meta, self._tags
        m.filename, m.width, m.height, m.bits = self.filename, self._width, self._height, self._bits
        if GDAL_NODATA in t:
            m.gdal_nodata = t[GDAL_NODATA][0]
        if GDAL_METADATA in t:
            m.gdal_metadata = t[GDAL_METADATA][0]
        if GEOASCIIPARAMS in t:
            m.gdal_ascii = t[GEOASCIIPARAMS][0]
        if GEOPIXELSCALE in t:
            m.scale = [float(v) for v in t[GEOPIXELSCALE]]
            if len(m.scale) > 2 and m.scale[2] == 0.0:
                m.scale[2] = 1.0
        if GEOTIEPOINTS in t:
            m.coords = [float(v) for v in t[GEOTIEPOINTS]]
        if GEODOUBLEPARAMS in t:
            m.params = [float(v) for v in t[GEODOUBLEPARAMS]]
        if GEOKEYDIRECTORY in t:
            m.keydir = [int(v) for v in t[GEOKEYDIRECTORY]]
        return True

    def read(self, filename) -> bool:                          # geotiff.hpp:182-189
        super().read(filename)
        self._set_nan()
        return True

    def write(self, filename) -> bool:                         # geotiff.hpp:191-240
        if self._tensor is None:
            raise ValueError("geotiff.write: no tensor")
        m, extra = self.meta, {}
        if m.scale:
            extra[GEOPIXELSCALE] = (12, tuple(float(v) for v in m.scale))
        if m.coords:
            extra[GEOTIEPOINTS] = (12, tuple(float(v) for v in m.coords))
        if m.params:
            extra[GEODOUBLEPARAMS] = (12, tuple(float(v) for v in m.params))
        if m.keydir:
            extra[GEOKEYDIRECTORY] = (3, tuple(int(v) & 0xFFFF for v in m.keydir))
        if m.gdal_nodata:
            extra[GDAL_NODATA] = (2, (m.gdal_nodata,))
        if m.gdal_metadata:
            extra[GDAL_METADATA] = (2, (m.gdal_metadata,))
        if m.gdal_ascii:
            extra[GEOASCIIPARAMS] = (2, (m.gdal_ascii,))
        _write_tiff(str(filename), self._tensor.numpy(), extra)
        return True

    def _nodata(self):
        if not self.meta.gdal_nodata:
            return None
        v = float(self.meta.gdal_nodata)                       # std::stof / std::stod
        return np.float32(v) if self._bits != 64 else np.float64(v)

    def _set_nan(self):                                        # geotiff.hpp:242-276
        nd = self._nodata()
        if nd is not None and self._tensor is not None:
            a = self._tensor.data
            a[a == float(nd)] = float("nan")

    def unsetnan(self):
        """NaN -> the file's nodata value.  (The reference's unsetNaN compares `buffer[i] == nan`, geotiff.hpp:285-311,
        which is never true, so there it is a no-op; this does what its name and comment say.)"""
        nd = self._nodata()
        if nd is not None and self._tensor is not None:
            a = self._tensor.data
            a[a != a] = float(nd)

    @property
    def scale(self):
        return (float(np.float32(self.meta.scale[0])), float(np.float32(self.meta.scale[1])))

    def _corners(self):
        o = np.array([self.meta.coords[3], self.meta.coords[4]], np.float32)
        f = o + np.array(self.meta.scale[:2], np.float32) * np.array([self._width, self._height], np.float32)
        return o, f

    @property
    def min(self):
        o, f = self._corners()
        return tuple(np.minimum(o, f).tolist())

    @property
    def max(self):
        o, f = self._corners()
        return tuple(np.maximum(o, f).tolist())
