"""soil.tiff / soil.geotiff (soillib_b200/io.py; reference io/tiff.hpp, io/geotiff.hpp): round trips, files written by an
independent encoder (Pillow, and hand-assembled tiled / big-endian / Deflate+predictor files), GeoTIFF tags, nodata."""
import struct
import zlib

import numpy as np
import pytest

import soillib_b200 as soil
from soillib_b200 import silt
from soillib_b200.io import missing_file

PIL = pytest.importorskip("PIL.Image")


def ramp(h, w, dtype=np.float32):
    rng = np.random.default_rng(h * 100 + w)
    return (rng.standard_normal((h, w)) * 100.0).astype(dtype)


@pytest.mark.parametrize("shape", [(1, 1), (7, 5), (64, 48), (300, 1000)])
@pytest.mark.parametrize("dtype", [np.float32, np.float64])
def test_write_read_round_trip(tmp_path, shape, dtype):
    a = ramp(*shape, dtype)
    path = tmp_path / "a.tif"
    t = soil.tiff(silt.tensor.from_numpy(a))
    assert (t.height, t.width, t.bits) == (shape[0], shape[1], 8 * a.itemsize) and tuple(t.shape) == shape
    assert t.write(str(path))
    back = soil.tiff(str(path))
    assert (back.height, back.width, back.bits) == (shape[0], shape[1], 8 * a.itemsize)
    assert back.tensor.numpy().dtype == dtype and np.array_equal(back.tensor.numpy(), a)
    if dtype == np.float32:                                    # an independent decoder reads what we wrote
        with PIL.open(path) as im:
            assert im.mode == "F" and np.array_equal(np.asarray(im), a)


def test_reads_files_written_by_pillow(tmp_path):
    a = ramp(33, 70)
    for comp in (None, "tiff_adobe_deflate"):
        path = tmp_path / f"p_{comp}.tif"
        try:
            PIL.fromarray(a).save(path, compression=comp)
        except Exception:                                       # this Pillow build has no libtiff encoder
            continue
        t = soil.tiff()
        assert t.peek(str(path)) and (t.width, t.height, t.bits) == (70, 33, 32) and t.tensor is None
        assert t.read(str(path)) and np.array_equal(t.tensor.numpy(), a)


def assemble(a, endian="<", tile=None, deflate=False, predictor=1, extra=()):
    """A classic TIFF built by hand: strips of 3 rows or tiles of `tile` (th, tw), optional Deflate + predictor."""
    h, w = a.shape
    e, size = endian, a.itemsize
    raw = a.astype(a.dtype.newbyteorder(e))

    def encode(block):
        rows, cols = block.shape
        b = np.frombuffer(block.tobytes(), np.uint8).reshape(rows, cols, size)
        if predictor == 3:
            order = range(size - 1, -1, -1) if e == "<" else range(size)
            planes = np.stack([b[:, :, k] for k in order], axis=1).reshape(rows, cols * size)
            d = planes.copy()
            d[:, 1:] = planes[:, 1:] - planes[:, :-1]
            b = d
        elif predictor == 2:
            v = block.view(np.dtype("u%d" % size).newbyteorder(e))
            d = v.copy()
            d[:, 1:] = v[:, 1:] - v[:, :-1]
            b = np.frombuffer(d.tobytes(), np.uint8)
        payload = b.tobytes()
        return zlib.compress(payload) if deflate else payload

    chunks = []
    if tile:
        th, tw = tile
        for r0 in range(0, h, th):
            for c0 in range(0, w, tw):
                blk = np.zeros((th, tw), raw.dtype)
                part = raw[r0:r0 + th, c0:c0 + tw]
                blk[:part.shape[0], :part.shape[1]] = part
                chunks.append(encode(blk))
    else:
        for r0 in range(0, h, 3):
            chunks.append(encode(np.ascontiguousarray(raw[r0:r0 + 3])))
    offs, pos = [], 8
    for c in chunks:
        offs.append(pos)
        pos += len(c) + (len(c) & 1)
    entries = [(256, 4, (w,)), (257, 4, (h,)), (258, 3, (8 * size,)), (259, 3, (8 if deflate else 1,)), (262, 3, (1,)),
               (277, 3, (1,)), (339, 3, (3,)), (317, 3, (predictor,))]
    if tile:
        entries += [(322, 4, (tile[1],)), (323, 4, (tile[0],)), (324, 4, tuple(offs)), (325, 4, tuple(len(c) for c in chunks))]
    else:
        entries += [(278, 4, (3,)), (273, 4, tuple(offs)), (279, 4, tuple(len(c) for c in chunks))]
    entries += list(extra)
    entries.sort()
    fmt = {2: "c", 3: "H", 4: "I", 12: "d"}
    ifd = pos
    blob_off, table, blob = ifd + 2 + 12 * len(entries) + 4, b"", b""
    for tag, typ, vals in entries:
        payload = (vals[0].encode() + b"\0") if typ == 2 else struct.pack(e + fmt[typ] * len(vals), *vals)
        count = len(payload) if typ == 2 else len(vals)
        if len(payload) <= 4:
            table += struct.pack(e + "HHI", tag, typ, count) + payload.ljust(4, b"\0")
        else:
            table += struct.pack(e + "HHII", tag, typ, count, blob_off + len(blob))
            blob += payload + (b"\0" if len(payload) & 1 else b"")
    body = b"".join(c + (b"\0" if len(c) & 1 else b"") for c in chunks)
    return (b"II" if e == "<" else b"MM") + struct.pack(e + "HI", 42, ifd) + body + struct.pack(e + "H", len(entries)) + \
        table + struct.pack(e + "I", 0) + blob


@pytest.mark.parametrize("endian", ["<", ">"])
@pytest.mark.parametrize("kind", ["strips", "tiles", "deflate", "deflate+fp", "tiles+deflate+fp", "deflate+hdiff"])
@pytest.mark.parametrize("dtype", [np.float32, np.float64])
def test_reads_hand_assembled_variants(tmp_path, endian, kind, dtype):
    a = ramp(37, 53, dtype)
    tile = (16, 32) if "tiles" in kind else None
    pred = 3 if "fp" in kind else (2 if "hdiff" in kind else 1)
    path = tmp_path / "v.tif"
    path.write_bytes(assemble(a, endian, tile, "deflate" in kind, pred))
    got = soil.tiff(str(path)).tensor.numpy()
    assert got.dtype == dtype and np.array_equal(got, a)


def test_geotiff_tags_and_nodata(tmp_path):
    a = ramp(20, 30)
    a[3, 4] = a[10, 11] = -9999.0
    extra = [(33550, 12, (2.0, 0.5, 0.0)), (33922, 12, (0.0, 0.0, 0.0, 1000.0, 5000.0, 0.0)),
             (34735, 3, (1, 1, 0, 1, 1024, 0, 1, 1)), (34737, 2, ("WGS 84|",)), (42113, 2, ("-9999",)),
             (42112, 2, ("<GDALMetadata></GDALMetadata>",))]
    path = tmp_path / "g.tif"
    path.write_bytes(assemble(a, "<", None, True, 3, extra))
    g = soil.geotiff(str(path))
    m = g.meta
    assert (m.width, m.height, m.bits, m.filename) == (30, 20, 32, str(path))
    assert m.scale == [2.0, 0.5, 1.0]                           # a zero z-scale reads as 1 (geotiff.hpp:163-165)
    assert m.coords[3:5] == [1000.0, 5000.0] and m.keydir == [1, 1, 0, 1, 1024, 0, 1, 1]
    assert m.gdal_nodata == "-9999" and m.gdal_ascii == "WGS 84|" and m.gdal_metadata.startswith("<GDALMetadata")
    assert g.scale == (2.0, 0.5) and g.min == (1000.0, 5000.0) and g.max == (1060.0, 5010.0) and m.max == g.max
    got = g.tensor.numpy()
    assert np.isnan(got[3, 4]) and np.isnan(got[10, 11]) and np.isnan(got).sum() == 2      # nodata -> NaN
    keep = ~np.isnan(got)
    assert np.array_equal(got[keep], a[keep])
    # write it back: the tags survive, NaN goes back to nodata on request
    g.unsetnan()
    out = tmp_path / "g2.tif"
    assert g.write(str(out))
    h = soil.geotiff()
    assert h.peek(str(out)) and h.meta.scale == [2.0, 0.5, 1.0] and h.meta.coords == m.coords and h.meta.keydir == m.keydir
    assert h.meta.gdal_nodata == "-9999" and h.meta.gdal_ascii == "WGS 84|"
    h2 = soil.tiff(str(out))
    assert np.array_equal(h2.tensor.numpy(), a)
    # a geotiff made from a tensor carries its shape as the origin (geotiff.hpp:71-75)
    fresh = soil.geotiff(silt.tensor.from_numpy(a))
    assert fresh.meta.coords[3:5] == [20.0, 30.0] and fresh.meta.scale == [1.0, 1.0, 1.0]
    fresh.meta.min = (7.0, 8.0)
    assert fresh.meta.coords[3:5] == [7.0, 8.0]


def test_errors(tmp_path):
    with pytest.raises(missing_file):
        soil.tiff(str(tmp_path / "nope.tif"))
    (tmp_path / "junk.tif").write_bytes(b"not a tiff at all")
    with pytest.raises(ValueError):
        soil.tiff(str(tmp_path / "junk.tif"))
    a = ramp(4, 4)
    blob = bytearray(assemble(a))
    (tmp_path / "lzw.tif").write_bytes(bytes(blob).replace(struct.pack("<HHI4s", 259, 3, 1, struct.pack("<H", 1).ljust(4, b"\0")),
                                                           struct.pack("<HHI4s", 259, 3, 1, struct.pack("<H", 5).ljust(4, b"\0"))))
    with pytest.raises(ValueError, match="LZW"):
        soil.tiff(str(tmp_path / "lzw.tif"))
    with pytest.raises(ValueError):
        soil.tiff(silt.tensor.from_numpy(np.zeros((2, 2, 2), np.float32)))
    with pytest.raises(ValueError):
        soil.tiff().write(str(tmp_path / "x.tif"))
