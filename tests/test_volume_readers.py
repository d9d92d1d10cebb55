"""On-disk volume formats of the host mirror's vis::VolumeReader: .pvm (plain and DDS-compressed),
.pvmold, .nrrd / .nhrd, .dat, next to the .raw / .syn readers of test_host_mirror.py.  CPU only.  The .pvm reader is pinned
against the reference's own Pvm class (libs/file_utils/pvm.cpp compiled into oracle/_ref); the differential data
streams are produced by the encoder below, written from the stream layout the decoder defines."""
import os

import numpy as np
import pytest

from cpp_volume_rendering_b200 import hostapi
from oracle import pyref as R


class _Bits:
    def __init__(self):
        self.acc = 0; self.n = 0; self.out = bytearray()

    def put(self, value, bits):
        if bits == 0:
            return
        self.acc = (self.acc << bits) | (value & ((1 << bits) - 1)); self.n += bits
        while self.n >= 8:
            self.n -= 8
            self.out.append((self.acc >> self.n) & 0xFF)
        self.acc &= (1 << self.n) - 1

    def done(self):
        if self.n:
            self.put(0, 8 - self.n)
        while len(self.out) % 4:
            self.out.append(0)
        return bytes(self.out)


def dds_encode(data, skip, strip, version, rng, max_run=127):
    """bytes -> 'DDS v3d\\n' / 'DDS v3e\\n' stream: planes of every skip-th byte, then runs of residuals whose width is the
    smallest the run needs (or wider, at random: any width that holds the residuals is a valid stream)."""
    data = bytes(data)
    if skip > 1:                       # block interleave of v3e spans skip * 2^24 bytes: one block for test-sized data
        data = b"".join(data[i::skip] for i in range(skip))
    res = []
    prev = 0
    for cnt, v in enumerate(data):
        pred = prev if (strip == 1 or cnt <= strip) else prev + data[cnt - strip] - data[cnt - strip - 1]
        r = (v - pred) % 256
        res.append(r - 256 if r >= 128 else r)
        prev = v
    b = _Bits()
    b.put(skip - 1, 2); b.put(strip - 1, 16)
    i = 0
    while i < len(res):
        n = int(rng.integers(1, max_run + 1))
        run = res[i:i + n]
        width = 0
        for wd in (0, 2, 3, 4, 5, 6, 7, 8):
            half = (1 << wd) // 2
            if all(-half <= r <= half - 1 for r in run) if wd else all(r == 0 for r in run):
                width = wd
                break
        if width < 8 and rng.random() < 0.2:
            width = min(8, max(2, width + 1))
        b.put(len(run), 7); b.put(width - 1 if width else 0, 3)
        for r in run:
            b.put(r + (1 << width) // 2, width)
        i += len(run)
    b.put(0, 7)
    return (b"DDS v3d\n" if version == 1 else b"DDS v3e\n") + b.done()


def _pvm_bytes(vox, version, scale=(1.0, 1.0, 1.0), comment=False):
    d, h, w = vox.shape
    comps = vox.dtype.itemsize
    body = vox.astype("<u2").tobytes() if comps == 2 else vox.tobytes()       # Pvm::PostProcessData: low byte first
    if version == 1:
        head = b"PVM\n" + (b"# a comment line\n" if comment else b"") + f"{w} {h} {d}\n{comps}\n".encode()
        return head + body
    head = f"PVM{version}\n{w} {h} {d}\n{scale[0]:g} {scale[1]:g} {scale[2]:g}\n{comps}\n".encode()
    tail = b"description\0courtesy\0parameter\0comment\0" if version == 3 else b""
    return head + body + tail


def _smooth(shape, dtype, seed):
    rng = np.random.default_rng(seed)
    z, y, x = np.mgrid[0:shape[0], 0:shape[1], 0:shape[2]]
    v = 0.5 + 0.3 * np.sin(x / 3.0) * np.cos(y / 4.0) + 0.15 * np.sin(z / 2.0) + 0.02 * rng.standard_normal(shape)
    mx = 255 if dtype == np.uint8 else 65535
    return np.clip(v * mx, 0, mx).astype(dtype)


@pytest.mark.parametrize("dtype", [np.uint8, np.uint16])
@pytest.mark.parametrize("version", [1, 2, 3])
def test_plain_pvm(tmp_path, dtype, version):
    vox = _smooth((5, 7, 9), dtype, version)
    p = str(tmp_path / "v.pvm")
    open(p, "wb").write(_pvm_bytes(vox, version, scale=(1.0, 0.5, 2.0), comment=True))
    assert np.array_equal(hostapi.read_volume(p), vox)
    assert hostapi.read_volume_scale(p) == ((1.0, 0.5, 2.0) if version > 1 else (1.0, 1.0, 1.0))
    if R.available():
        ref, sc = R.pvm_read(p)
        assert np.array_equal(ref, vox) and sc == hostapi.read_volume_scale(p)


@pytest.mark.parametrize("dtype,shape", [(np.uint8, (6, 11, 13)), (np.uint16, (4, 10, 12)), (np.uint8, (1, 1, 3)), (np.uint8, (3, 40, 50))])
@pytest.mark.parametrize("skip,strip_kind,version", [(1, "one", 1), (1, "row", 2), (2, "row", 2), (2, "one", 1), (3, "slice", 2), (4, "row", 1)])
def test_dds_compressed_pvm(tmp_path, dtype, shape, skip, strip_kind, version):
    """Differential data streams: every skip (1..4), strip = 1 / a row / a slice, both container versions, random run lengths
    and widths; mirror == original voxels == the reference's decoder."""
    vox = _smooth(shape, dtype, skip * 7 + version)
    strip = {"one": 1, "row": shape[2], "slice": shape[2] * shape[1]}[strip_kind]
    rng = np.random.default_rng(shape[2] + skip)
    p = str(tmp_path / "c.pvm")
    open(p, "wb").write(dds_encode(_pvm_bytes(vox, 3, scale=(0.5, 0.5, 1.5)), skip, strip, version, rng))
    got = hostapi.read_volume(p)
    assert got.dtype == vox.dtype and np.array_equal(got, vox)
    assert hostapi.read_volume_scale(p) == (0.5, 0.5, 1.5)
    if R.available():
        ref, sc = R.pvm_read(p)
        assert np.array_equal(ref, got) and sc == (0.5, 0.5, 1.5)


@pytest.mark.parametrize("dtype,shape", [(np.uint8, (5, 7, 9)), (np.uint16, (4, 10, 12)), (np.uint8, (3, 40, 50))])
@pytest.mark.parametrize("compressed", [False, True])
def test_pvmold(tmp_path, dtype, shape, compressed):
    """`.pvmold` (reader.cpp:227-281 over PvmOld): the .pvm container, 16-bit voxels high byte first, stretched from the volume's
    [min, max] to the full range in float32, truncated, normalised.  Mirror == the arithmetic restated here == the reference's PvmOld."""
    vox = _smooth(shape, dtype, 11)
    body = _pvm_bytes(vox, 3, scale=(0.5, 1.0, 1.5))
    p = str(tmp_path / "v.pvmold")
    open(p, "wb").write(dds_encode(body, 2, shape[2], 2, np.random.default_rng(5)) if compressed else body)
    got = hostapi.read_volume(p)
    assert got.dtype == np.float32 and got.shape == shape
    assert hostapi.read_volume_scale(p) == (0.5, 1.0, 1.5)
    comps = vox.dtype.itemsize
    v = (vox.astype("<u2").byteswap() if comps == 2 else vox).astype(np.float32)     # the file holds the low byte first; PvmOld reads it as the high one
    top = np.float32(255.0 if comps == 1 else 65535.0)
    stretched = ((v - v.min()) / (v.max() - v.min())).astype(np.float32) * top
    want = (np.trunc(stretched.astype(np.float64)).astype(np.float32).astype(np.float64) / float(top)).astype(np.float32)
    assert np.array_equal(got, want)
    if R.available():
        ref, sc = R.pvmold_read(p)
        assert np.array_equal(ref, got) and sc == (0.5, 1.0, 1.5)


def test_pvmold_constant_volume_is_refused(tmp_path):
    p = str(tmp_path / "flat.pvmold")
    open(p, "wb").write(_pvm_bytes(np.full((2, 3, 4), 7, np.uint8), 2))
    with pytest.raises(RuntimeError):
        hostapi.read_volume(p)


def test_dds_compresses_smooth_data(tmp_path):
    """Sanity of the encoder used above: the stream of a smooth volume is shorter than the plain file."""
    vox = _smooth((8, 32, 32), np.uint8, 0)
    plain = _pvm_bytes(vox, 2)
    assert len(dds_encode(plain, 1, 32, 2, np.random.default_rng(0))) < len(plain)


def test_bad_pvm_files(tmp_path):
    p = str(tmp_path / "bad.pvm")
    for content in [b"", b"PVX\n1 1 1\n1\n\0", b"PVM\n2 2 2\n1\n\0\0\0", b"PVM2\n2 2 2\n1 1 0\n1\n" + bytes(8), b"PVM\n2 2 2\n3\n" + bytes(24)]:
        open(p, "wb").write(content)
        with pytest.raises(RuntimeError):
            hostapi.read_volume(p)
    with pytest.raises(RuntimeError):
        hostapi.read_volume(str(tmp_path / "missing.pvm"))


@pytest.mark.parametrize("dtype,tname", [(np.uint8, "uint8"), (np.uint16, "uint16")])
@pytest.mark.parametrize("ext", ["nrrd", "nhrd"])
def test_nrrd_detached_header(tmp_path, dtype, tname, ext):
    vox = _smooth((4, 6, 8), dtype, 3)
    vox.tofile(str(tmp_path / "vol.raw.bin"))
    hdr = ("NRRD0004\n# a comment\ntype: %s\ndimension: 3\nspace: left-posterior-superior\nsizes: 8 6 4\n"
           "space directions: (2,0,0) (0,2,0) (0,0,2)\nkinds: domain domain domain\nendian: little\nencoding: raw\n"
           "space origin: (0,0,0)\ndata file: vol.raw.bin\n") % tname
    p = str(tmp_path / ("vol." + ext))
    open(p, "w").write(hdr)
    assert np.array_equal(hostapi.read_volume(p), vox)
    assert hostapi.read_volume_scale(p) == (1.0, 1.0, 1.0)          # reader.cpp:447-451: directions parsed, never applied
    open(p, "w").write(hdr.replace("vol.raw.bin", "absent.bin"))
    with pytest.raises(RuntimeError):
        hostapi.read_volume(p)


@pytest.mark.parametrize("dtype,fmt", [(np.uint8, "UCHAR"), (np.uint16, "USHORT")])
def test_dat_header(tmp_path, dtype, fmt):
    vox = _smooth((5, 4, 7), dtype, 4)
    vox.tofile(str(tmp_path / "engine.raw"))
    p = str(tmp_path / "engine.dat")
    open(p, "w").write("ObjectFileName: engine.raw\r\nTaggedFileName: ---\r\nResolution:     7 4 5\r\nSliceThickness: 1.0 1.0 2.5\r\n"
                       "Format:         %s\r\nNbrTags:        0\r\nObjectType:     TEXTURE_VOLUME_OBJECT\r\nObjectModel:    RGBA\r\nGridType:       EQUIDISTANT\r\n" % fmt)
    assert np.array_equal(hostapi.read_volume(p), vox)
    assert hostapi.read_volume_scale(p) == (1.0, 1.0, 2.5)
    os.remove(str(tmp_path / "engine.raw"))
    with pytest.raises(RuntimeError):
        hostapi.read_volume(p)
