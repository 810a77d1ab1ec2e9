"""TEST INFRASTRUCTURE ONLY -- oracle-side readers for the MANIAC inputs.

read_pcd_xyz   restates o3d.io.read_point_cloud(file, remove_nan_points=False)
               (manipulation_analyser.py:1441) for PCD v0.7 files with float32
               x/y/z fields: DATA ascii | binary | binary_compressed.  In
               binary_compressed the LZF payload holds the fields in SoA order
               (all x, all y, all z, all rgba) -- SURVEY.md section 7.2(7).
read_labels    restates pd.read_csv(label_file, delim_whitespace=True, dtype=float64,
               header=None).to_numpy() (manipulation_analyser.py:1391).
"""
import ctypes
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None


def _lzf():
    global _LIB
    if _LIB is None:
        so = os.path.join(_HERE, "_ref", "liboracle_lzf.so")
        src = os.path.join(_HERE, "lzf_dec.c")
        if not os.path.exists(so) or os.path.getmtime(so) < os.path.getmtime(src):
            os.makedirs(os.path.dirname(so), exist_ok=True)
            subprocess.check_call(["gcc", "-O2", "-shared", "-fPIC", src, "-o", so])
        _LIB = ctypes.CDLL(so)
        _LIB.oracle_lzf_decompress.restype = ctypes.c_long
        _LIB.oracle_lzf_decompress.argtypes = [ctypes.c_void_p, ctypes.c_size_t,
                                               ctypes.c_void_p, ctypes.c_size_t]
        _LIB.oracle_lzf_compress.restype = ctypes.c_long
        _LIB.oracle_lzf_compress.argtypes = [ctypes.c_void_p, ctypes.c_size_t, ctypes.c_void_p, ctypes.c_size_t]
    return _LIB


def lzf_compress(data):
    """bytes -> LZF stream (oracle/lzf_dec.c: oracle_lzf_compress)."""
    src = np.frombuffer(bytes(data), np.uint8)
    out = np.empty(len(src) + len(src) // 16 + 64, np.uint8)
    n = _lzf().oracle_lzf_compress(src.ctypes.data, len(src), out.ctypes.data, len(out))
    assert n > 0
    return out[:n].tobytes()


def pcd_binary_compressed(xyz, width, height, rgba=True):
    """The bytes of a PCD v0.7 binary_compressed file as PCL writes the MANIAC clouds: fields x y z (rgba), SoA payload
    (all x, all y, all z, all rgba) compressed as ONE LZF stream behind two uint32 (compressed, uncompressed size)."""
    xyz = np.ascontiguousarray(xyz, np.float32)
    n = len(xyz)
    planes = [xyz[:, 0].tobytes(), xyz[:, 1].tobytes(), xyz[:, 2].tobytes()]
    if rgba:
        planes.append(np.zeros(n, np.uint32).tobytes())
    raw = b"".join(planes)
    comp = lzf_compress(raw)
    extra = (" rgba", " 4", " U", " 1") if rgba else ("", "", "", "")
    head = ("# .PCD v0.7 - Point Cloud Data file format\nVERSION 0.7\nFIELDS x y z%s\nSIZE 4 4 4%s\nTYPE F F F%s\nCOUNT 1 1 1%s\n"
            "WIDTH %d\nHEIGHT %d\nVIEWPOINT 0 0 0 1 0 0 0\nPOINTS %d\nDATA binary_compressed\n"
            % (extra + (width, height, n))).encode()
    return head + np.array([len(comp), len(raw)], np.uint32).tobytes() + comp


def lzf_decompress_py(data, out_len):
    """Pure-Python LZF (slow; only for cross-checking the C decoder on small inputs)."""
    out = bytearray()
    i = 0
    n = len(data)
    while i < n:
        ctrl = data[i]
        i += 1
        if ctrl < 32:
            out += data[i:i + ctrl + 1]
            i += ctrl + 1
        else:
            length = ctrl >> 5
            if length == 7:
                length += data[i]
                i += 1
            ref = len(out) - ((ctrl & 0x1F) << 8) - 1 - data[i]
            i += 1
            for _ in range(length + 2):
                out.append(out[ref])
                ref += 1
    assert len(out) == out_len, (len(out), out_len)
    return bytes(out)


def read_pcd_xyz(path):
    """-> (N,3) float32 array with NaNs preserved, plus (width, height)."""
    with open(path, "rb") as f:
        raw = f.read()
    hdr = {}
    pos = 0
    while True:
        end = raw.index(b"\n", pos)
        line = raw[pos:end].decode("ascii", "replace").strip()
        pos = end + 1
        if not line or line.startswith("#"):
            continue
        key, _, val = line.partition(" ")
        hdr[key] = val.split()
        if key == "DATA":
            break
    fields = hdr["FIELDS"]
    sizes = [int(s) for s in hdr["SIZE"]]
    types = hdr["TYPE"]
    counts = [int(c) for c in hdr.get("COUNT", ["1"] * len(fields))]
    width, height = int(hdr["WIDTH"][0]), int(hdr["HEIGHT"][0])
    npts = int(hdr["POINTS"][0])
    mode = hdr["DATA"][0]
    for name in "xyz":
        k = fields.index(name)
        assert sizes[k] == 4 and types[k] == "F" and counts[k] == 1, "x/y/z must be float32"
    fsz = [s * c for s, c in zip(sizes, counts)]
    stride = sum(fsz)
    offs = np.concatenate(([0], np.cumsum(fsz)))[:-1]
    xyz = np.empty((npts, 3), np.float32)
    if mode == "binary_compressed":
        comp, uncomp = np.frombuffer(raw, np.uint32, 2, pos)
        buf = np.empty(int(uncomp), np.uint8)
        src = np.frombuffer(raw, np.uint8, int(comp), pos + 8)
        got = _lzf().oracle_lzf_decompress(src.ctypes.data, int(comp), buf.ctypes.data, int(uncomp))
        assert got == uncomp, (got, uncomp)
        for a, name in enumerate("xyz"):
            k = fields.index(name)
            xyz[:, a] = np.frombuffer(buf, np.float32, npts, int(offs[k]) * npts)
    elif mode == "binary":
        rec = np.frombuffer(raw, np.uint8, npts * stride, pos).reshape(npts, stride)
        for a, name in enumerate("xyz"):
            k = fields.index(name)
            xyz[:, a] = rec[:, offs[k]:offs[k] + 4].copy().view(np.float32)[:, 0]
    elif mode == "ascii":
        tab = np.loadtxt(raw[pos:].decode("ascii").splitlines(), dtype=np.float64, ndmin=2)
        col = np.concatenate(([0], np.cumsum(counts)))[:-1]
        for a, name in enumerate("xyz"):
            xyz[:, a] = tab[:, col[fields.index(name)]].astype(np.float32)
    else:
        raise ValueError("unsupported PCD DATA mode %r" % mode)
    return xyz, (width, height)


def read_labels(path):
    """-> (rows, cols) float64, like pd.read_csv(..., delim_whitespace=True).to_numpy()."""
    with open(path, "rb") as f:
        rows = [np.array(line.split(), dtype=np.float64) for line in f if line.strip()]
    return np.vstack(rows)
