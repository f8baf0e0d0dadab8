"""Minimal reader (and writer, for tests) of VTK XML StructuredGrid files with inline binary arrays.

The reference loads its stored finite-volume solutions with `pv.read("data/validation/fv/Re*/solution.vts")`
(src/solvers/base.py:1004-1013); PyVista/VTK are not available where this package runs, and all that call needs is
`mesh.points` and `mesh.point_data["u" | "v"]`.  Those files are `<VTKFile type="StructuredGrid" header_type="UInt32"
compressor="vtkZLibDataCompressor">` with `format="binary"` DataArrays: base64(header) + base64(zlib blocks), the
header being [n_blocks, block_size, last_block_size, compressed_size_0, ...].  Standard library only.
"""
from __future__ import annotations

import base64
import xml.etree.ElementTree as ET
import zlib
from dataclasses import dataclass, field
from typing import Dict

import numpy as np

_DTYPES = {"Float64": "f8", "Float32": "f4", "Int64": "i8", "Int32": "i4", "UInt64": "u8", "UInt32": "u4", "UInt8": "u1",
           "Int8": "i1", "Int16": "i2", "UInt16": "u2"}


@dataclass
class StructuredGrid:
    """What the reference reads from / writes through a `pv.StructuredGrid`: points (n, 3), point_data by name, the point
    extent, field_data (numeric arrays and strings), `grid["name"]` access and `save(path)` (main.py:116)."""

    points: np.ndarray
    point_data: Dict[str, np.ndarray] = field(default_factory=dict)
    dimensions: tuple = (0, 0, 0)
    field_data: Dict[str, object] = field(default_factory=dict)

    def __getitem__(self, name):
        return self.point_data[name]

    def __setitem__(self, name, values):
        self.point_data[name] = np.asarray(values)

    @property
    def n_points(self) -> int:
        return int(self.points.shape[0])

    def save(self, path: str) -> None:
        write_vts(path, self.points, self.dimensions, self.point_data, self.field_data)


def _b64_chunk_len(nbytes: int) -> int:
    return (nbytes + 2) // 3 * 4


def _decode_binary(text: str, dtype: np.dtype, header_dtype: np.dtype, compressed: bool, order: str) -> np.ndarray:
    s = "".join(text.split())
    hsz = header_dtype.itemsize
    if not compressed:
        raw = base64.b64decode(s)
        n = int(np.frombuffer(raw[:hsz], dtype=header_dtype)[0])
        return np.frombuffer(raw[hsz:hsz + n], dtype=dtype.newbyteorder(order)).astype(dtype)
    # first the fixed part of the header tells how many blocks follow; header and payload are encoded separately
    first = base64.b64decode(s[:_b64_chunk_len(3 * hsz)])
    nblocks = int(np.frombuffer(first[:hsz], dtype=header_dtype)[0])
    if nblocks == 0:
        return np.zeros(0, dtype=dtype)
    hlen = _b64_chunk_len((3 + nblocks) * hsz)
    head = np.frombuffer(base64.b64decode(s[:hlen])[:(3 + nblocks) * hsz], dtype=header_dtype)
    sizes = [int(x) for x in head[3:3 + nblocks]]
    payload = base64.b64decode(s[hlen:])
    out, pos = [], 0
    for sz in sizes:
        out.append(zlib.decompress(payload[pos:pos + sz]))
        pos += sz
    raw = b"".join(out)
    expect = (nblocks - 1) * int(head[1]) + (int(head[2]) if head[2] else int(head[1]))
    if len(raw) != expect:
        raise ValueError(f"VTK binary array: decompressed {len(raw)} bytes, header promises {expect}")
    return np.frombuffer(raw, dtype=dtype.newbyteorder(order)).astype(dtype)


def read_vts(path: str) -> StructuredGrid:
    """Parse a .vts file: points and every PointData DataArray (multi-component arrays come back 2-D)."""
    root = ET.parse(path).getroot()
    if root.tag != "VTKFile" or root.get("type") != "StructuredGrid":
        raise ValueError(f"{path}: not a VTK XML StructuredGrid file")
    order = "<" if root.get("byte_order", "LittleEndian") == "LittleEndian" else ">"
    header_dtype = np.dtype(_DTYPES[root.get("header_type", "UInt32")]).newbyteorder(order)
    comp = root.get("compressor")
    if comp not in (None, "vtkZLibDataCompressor"):
        raise ValueError(f"{path}: unsupported compressor {comp}")
    grid = root.find("StructuredGrid")
    piece = grid.find("Piece")
    ext = [int(x) for x in piece.get("Extent").split()]
    dims = (ext[1] - ext[0] + 1, ext[3] - ext[2] + 1, ext[5] - ext[4] + 1)

    def array(el):
        dt = np.dtype(_DTYPES[el.get("type")])
        fmt = el.get("format", "ascii")
        if fmt == "binary":
            a = _decode_binary(el.text, dt, header_dtype, comp is not None, order)
        elif fmt == "ascii":
            a = np.array(el.text.split(), dtype=dt)
        else:
            raise ValueError(f"{path}: DataArray format '{fmt}' is not supported (appended data)")
        nc = int(el.get("NumberOfComponents", "1"))
        return a.reshape(-1, nc) if nc > 1 else a

    pts = array(piece.find("Points").find("DataArray"))
    npts = dims[0] * dims[1] * dims[2]
    if pts.shape != (npts, 3):
        raise ValueError(f"{path}: {pts.shape} points for extent {dims}")
    out = StructuredGrid(points=pts, dimensions=dims)
    fd = grid.find("FieldData")
    if fd is not None:
        for el in list(fd):
            if el.get("type") == "String":   # <Array type="String">: the bytes of the string(s), NUL-terminated
                if el.get("format", "ascii") == "binary":
                    raw = _decode_binary(el.text, np.dtype("u1"), header_dtype, comp is not None, order).tobytes()
                else:
                    raw = bytes(int(x) for x in el.text.split())
                out.field_data[el.get("Name")] = raw.rstrip(b"\0").decode()
            else:
                out.field_data[el.get("Name")] = array(el)
    pd = piece.find("PointData")
    if pd is not None:
        for el in pd.findall("DataArray"):
            a = array(el)
            if a.shape[0] != npts:
                raise ValueError(f"{path}: point array {el.get('Name')} has {a.shape[0]} tuples, expected {npts}")
            out.point_data[el.get("Name")] = a
    return out


def _encode_binary(a: np.ndarray, block: int = 32768) -> str:
    raw = np.ascontiguousarray(a).tobytes()
    blocks = [raw[i:i + block] for i in range(0, len(raw), block)] or [b""]
    comp = [zlib.compress(b) for b in blocks]
    last = len(blocks[-1]) if len(blocks[-1]) != block else 0
    head = np.array([len(blocks), block, last] + [len(c) for c in comp], dtype="<u4").tobytes()
    return base64.b64encode(head).decode() + base64.b64encode(b"".join(comp)).decode()


def _range_attrs(a: np.ndarray) -> str:
    """RangeMin / RangeMax as VTK writes them: of the values, or of the tuple norms for multi-component arrays."""
    if a.size == 0:
        return ""
    m = np.linalg.norm(a, axis=1) if a.ndim == 2 else a
    return f' RangeMin="{m.min()!r}" RangeMax="{m.max()!r}"'.replace("np.float64(", "").replace("np.int64(", "").replace(")", "")


def write_vts(path: str, points: np.ndarray, dims, point_data: Dict[str, np.ndarray], field_data: Dict[str, object] = None) -> None:
    """Write the flavour of .vts `pv.StructuredGrid.save` produces (main.py:116; zlib, UInt32 headers, inline binary,
    FieldData in front of the piece, RangeMin / RangeMax attributes), so that ParaView, PyVista and `read_vts` read it."""
    ext = f"0 {dims[0] - 1} 0 {dims[1] - 1} 0 {dims[2] - 1}"
    names = {np.dtype("f8"): "Float64", np.dtype("f4"): "Float32", np.dtype("i8"): "Int64", np.dtype("i4"): "Int32"}
    lines = ['<?xml version="1.0"?>',
             '<VTKFile type="StructuredGrid" version="0.1" byte_order="LittleEndian" header_type="UInt32" '
             'compressor="vtkZLibDataCompressor">',
             f'  <StructuredGrid WholeExtent="{ext}">']
    if field_data:
        lines.append("    <FieldData>")
        for name, val in field_data.items():
            if isinstance(val, str):
                raw = np.frombuffer(val.encode() + b"\0", dtype="u1")
                lines += [f'      <Array type="String" Name="{name}" NumberOfTuples="1" format="binary">',
                          "        " + _encode_binary(raw), "      </Array>"]
            else:
                a = np.atleast_1d(np.asarray(val))
                if a.dtype not in names:
                    a = a.astype(np.int64 if np.issubdtype(a.dtype, np.integer) else np.float64)
                lines += [f'      <DataArray type="{names[a.dtype]}" Name="{name}" NumberOfTuples="{a.shape[0]}" '
                          f'format="binary"{_range_attrs(a)}>', "        " + _encode_binary(a), "      </DataArray>"]
        lines.append("    </FieldData>")
    scal = next((k for k, a in point_data.items() if np.ndim(a) == 1), None)
    lines += [f'  <Piece Extent="{ext}">', f'    <PointData Scalars="{scal}">' if scal else "    <PointData>"]
    for name, a in point_data.items():
        a = np.asarray(a)
        nc = "" if a.ndim == 1 else f' NumberOfComponents="{a.shape[1]}"'
        lines += [f'      <DataArray type="{names[a.dtype]}" Name="{name}"{nc} format="binary"{_range_attrs(a)}>',
                  "        " + _encode_binary(a), "      </DataArray>"]
    pts = np.asarray(points, dtype=np.float64)
    lines += ["    </PointData>", "    <CellData>", "    </CellData>", "    <Points>",
              f'      <DataArray type="Float64" Name="Points" NumberOfComponents="3" format="binary"{_range_attrs(pts)}>',
              "        " + _encode_binary(pts), "      </DataArray>", "    </Points>", "  </Piece>", "  </StructuredGrid>",
              "</VTKFile>"]
    with open(path, "w") as f:
        f.write("\n".join(lines) + "\n")
