"""Minimal HDF5 *writer* for the test-suite: produces files laid out the way h5py/Keras write
`model.save(..., save_format='h5')` output with the default ("earliest") library format, from the
published HDF5 File Format Specification v1.1/2.0:

  superblock v0 -> root symbol-table entry; groups = v1 object header + symbol-table message ->
  v1 B-tree (type 0) + local heap + SNOD leaves (<= 8 entries each); datasets = v1 object header with
  dataspace v1, IEEE float datatype (class 1), layout v3 contiguous or chunked (v1 B-tree type 1, optional
  shuffle + deflate pipeline v1); attributes v1 with fixed-length (class 3) or variable-length (class 9,
  global heap "GCOL") strings; optionally placed in an object-header continuation block.

There is no HDF5 library in this image (no h5py, no libhdf5), so host/Hdf5File.cpp cannot be checked
against the real thing here: this writer and that reader are two independent readings of the same
specification. DESIGN.md marks the HDF5 path "parity unpinned".
"""
from __future__ import annotations

import struct
import zlib

import numpy as np

UNDEF = 0xFFFFFFFFFFFFFFFF


def _pad8(b: bytes) -> bytes:
    return b + b"\0" * (-len(b) % 8)


class H5Writer:
    def __init__(self):
        self.buf = bytearray(96)            # superblock v0 with 8-byte offsets/lengths
        self.root = {"groups": {}, "datasets": {}, "attrs": []}

    # ---- tree description -------------------------------------------------------------------
    def group(self, path: str) -> dict:
        node = self.root
        for part in [p for p in path.split("/") if p]:
            node = node["groups"].setdefault(part, {"groups": {}, "datasets": {}, "attrs": []})
        return node

    def dataset(self, path: str, array: np.ndarray, chunks=None, gzip=None, shuffle=False):
        parent, name = path.rsplit("/", 1)
        self.group(parent)["datasets"][name] = dict(array=np.ascontiguousarray(array), chunks=chunks, gzip=gzip, shuffle=shuffle)

    def attr(self, path: str, name: str, value: str, vlen: bool = False, continuation: bool = False):
        self.group(path)["attrs"].append(dict(name=name, value=value.encode("utf8"), vlen=vlen, continuation=continuation))

    # ---- low level --------------------------------------------------------------------------
    def _alloc(self, data: bytes, align: int = 8) -> int:
        while len(self.buf) % align:
            self.buf.append(0)
        addr = len(self.buf)
        self.buf += data
        return addr

    @staticmethod
    def _msg(mtype: int, body: bytes, flags: int = 0) -> bytes:
        body = _pad8(body)
        return struct.pack("<HHB3x", mtype, len(body), flags) + body

    def _object_header(self, messages: list[bytes], continued: list[bytes] | None = None) -> int:
        n = len(messages)
        cont_patch = None
        if continued:
            # continuation message: offset + length, patched once the block is placed
            messages = messages + [self._msg(0x10, struct.pack("<QQ", 0, 0))]
            n += 1 + len(continued)
        body = b"".join(messages)
        addr = self._alloc(struct.pack("<BxHII4x", 1, n, 1, len(body)) + body)
        if continued:
            block = b"".join(continued)
            caddr = self._alloc(block)
            cont_patch = addr + 16 + len(body) - 16
            struct.pack_into("<QQ", self.buf, cont_patch, caddr, len(block))
        return addr

    def _global_heap(self, payload: bytes) -> tuple[int, int]:
        obj = struct.pack("<HH4xQ", 1, 1, len(payload)) + _pad8(payload)
        size = max(4096, 16 + len(obj) + 16)
        size += -size % 8
        free = size - 16 - len(obj)
        block = b"GCOL" + struct.pack("<B3xQ", 1, size) + obj + struct.pack("<HH4xQ", 0, 0, free) + b"\0" * (free - 16)
        return self._alloc(block), 1

    def _attr_msg(self, a: dict) -> bytes:
        name = a["name"].encode() + b"\0"
        space = struct.pack("<BBB5x", 1, 0, 0)
        if a["vlen"]:
            dtype = struct.pack("<BBBBI", 0x19, 0x01, 0x01, 0, 16) + struct.pack("<BBBBI", 0x13, 0x00, 0, 0, 1)
            addr, idx = self._global_heap(a["value"])
            data = struct.pack("<IQI", len(a["value"]), addr, idx)
        else:
            dtype = struct.pack("<BBBBI", 0x13, 0x01, 0, 0, max(1, len(a["value"])))      # null-padded, ASCII
            data = a["value"] or b"\0"
        body = struct.pack("<BxHHH", 1, len(name), len(dtype), len(space)) + _pad8(name) + _pad8(dtype) + _pad8(space) + data
        return self._msg(0x0C, body)

    @staticmethod
    def _float_type(dt: np.dtype) -> bytes:
        if dt == np.float16:
            size, sign, eloc, esz, msz, bias = 2, 15, 10, 5, 10, 15
        elif dt == np.float32:
            size, sign, eloc, esz, msz, bias = 4, 31, 23, 8, 23, 127
        elif dt == np.float64:
            size, sign, eloc, esz, msz, bias = 8, 63, 52, 11, 52, 1023
        else:
            raise TypeError(dt)
        return struct.pack("<BBBBI", 0x11, 0x20, sign, 0, size) + struct.pack("<HHBBBBI", 0, size * 8, eloc, esz, 0, msz, bias)

    def _write_dataset(self, d: dict) -> int:
        arr = d["array"]
        rank = arr.ndim
        msgs = [self._msg(0x01, struct.pack("<BBB5x", 1, rank, 0) + b"".join(struct.pack("<Q", s) for s in arr.shape)),
                self._msg(0x03, self._float_type(arr.dtype), flags=1)]
        es = arr.dtype.itemsize
        if d["chunks"] is None:
            raw = arr.tobytes()
            addr = self._alloc(raw) if raw else UNDEF
            msgs.append(self._msg(0x08, struct.pack("<BBQQ", 3, 1, addr, len(raw))))
        else:
            cd = tuple(d["chunks"])
            filters = []
            if d["shuffle"]:
                filters.append((2, b"shuffle\0", [es]))
            if d["gzip"] is not None:
                filters.append((1, b"deflate\0", [d["gzip"]]))
            if filters:
                body = struct.pack("<BB6x", 1, len(filters))
                for fid, name, cdv in filters:
                    body += struct.pack("<HHHH", fid, len(name), 1, len(cdv)) + _pad8(name) + b"".join(struct.pack("<I", v) for v in cdv)
                    if len(cdv) % 2:
                        body += b"\0" * 4
                msgs.append(self._msg(0x0B, body))
            entries = []
            grid = [range(0, arr.shape[i], cd[i]) for i in range(rank)]
            for origin in np.ndindex(*[len(g) for g in grid]):
                o = [grid[i][origin[i]] for i in range(rank)]
                chunk = np.zeros(cd, dtype=arr.dtype)
                sl = tuple(slice(o[i], min(o[i] + cd[i], arr.shape[i])) for i in range(rank))
                chunk[tuple(slice(0, s.stop - s.start) for s in sl)] = arr[sl]
                raw = chunk.tobytes()
                if d["shuffle"]:
                    raw = np.frombuffer(raw, np.uint8).reshape(-1, es).T.tobytes()
                if d["gzip"] is not None:
                    raw = zlib.compress(raw, d["gzip"])
                entries.append((len(raw), o, self._alloc(raw)))
            node = b"TREE" + struct.pack("<BBHQQ", 1, 0, len(entries), UNDEF, UNDEF)
            for size, o, caddr in entries:
                node += struct.pack("<II", size, 0) + b"".join(struct.pack("<Q", x) for x in o) + struct.pack("<Q", 0) + struct.pack("<Q", caddr)
            node += struct.pack("<II", 0, 0) + b"".join(struct.pack("<Q", s) for s in arr.shape) + struct.pack("<Q", 0)
            baddr = self._alloc(node)
            msgs.append(self._msg(0x08, struct.pack("<BBBQ", 3, 2, rank + 1, baddr) + b"".join(struct.pack("<I", c) for c in cd) + struct.pack("<I", es)))
        return self._object_header(msgs)

    def _write_group(self, g: dict) -> tuple[int, int, int]:
        children = []
        for name, sub in g["groups"].items():
            hdr, bt, hp = self._write_group(sub)
            children.append((name, hdr, 1, bt, hp))
        for name, d in g["datasets"].items():
            children.append((name, self._write_dataset(d), 0, 0, 0))
        children.sort(key=lambda c: c[0].encode())
        heap = bytearray(8)                               # offset 0 = the empty name
        offs = {}
        for name, *_ in children:
            offs[name] = len(heap)
            heap += _pad8(name.encode() + b"\0")
        heap += b"\0" * 16                                # a free block, like the library leaves
        haddr_data = self._alloc(bytes(heap))
        hp = self._alloc(b"HEAP" + struct.pack("<B3xQQQ", 0, len(heap), len(heap) - 16, haddr_data))
        leaves = []
        for i in range(0, len(children), 8):
            part = children[i:i + 8]
            snod = b"SNOD" + struct.pack("<BxH", 1, len(part))
            for name, hdr, cache, bt_, hp_ in part:
                snod += struct.pack("<QQI4x", offs[name], hdr, cache) + (struct.pack("<QQ", bt_, hp_) if cache else b"\0" * 16)
            leaves.append((self._alloc(snod), offs[part[-1][0]]))
        node = b"TREE" + struct.pack("<BBHQQ", 0, 0, len(leaves), UNDEF, UNDEF) + struct.pack("<Q", 0)
        for addr, lastoff in leaves:
            node += struct.pack("<QQ", addr, lastoff)
        bt = self._alloc(node)
        plain = [a for a in g["attrs"] if not a["continuation"]]
        cont = [a for a in g["attrs"] if a["continuation"]]
        msgs = [self._msg(0x11, struct.pack("<QQ", bt, hp))] + [self._attr_msg(a) for a in plain]
        hdr = self._object_header(msgs, [self._attr_msg(a) for a in cont] or None)
        return hdr, bt, hp

    def write(self, path: str):
        hdr, bt, hp = self._write_group(self.root)
        sb = b"\x89HDF\r\n\x1a\n" + struct.pack("<BBBBBBBBHHI", 0, 0, 0, 0, 0, 8, 8, 0, 4, 16, 0)
        sb += struct.pack("<QQQQ", 0, UNDEF, len(self.buf), UNDEF)
        sb += struct.pack("<QQI4xQQ", 0, hdr, 1, bt, hp)
        assert len(sb) == 96
        self.buf[:96] = sb
        with open(path, "wb") as f:
            f.write(self.buf)


def write_keras_nif(path: str, model, vlen_attrs: bool = False, dtype=np.float16, chunked: bool = False):
    """A Keras-style weights file for a NifModel: /model_weights/<layer>/<layer>/{kernel:0,bias:0} and the
    root attributes model_config / keras_version / backend that src/keras/Hdf5Model.cpp:62-94 reads."""
    import json
    w = H5Writer()
    layers = [{"class_name": "InputLayer", "config": {"name": "input_1", "dtype": "float16"}}]
    nfeat = 4 * model.embedding_dimension
    width = nfeat
    for i, l in enumerate(model.layers):
        name = "dense" if i == 0 else f"dense_{i}"
        k = np.asarray(l["kernel"], dtype=dtype)
        if k.shape[0] != width:
            layers.append({"class_name": "Concatenate", "config": {"name": f"concatenate_{i}", "axis": -1}})
        layers.append({"class_name": "Dense", "config": {"name": name, "activation": "relu" if l["relu"] else "linear",
                                                          "dtype": np.dtype(dtype).name, "units": int(k.shape[1]),
                                                          "use_bias": l.get("bias") is not None}})
        kw = dict(chunks=(min(64, k.shape[0]), min(48, k.shape[1])), gzip=4, shuffle=True) if chunked else {}
        w.dataset(f"/model_weights/{name}/{name}/kernel:0", k, **kw)
        if l.get("bias") is not None:
            w.dataset(f"/model_weights/{name}/{name}/bias:0", np.asarray(l["bias"], dtype=dtype))
        width = k.shape[1]
    for extra in ("input_1", "concatenate"):
        w.group(f"/model_weights/{extra}")
    cfg = {"class_name": "Functional", "config": {"name": "model", "layers": layers}}
    w.attr("/", "keras_version", "2.6.0", vlen=vlen_attrs)
    w.attr("/", "backend", "tensorflow", vlen=vlen_attrs)
    w.attr("/", "model_config", json.dumps(cfg), vlen=vlen_attrs, continuation=True)
    w.write(path)
