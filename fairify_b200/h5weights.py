"""Minimal pure-Python reader for the Keras HDF5 weight files Fairify ships.

The reference loads its models with ``tensorflow.keras.models.load_model`` and then reads
``layer.get_weights()`` (reference ``src/AC/Verify-AC.py:95-99``).  Neither tensorflow nor h5py exists in
this image, and the hot path only needs the Dense kernels/biases, so this module walks the small subset
of HDF5 those files use: superblock v0, v1 object headers, old-style symbol-table groups, contiguous or
compact little-endian float32 datasets.  It is a host-side tool (fixture generation, the ``load_model``
shim of the drop-in ``utils`` package); nothing on the GPU path depends on it.
"""
from __future__ import annotations

import json
import struct

import numpy as np

_SIG = b"\x89HDF\r\n\x1a\n"


class H5Error(RuntimeError):
    pass


class _File:
    def __init__(self, buf: bytes):
        if buf[:8] != _SIG:
            raise H5Error("not an HDF5 file")
        if buf[8] != 0:
            raise H5Error(f"unsupported superblock version {buf[8]}")
        self.b = buf
        self.O = buf[13]
        self.L = buf[14]
        if self.O != 8 or self.L != 8:
            raise H5Error("only 8-byte offsets/lengths supported")
        root_entry = 24 + 4 * self.O
        self.root_header = self._u(root_entry + self.O, self.O)

    def _u(self, off, n):
        return int.from_bytes(self.b[off:off + n], "little")

    # -- object headers ---------------------------------------------------
    def messages(self, addr):
        """Yield (type, payload_offset, payload_size) for a v1 object header."""
        b = self.b
        if b[addr] != 1:
            raise H5Error(f"unsupported object header version {b[addr]} at {addr}")
        nmsg = self._u(addr + 2, 2)
        size = self._u(addr + 8, 4)
        blocks = [(addr + 16, size)]
        seen = 0
        while blocks and seen < nmsg:
            pos, left = blocks.pop(0)
            end = pos + left
            while pos + 8 <= end and seen < nmsg:
                mtype = self._u(pos, 2)
                msize = self._u(pos + 2, 2)
                payload = pos + 8
                seen += 1
                if mtype == 0x0010:  # continuation
                    blocks.append((self._u(payload, self.O), self._u(payload + self.O, self.L)))
                else:
                    yield mtype, payload, msize
                pos = payload + msize

    # -- groups -------------------------------------------------------------
    def children(self, addr):
        """Return {name: object_header_address} for an old-style group."""
        for mtype, p, _ in self.messages(addr):
            if mtype == 0x0011:
                btree = self._u(p, self.O)
                heap = self._u(p + self.O, self.O)
                heap_data = self._heap_data(heap)
                out = {}
                self._walk_btree(btree, heap_data, out)
                return out
        return {}

    def _heap_data(self, addr):
        if self.b[addr:addr + 4] != b"HEAP":
            raise H5Error("bad local heap signature")
        return self._u(addr + 8 + 2 * self.L, self.O)

    def _name(self, heap_data, off):
        start = heap_data + off
        end = self.b.index(b"\0", start)
        return self.b[start:end].decode("utf-8")

    def _walk_btree(self, addr, heap_data, out):
        b = self.b
        if b[addr:addr + 4] != b"TREE":
            raise H5Error("bad B-tree signature")
        level = b[addr + 5]
        used = self._u(addr + 6, 2)
        base = addr + 8 + 2 * self.O
        for k in range(used):
            child = self._u(base + self.L + k * (self.L + self.O), self.O)
            if level > 0:
                self._walk_btree(child, heap_data, out)
            else:
                self._read_snod(child, heap_data, out)

    def _read_snod(self, addr, heap_data, out):
        if self.b[addr:addr + 4] != b"SNOD":
            raise H5Error("bad symbol node signature")
        n = self._u(addr + 6, 2)
        ent = addr + 8
        esize = 2 * self.O + 24
        for k in range(n):
            e = ent + k * esize
            out[self._name(heap_data, self._u(e, self.O))] = self._u(e + self.O, self.O)

    # -- datasets -----------------------------------------------------------
    def dataset(self, addr):
        """Return a numpy array if `addr` is a float32 dataset, else None."""
        dims = None
        dtype_ok = False
        data = None
        for mtype, p, msize in self.messages(addr):
            b = self.b
            if mtype == 0x0001:
                ver, rank, flags = b[p], b[p + 1], b[p + 2]
                q = p + (8 if ver == 1 else 4)
                dims = [self._u(q + i * self.L, self.L) for i in range(rank)]
            elif mtype == 0x0003:
                cls = b[p] & 0x0F
                size = self._u(p + 4, 4)
                dtype_ok = (cls == 1 and size == 4 and (b[p + 1] & 1) == 0)
            elif mtype == 0x0008:
                ver = b[p]
                if ver != 3:
                    raise H5Error(f"unsupported layout version {ver}")
                lclass = b[p + 1]
                if lclass == 1:
                    a = self._u(p + 2, self.O)
                    n = self._u(p + 2 + self.O, self.L)
                    data = (a, n)
                elif lclass == 0:
                    n = self._u(p + 2, 2)
                    data = (p + 4, n)
                else:
                    raise H5Error("chunked datasets not supported")
        if dims is None or data is None:
            return None
        if not dtype_ok:
            return None
        count = int(np.prod(dims)) if dims else 1
        a, n = data
        if n < 4 * count:
            raise H5Error("dataset shorter than its dataspace")
        arr = np.frombuffer(self.b, dtype="<f4", count=count, offset=a).reshape(dims)
        return arr.copy()


def _layer_names(buf: bytes):
    """Dense layer names + units in model order, from the embedded model_config JSON."""
    i = buf.find(b'{"class_name"')
    if i < 0:
        raise H5Error("model_config JSON not found")
    text = buf[i:].decode("latin1")
    cfg, _ = json.JSONDecoder().raw_decode(text)
    out = []
    for layer in cfg["config"]["layers"]:
        if layer["class_name"] == "InputLayer":
            continue
        if layer["class_name"] != "Dense":
            raise H5Error(f"unsupported layer type {layer['class_name']}")
        c = layer["config"]
        out.append((c["name"], int(c["units"]), c.get("activation")))
    return out


def read_keras_dense_weights(path):
    """Return (w, b): lists of float32 arrays, w[l] of shape (in, out), b[l] of shape (out,).

    Same values `model.layers[i].get_weights()` yields in the reference drivers
    (`src/AC/Verify-AC.py:97-99`).
    """
    with open(path, "rb") as fh:
        buf = fh.read()
    f = _File(buf)
    root = f.children(f.root_header)
    if "model_weights" not in root:
        raise H5Error("no /model_weights group")
    mw = f.children(root["model_weights"])
    w, b = [], []
    for name, units, _act in _layer_names(buf):
        if name not in mw:
            raise H5Error(f"layer {name} missing from /model_weights")
        inner = f.children(mw[name])
        # /model_weights/<layer>/<layer>/{kernel:0,bias:0}
        # (the inner group is usually named like the layer, but e.g. BM-4 has dense_4/dense_4_1)
        grp = inner
        while "kernel:0" not in grp:
            if len(grp) != 1:
                raise H5Error(f"layer {name}: cannot locate kernel:0")
            grp = f.children(next(iter(grp.values())))
        k = f.dataset(grp["kernel:0"])
        bias = f.dataset(grp["bias:0"])
        if k is None or bias is None:
            raise H5Error(f"layer {name}: kernel/bias not float32 contiguous")
        if k.shape[1] != units or bias.shape != (units,):
            raise H5Error(f"layer {name}: shape mismatch {k.shape} {bias.shape}")
        w.append(k.astype(np.float32))
        b.append(bias.astype(np.float32))
    for l in range(1, len(w)):
        if w[l].shape[0] != w[l - 1].shape[1]:
            raise H5Error("layer chain shape mismatch")
    return w, b
