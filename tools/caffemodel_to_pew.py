#!/usr/bin/env python
"""Converts a trained Caffe model (binary NetParameter, *.caffemodel) into the PEW1 weight file pe_net_load_weights reads
(SURVEY 8f rank 3), without protobuf / Caffe: a minimal reader of the protobuf wire format for the few fields needed.

    python tools/caffemodel_to_pew.py net.caffemodel weights.pew [--prototxt network/manifold_rgbdnor_16.prototxt]

caffe.proto field numbers used:  NetParameter.layer = 100 (LayerParameter: name 1, type 2, blobs 7) and the legacy
NetParameter.layers = 2 (V1LayerParameter: name 4, blobs 6);  BlobProto: data 5 (float, packed or not), double_data 8,
shape 7 (BlobShape.dim = 1), legacy num / channels / height / width = 1..4.
The blobs of all layers, in file order, are the net's learnable blobs in Caffe order (weight, bias; PReLU slope) - the
order pe_net_blob_info reports.  With --prototxt the blob count and sizes are checked against the net definition.
PEW1: "PEW1" | int32 n_blobs | per blob: int64 count, count x f32 (little endian)."""
import struct
import sys

import numpy as np


def _varint(buf, pos):
    out, shift = 0, 0
    while True:
        b = buf[pos]
        pos += 1
        out |= (b & 0x7F) << shift
        if not b & 0x80:
            return out, pos
        shift += 7


def fields(buf):
    """yields (field number, wire type, value) of one message; value is an int, or a memoryview for length-delimited"""
    pos, n = 0, len(buf)
    while pos < n:
        key, pos = _varint(buf, pos)
        num, wt = key >> 3, key & 7
        if wt == 0:
            val, pos = _varint(buf, pos)
        elif wt == 1:
            val, pos = buf[pos:pos + 8], pos + 8
        elif wt == 2:
            ln, pos = _varint(buf, pos)
            val, pos = buf[pos:pos + ln], pos + ln
        elif wt == 5:
            val, pos = buf[pos:pos + 4], pos + 4
        else:
            raise ValueError(f"unsupported wire type {wt}")
        yield num, wt, val


def parse_blob(buf):
    data, shape, legacy = [], None, {}
    for num, wt, val in fields(buf):
        if num == 5:        # float data
            data.append(np.frombuffer(bytes(val), "<f4"))
        elif num == 8:      # double_data
            data.append(np.frombuffer(bytes(val), "<f8").astype(np.float32))
        elif num == 7:      # BlobShape
            for n2, w2, v2 in fields(val):
                if n2 == 1:
                    if w2 == 2:
                        dims, p = [], 0
                        while p < len(v2):
                            d, p = _varint(v2, p)
                            dims.append(d)
                        shape = (shape or []) + dims
                    else:
                        shape = (shape or []) + [v2]
        elif num in (1, 2, 3, 4) and wt == 0:
            legacy[num] = val
    arr = np.concatenate(data) if data else np.zeros(0, np.float32)
    if shape is None and legacy:
        shape = [legacy.get(i, 1) for i in (1, 2, 3, 4)]
    return arr, shape


def parse_caffemodel(raw):
    """-> list of (layer name, [(array, shape), ...]) for the layers that carry blobs, in file order"""
    buf = memoryview(raw)
    layers = []
    for num, wt, val in fields(buf):
        if num not in (100, 2) or wt != 2:
            continue
        name_field, blob_field = (1, 7) if num == 100 else (4, 6)
        name, blobs = "", []
        for n2, w2, v2 in fields(val):
            if n2 == name_field and w2 == 2:
                name = bytes(v2).decode("utf-8", "replace")
            elif n2 == blob_field and w2 == 2:
                blobs.append(parse_blob(v2))
        if blobs:
            layers.append((name, blobs))
    return layers


def write_pew(path, arrays):
    with open(path, "wb") as f:
        f.write(b"PEW1" + struct.pack("<i", len(arrays)))
        for a in arrays:
            a = np.ascontiguousarray(a, "<f4")
            f.write(struct.pack("<q", a.size))
            f.write(a.tobytes())


def main(argv):
    if len(argv) < 3:
        print(__doc__)
        return 2
    layers = parse_caffemodel(open(argv[1], "rb").read())
    arrays = [a for _, blobs in layers for a, _ in blobs]
    for name, blobs in layers:
        print(f"{name}: " + ", ".join(f"{tuple(s) if s else (a.size,)}" for a, s in blobs))
    if "--prototxt" in argv:
        import importlib
        import os
        sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
        # the library's own prototxt reader through the C ABI (pe_net_create_from_prototxt + pe_net_blob_info; no GPU needed)
        capi = importlib.import_module("3dposeestimation_b200.capi")
        net = capi.Net(path=argv[argv.index("--prototxt") + 1])
        want = [cnt for _, cnt in net.blob_info()]
        got = [a.size for a in arrays]
        if want != got:
            print(f"blob sizes {got} do not match the prototxt's {want}")
            return 1
    write_pew(argv[2], arrays)
    print(f"wrote {argv[2]}: {len(arrays)} blobs, {sum(a.size for a in arrays)} floats")
    return 0


if __name__ == "__main__":
    sys.exit(main(sys.argv))
