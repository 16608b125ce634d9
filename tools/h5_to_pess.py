#!/usr/bin/env python
"""Dumps a reference sample file (HDF5 datasets "data" [N,C,H,W] f32 and "label" [N,L] f32, written by
hdf5Handler::write, reference src/hdf5handler.cpp:71-132) into the flat PESS0001 file pe_samples_read reads:
the same two arrays behind a 32-byte header.  Needs h5py, which this build box does not have: best effort, untested here.

    python tools/h5_to_pess.py samples.h5 samples.pess"""
import struct
import sys

import numpy as np


def write_pess(path, data_nchw, labels):
    data = np.ascontiguousarray(data_nchw, "<f4")
    lab = np.ascontiguousarray(labels, "<f4").reshape(data.shape[0], -1)
    n, c, h, w = data.shape
    with open(path, "wb") as f:
        f.write(b"PESS0001" + struct.pack("<qiiii", n, c, h, w, lab.shape[1]))
        f.write(data.tobytes())
        f.write(lab.tobytes())


def main(argv):
    if len(argv) != 3:
        print(__doc__)
        return 2
    import h5py
    with h5py.File(argv[1], "r") as f:
        write_pess(argv[2], f["data"][...], f["label"][...])
    return 0


if __name__ == "__main__":
    sys.exit(main(sys.argv))
