"""What the reference's Python decoder does with two ill-formed pglz streams that this library refuses
(build container only: needs /root/reference):

    python tests/golden/make_golden_deviation.py

  * a match with offset 0;
  * a match whose offset reaches before the start of the output.
pglz_decompress_py (/root/reference/cerebro_gpdb/pg_page_reader.py:191-231) indexes a zero-filled bytearray with
dest[dp - off]: offset 0 re-reads the byte being written (zero), a negative index wraps round to the END of the
buffer (still zero when it is read) — so it returns bytes and raises nothing as long as dp and sp end exactly.
The C decoder (pg_lzcompress.c:49-143) reads memory outside `dest` for the second stream.  No compressor emits such
streams; libcds_da reports them as corrupt.  Writes tests/golden/pglz_deviation.npz.
"""
import os
import struct
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
from make_golden import import_reference          # noqa: E402


def varlena(stream, rawsize):
    return struct.pack(">I", 0x40000000 | (8 + len(stream))) + struct.pack("<I", rawsize) + bytes(stream)


def main():
    ref_reader, _ = import_reference()
    cases = {
        # literal 'A', then a match of length 3 with offset 0, then literal 'B'   (control byte 0b010)
        "offset_zero": varlena(bytes([0b00000010, ord("A"), 0x00, 0x00, ord("B")]), 5),
        # literal 'A', then a match of length 4 with offset 100 (only 1 byte of output exists)
        "offset_before_start": varlena(bytes([0b00000010, ord("A"), 0x01, 100]), 5),
    }
    # the same match inside a 200-byte value: index 1 - 100 is a valid NEGATIVE index there and wraps to dest[101],
    # which is still zero when it is read: decodes without raising
    body = bytearray([0b00000010, ord("A"), 0x01, 100])
    rest = bytes((7 * i + 3) & 0xFF for i in range(195))
    for k in range(6):                                     # the first group still has room for 6 literals
        body.append(rest[k])
    pos = 6
    while pos < len(rest):
        body.append(0)
        body += rest[pos:pos + 8]
        pos += 8
    cases["offset_wraps_to_unwritten"] = varlena(bytes(body), 200)
    arrays = {"names": np.array(list(cases))}
    for i, (name, datum) in enumerate(cases.items()):
        dest = bytearray(struct.unpack_from("<I", datum, 4)[0])
        try:
            out = ref_reader.pglz_decompress(bytearray(datum), dest, "py")
            arrays["raised_%d" % i] = np.array([""])
            arrays["out_%d" % i] = np.frombuffer(bytes(out), dtype=np.uint8)
        except Exception as e:  # noqa
            arrays["raised_%d" % i] = np.array([type(e).__name__ + ": " + str(e)])
            arrays["out_%d" % i] = np.zeros(0, dtype=np.uint8)
        arrays["datum_%d" % i] = np.frombuffer(datum, dtype=np.uint8)
        print(name, "->", arrays["raised_%d" % i][0] or bytes(arrays["out_%d" % i]))
    np.savez_compressed(os.path.join(HERE, "pglz_deviation.npz"), **arrays)


if __name__ == "__main__":
    main()
