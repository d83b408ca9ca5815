"""Minimal proto2 wire reader/writer for the reference's .avrec container (recode.proto:1-19).

Test-side helper only: lets the tests pull `Block.cabac` payloads out of a container written by
the unmodified reference (oracle/_ref/recode) or by the product CLI, and compare them.
"""


def _varint(buf, i):
    v = 0
    s = 0
    while True:
        b = buf[i]
        i += 1
        v |= (b & 0x7F) << s
        s += 7
        if not b & 0x80:
            return v, i


def parse_avrec(buf):
    """-> list of dict(size, literal, skip_coded, cabac, length_parity, last_byte); absent fields omitted."""
    names = {1: "size", 2: "literal", 3: "skip_coded", 4: "cabac", 5: "length_parity", 6: "last_byte"}
    blocks = []
    i = 0
    while i < len(buf):
        tag, i = _varint(buf, i)
        assert tag & 7 == 2, "Recoded has only length-delimited fields"
        n, i = _varint(buf, i)
        body = buf[i:i + n]
        i += n
        if tag >> 3 != 2:
            continue  # metadata (never written, recode.cpp has no mutable_metadata)
        blk = {}
        j = 0
        while j < len(body):
            t, j = _varint(body, j)
            f, wt = t >> 3, t & 7
            if wt == 0:
                v, j = _varint(body, j)
            else:
                ln, j = _varint(body, j)
                v = bytes(body[j:j + ln])
                j += ln
            blk[names[f]] = v
        blocks.append(blk)
    return blocks
