"""Re-wrappings of a baseline JPEG file that every reader must take alike (camera files look like this): extra APPn /
COM segments (Exif, ICC, comments), all Huffman tables in ONE DHT segment and both quantisation tables in ONE DQT,
fill bytes (0xFF 0xFF) in front of markers, bytes behind EOI.  The entropy-coded data is never touched."""


def segments(data: bytes):
    """-> [(marker, payload incl. the two length bytes)] up to and including SOS, then the rest (scan + EOI ...)"""
    assert data[:2] == b"\xff\xd8"
    p, segs = 2, []
    while True:
        assert data[p] == 0xFF
        m = data[p + 1]
        n = (data[p + 2] << 8) | data[p + 3]
        segs.append((m, data[p + 2: p + 2 + n]))
        p += 2 + n
        if m == 0xDA:
            return segs, data[p:]


def rebuild(segs, rest, fill=False):
    out = bytearray(b"\xff\xd8")
    for m, payload in segs:
        if fill:
            out += b"\xff\xff"
        out += bytes([0xFF, m]) + payload
    return bytes(out) + rest


def with_extra_segments(data: bytes) -> bytes:
    segs, rest = segments(data)
    exif = b"Exif\x00\x00" + bytes(range(256)) * 40                      # a 10 KB APP1 with 0xFF bytes inside
    app1 = (0xE1, (len(exif) + 2).to_bytes(2, "big") + exif)
    com = (0xFE, (7 + 2).to_bytes(2, "big") + b"comment")
    return rebuild([segs[0], app1, com] + segs[1:], rest)


def with_merged_tables(data: bytes) -> bytes:
    segs, rest = segments(data)
    dht = b"".join(p[2:] for m, p in segs if m == 0xC4)
    dqt = b"".join(p[2:] for m, p in segs if m == 0xDB)
    out, done = [], set()
    for m, p in segs:
        if m in (0xC4, 0xDB):
            if m not in done:
                body = dht if m == 0xC4 else dqt
                out.append((m, (len(body) + 2).to_bytes(2, "big") + body))
                done.add(m)
        else:
            out.append((m, p))
    return rebuild(out, rest)


def with_fill_bytes_and_tail(data: bytes) -> bytes:
    segs, rest = segments(data)
    return rebuild(segs, rest, fill=True) + b"\x00\x01trailing bytes after EOI\xff\xd8\xff\xe0 not a second image"


VARIANTS = {"extra_segments": with_extra_segments, "merged_tables": with_merged_tables, "fill_and_tail": with_fill_bytes_and_tail}
