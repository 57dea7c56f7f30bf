"""A minimal BAM / BGZF writer for the tests of hsq_index_bam (SAM/BAM format specification 4.1-4.2), written independently of the
parser in oracle/oracle.py.  Records are dicts: name, flag, ref_id, pos (0-based, -1 unmapped), mapq, cigar [(op_char, len)], seq,
qual (bytes of phred values), next_ref_id, next_pos, tlen, tags (raw bytes)."""
import struct
import zlib

SEQ_CODES = "=ACMGRSVTWYHKDBN"
CIGAR_OPS = "MIDNSHP=X"
EOF_BLOCK = bytes.fromhex("1f8b08040000000000ff0600424302001b0003000000000000000000")


def reg2bin(beg, end):
    end -= 1
    for shift, base in ((14, 4681), (17, 585), (20, 73), (23, 9), (26, 1)):
        if beg >> shift == end >> shift:
            return base + (beg >> shift)
    return 0


def encode_record(r):
    name = r["name"].encode() + b"\0"
    cigar = b"".join(struct.pack("<I", (n << 4) | CIGAR_OPS.index(op)) for op, n in r["cigar"])
    seq = r["seq"]
    packed = bytearray((len(seq) + 1) // 2)
    for i, ch in enumerate(seq):
        packed[i // 2] |= SEQ_CODES.index(ch) << (4 if i % 2 == 0 else 0)
    qual = bytes(r["qual"])
    assert len(qual) == len(seq)
    ref_len = sum(n for op, n in r["cigar"] if op in "MDN=X")
    pos = r["pos"]
    body = struct.pack("<iiBBHHHiiii", r["ref_id"], pos, len(name), r["mapq"], reg2bin(max(pos, 0), max(pos, 0) + max(ref_len, 1)),
                       len(r["cigar"]), r["flag"], len(seq), r["next_ref_id"], r["next_pos"], r["tlen"])
    body += name + cigar + bytes(packed) + qual + r.get("tags", b"")
    return struct.pack("<i", len(body)) + body


def bam_stream(text, refs, records):
    """the uncompressed BAM stream and the offset of each record in it"""
    out = bytearray(b"BAM\1" + struct.pack("<i", len(text)) + text + struct.pack("<i", len(refs)))
    for name, length in refs:
        nm = name.encode() + b"\0"
        out += struct.pack("<i", len(nm)) + nm + struct.pack("<i", length)
    offsets = []
    for r in records:
        offsets.append(len(out))
        out += encode_record(r)
    return bytes(out), offsets


def stored(data, pieces=1):
    """raw deflate of `data` as `pieces` stored blocks (level 0 by hand)"""
    n = len(data)
    cuts = [n * k // pieces for k in range(pieces + 1)]
    out = b""
    for k in range(pieces):
        part = data[cuts[k]:cuts[k + 1]]
        out += bytes([1 if k == pieces - 1 else 0]) + struct.pack("<HH", len(part), len(part) ^ 0xFFFF) + part
    return out


def bgzf(stream, block=65280, level=0, pieces=None, eof=True):
    """BGZF file of the stream.  level 0 through zlib (what htsjdk's Deflater does), or `pieces` stored blocks per BGZF block by hand."""
    out = bytearray()
    for b in range(0, len(stream), block):
        data = stream[b:b + block]
        if pieces:
            cdata = stored(data, pieces)
        else:
            c = zlib.compressobj(level, zlib.DEFLATED, -15)
            cdata = c.compress(data) + c.flush()
        total = 18 + len(cdata) + 8
        assert total <= 65536
        out += b"\x1f\x8b\x08\x04\0\0\0\0\0\xff\x06\0BC\x02\0" + struct.pack("<H", total - 1) + cdata
        out += struct.pack("<II", zlib.crc32(data) & 0xFFFFFFFF, len(data))
    if eof:
        out += EOF_BLOCK
    return bytes(out)
