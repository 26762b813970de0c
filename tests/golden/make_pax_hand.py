#!/usr/bin/env python
"""Hand-derived PAX fixtures (tests/golden/pax_hand_derived.json), INDEPENDENT of oracle/gotar.c.

Every byte below is put together in this file, field by field, from the published go1.11 archive/tar writer
(writer.go: WriteHeader -> writeUSTARHeader / writePAXHeader, writeRawFile, templateV7Plus, formatPAXRecord;
strconv.go: formatString, formatOctal; format.go: block layout, SetFormat, ComputeChecksum) -- no import of the oracle,
no shared helper. The tests then hold the oracle AND the GPU kernels against these bytes: it pins the PAX path (name of
the PaxHeaders.0 entry, record length self-reference, truncated name field without terminator, zeroed overflow field,
the PAX block's NUL devmajor/devminor) against a second, separately written derivation.

Case A: a 101-byte name without '/' (splitUSTARPath cannot split it)  -> PAX record "path".
Case B: uid 2097152 = 0o10000000 (one more than the 7 octal digits of the USTAR uid field hold) -> PAX record "uid".
"""
import json
import os

HERE = os.path.dirname(os.path.abspath(__file__))
MTIME = 1_546_300_800           # 2019-01-01T00:00:00Z, octal 13412531400


def octal(width: int, v: int) -> bytes:
    # formatOctal (strconv.go): zero-padded octal in width-1 digits, NUL terminated
    s = ("%o" % v).encode()
    assert len(s) <= width - 1
    return b"0" * (width - 1 - len(s)) + s + b"\0"


def put(blk: bytearray, off: int, width: int, val: bytes) -> None:
    assert len(val) <= width
    blk[off:off + len(val)] = val


def checksum(blk: bytearray) -> None:
    # ComputeChecksum: unsigned byte sum with the chksum field read as eight spaces; written as 6 octal digits, NUL, ' '
    blk[148:156] = b" " * 8
    s = sum(blk)
    blk[148:156] = ("%06o" % s).encode() + b"\0 "


def pax_record(key: bytes, val: bytes) -> bytes:
    # formatPAXRecord: "%d %s=%s\n" where %d counts the whole record including its own digits
    body = b" " + key + b"=" + val + b"\n"
    n = len(body)
    n += len(str(n))
    if len(str(n)) + len(body) != n:          # the digit count changed the digit count
        n = len(body) + len(str(n))
    rec = str(n).encode() + body
    assert len(rec) == n
    return rec


def pax_block(entry_name: bytes, records: bytes) -> bytes:
    # writeRawFile(name, data, TypeXHeader, FormatPAX): only name, mode 0, uid 0, gid 0, size, mtime 0, typeflag and
    # the magic/version are set; uname, gname, devmajor, devminor and prefix stay NUL. The name is cut to 100 bytes and
    # a trailing '/' trimmed.
    name = entry_name[:100].rstrip(b"/")
    b = bytearray(512)
    put(b, 0, 100, name)
    put(b, 100, 8, octal(8, 0))
    put(b, 108, 8, octal(8, 0))
    put(b, 116, 8, octal(8, 0))
    put(b, 124, 12, octal(12, len(records)))
    put(b, 136, 12, octal(12, 0))
    b[156] = ord("x")
    put(b, 257, 6, b"ustar\0")
    put(b, 263, 2, b"00")
    checksum(b)
    pad = (-len(records)) % 512
    return bytes(b) + records + b"\0" * pad


def main_block(name_field: bytes, mode: int, uid_field: int, gid: int, size: int) -> bytes:
    # templateV7Plus + the USTAR fields: name (copied without terminator when it fills the field), numeric fields as
    # octal, uname/gname, devmajor/devminor ALWAYS formatted ("0000000\0"), prefix empty
    b = bytearray(512)
    put(b, 0, 100, name_field[:100])
    put(b, 100, 8, octal(8, mode))
    put(b, 108, 8, octal(8, uid_field))
    put(b, 116, 8, octal(8, gid))
    put(b, 124, 12, octal(12, size))
    put(b, 136, 12, octal(12, MTIME))
    b[156] = ord("0")
    put(b, 257, 6, b"ustar\0")
    put(b, 263, 2, b"00")
    put(b, 265, 32, b"dev")
    put(b, 297, 32, b"dev")
    put(b, 329, 8, octal(8, 0))
    put(b, 337, 8, octal(8, 0))
    checksum(b)
    return bytes(b)


def stream(pax: bytes, main: bytes, payload: bytes) -> bytes:
    pad = (-len(payload)) % 512
    return pax + main + payload + b"\0" * pad + b"\0" * 1024      # Writer.Close(): two zero blocks


def main():
    out = {}
    # A: name = 101 x 'n', payload "a", mode 0644, uid = gid = 1000
    name = b"n" * 101
    rec = pax_record(b"path", name)
    assert rec[:9] == b"111 path=" and len(rec) == 111
    a = stream(pax_block(b"PaxHeaders.0/" + name, rec), main_block(name, 0o644, 1000, 1000, 1), b"a")
    out["name_101_nosplit"] = {"manifest": {"name": name.decode(), "data": "a", "mode": 0o644, "uid": 1000, "gid": 1000,
                                            "uname": "dev", "gname": "dev", "mtime": MTIME},
                               "stream_hex": a.hex(), "bytes": len(a)}
    # B: name "f", payload "x", uid 2097152: the record carries the uid, the USTAR field is written as zero
    rec = pax_record(b"uid", b"2097152")
    assert rec == b"15 uid=2097152\n"
    b = stream(pax_block(b"PaxHeaders.0/f", rec), main_block(b"f", 0o644, 0, 1000, 1), b"x")
    out["uid_2097152"] = {"manifest": {"name": "f", "data": "x", "mode": 0o644, "uid": 0o10000000, "gid": 1000,
                                       "uname": "dev", "gname": "dev", "mtime": MTIME},
                          "stream_hex": b.hex(), "bytes": len(b)}
    with open(os.path.join(HERE, "pax_hand_derived.json"), "w") as f:
        json.dump(out, f, indent=1, sort_keys=True)
    print({k: v["bytes"] for k, v in out.items()})


if __name__ == "__main__":
    main()
