"""TEST INFRASTRUCTURE — Python restatement of the PrestoPage wire format (velox/serializers/PrestoSerializer.cpp), written
independently of velox-1_b200/host/PrestoPage.cpp to pin its bytes.  Only tests/ may import this.

  page   (:1947-1949)  numRows i32 | codec i8 | uncompressedSize i32 | size i32 | checksum i64 | data
  data                 numColumns i32 | columns
  column (:1018-1023)  len i32 | encoding name | body
  fixed  (:1142-1146)  positions i32 | nulls | non-null values        names: typeToEncodingName (:102-137)
  varchar(:1133-1140)  positions i32 | end offsets i32[n] | nulls | totalBytes i32 | bytes
  row    (:1097-1112)  numChildren i32 | children over non-null rows | positions i32 | offsets i32[n + 1] | nulls
  nulls  (:1149-1158)  0, or 1 + bits (1 = NULL), most significant bit first (nulls_ is a reversed bit stream, :943)
  checksum (:34-45)    zlib CRC-32 over data, the codec byte, numRows, uncompressedSize; codec bit 4 set

A column is (type, values) with None for NULL; types: 'bool', 'i8', 'i16', 'i32', 'i64', 'f32', 'f64', 'varchar',
('row', [child types]) whose values are tuples or None."""
import struct
import zlib

_FIXED = {"bool": ("BYTE_ARRAY", "<b"), "i8": ("BYTE_ARRAY", "<b"), "i16": ("SHORT_ARRAY", "<h"), "i32": ("INT_ARRAY", "<i"),
          "f32": ("INT_ARRAY", "<f"), "i64": ("LONG_ARRAY", "<q"), "f64": ("LONG_ARRAY", "<d")}


def _name(n):
    return struct.pack("<i", len(n)) + n.encode()


def _nulls(values):
    if all(v is not None for v in values):
        return b"\x00"
    bits = bytearray((len(values) + 7) // 8)
    for i, v in enumerate(values):
        if v is None:
            bits[i // 8] |= 0x80 >> (i % 8)
    return b"\x01" + bytes(bits)


def column(typ, values):
    n = len(values)
    if isinstance(typ, tuple):  # ('row', [child types])
        present = [v for v in values if v is not None]
        out = _name("ROW") + struct.pack("<i", len(typ[1]))
        for c, ct in enumerate(typ[1]):
            out += column(ct, [v[c] for v in present])
        out += struct.pack("<i", n) + struct.pack("<i", 0)
        off = 0
        for v in values:
            off += 0 if v is None else 1
            out += struct.pack("<i", off)
        return out + _nulls(values)
    if typ == "varchar":
        blob, ends = b"", b""
        for v in values:
            if v is not None:
                blob += v
            ends += struct.pack("<i", len(blob))
        return _name("VARIABLE_WIDTH") + struct.pack("<i", n) + ends + _nulls(values) + struct.pack("<i", len(blob)) + blob
    enc, fmt = _FIXED[typ]
    body = b"".join(struct.pack(fmt, (1 if v else 0) if typ == "bool" else v) for v in values if v is not None)
    return _name(enc) + struct.pack("<i", n) + _nulls(values) + body


def page(columns, num_rows, with_checksum=False):
    data = struct.pack("<i", len(columns)) + b"".join(column(t, v) for t, v in columns)
    codec = 4 if with_checksum else 0
    checksum = 0
    if with_checksum:
        checksum = zlib.crc32(struct.pack("<bii", codec, num_rows, len(data)), zlib.crc32(data))
    return struct.pack("<ibiiq", num_rows, codec, len(data), len(data), checksum) + data
