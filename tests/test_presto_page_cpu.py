"""The host library's PrestoPage bytes (velox-1_b200/host/PrestoPage.cpp, dumped by `host_tests presto`) against the
independent Python restatement of velox/serializers/PrestoSerializer.cpp's layout (oracle/presto_page.py).  No GPU needed.
The hand-computed byte strings and the round trips live in `host_tests cpu` ("PrestoPage" section)."""
import os
import subprocess
import tempfile

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HOST_TESTS = os.path.join(ROOT, "tests", "host", "host_tests")


def test_known_small_pages():
    import presto_page as pp
    # [1, NULL, 3] as BIGINT: 40 data bytes, null row 1 = bit 6 of the first null byte
    got = pp.page([("i64", [1, None, 3])], 3)
    assert got[:21] == (3).to_bytes(4, "little") + b"\x00" + (40).to_bytes(4, "little") * 2 + bytes(8)
    assert got[21:] == (1).to_bytes(4, "little") + (10).to_bytes(4, "little") + b"LONG_ARRAY" + (3).to_bytes(4, "little") + \
        b"\x01\x40" + (1).to_bytes(8, "little") + (3).to_bytes(8, "little")
    # nine rows, the last one NULL: second null byte has its top bit set
    assert pp.column("i8", [0] * 8 + [None])[-10:-8] == b"\x00\x80"


@pytest.mark.skipif(not os.path.exists(HOST_TESTS), reason="tests/host/host_tests not built (build() makes it)")
def test_host_library_bytes_match_the_restatement():
    import presto_page as pp
    with tempfile.TemporaryDirectory() as tmp:
        path = os.path.join(tmp, "page.bin")
        subprocess.run([HOST_TESTS, "presto", path], check=True, stdout=subprocess.PIPE)
        got = open(path, "rb").read()
    cols = [
        ("i64", [5, -1, 1 << 40, None, 77]),
        ("varchar", [b"AIR", b"", b"REG AIR", None, b"TRUCK"]),
        (("row", ["f64", "i64"]), [(0.5, 1), None, (1e300, 3), (3.25, 4), (2.0, 5)]),
        ("i32", [1, 2, 3, 4, 5]),
        ("bool", [True, False, None, False, True]),
    ]
    want = pp.page(cols, 5, with_checksum=True)
    assert got == want, (len(got), len(want))
