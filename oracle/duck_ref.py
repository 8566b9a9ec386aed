"""TEST INFRASTRUCTURE — ctypes wrapper over oracle/_ref/libduckdb_ref.so, the reference's own vendored DuckDB
(velox/external/duckdb) compiled by oracle/Makefile.  The reference's operator tests assert every answer against this
engine (velox/exec/tests/utils/QueryAssertions.cpp: DuckDbQueryRunner), so the same SQL here pins our oracle to the
reference's test oracle and gives SURVEY.md section 8(d)(ii)'s second CPU baseline.  Never imported by the product."""
import ctypes as C
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
SO = os.path.join(HERE, "_ref", "libduckdb_ref.so")

INTEGER, BIGINT, DOUBLE, DATE, VARCHAR_CODES = range(5)
_SQL_TYPE = {INTEGER: "INTEGER", BIGINT: "BIGINT", DOUBLE: "DOUBLE", DATE: "DATE", VARCHAR_CODES: "VARCHAR"}
_NP_TYPE = {INTEGER: np.int32, BIGINT: np.int64, DOUBLE: np.float64, DATE: np.int32, VARCHAR_CODES: np.int32}

_lib = None


def available():
    return os.path.exists(SO)


def lib():
    global _lib
    if _lib is None:
        if not available():
            raise FileNotFoundError(f"{SO} missing: run `make -C oracle -j8 ref` where /root/reference exists")
        L = C.CDLL(SO)
        L.rdb_open.restype = C.c_void_p
        L.rdb_open.argtypes = [C.c_int]
        L.rdb_close.argtypes = [C.c_void_p]
        L.rdb_error.restype = C.c_char_p
        L.rdb_error.argtypes = [C.c_void_p]
        L.rdb_exec.argtypes = [C.c_void_p, C.c_char_p]
        L.rdb_append_columns.argtypes = [C.c_void_p, C.c_char_p, C.c_int, C.POINTER(C.c_int), C.POINTER(C.c_void_p),
                                         C.POINTER(C.c_void_p), C.POINTER(C.c_void_p), C.c_int64]
        L.rdb_query.restype = C.c_int64
        L.rdb_query.argtypes = [C.c_void_p, C.c_char_p, C.POINTER(C.c_int)]
        L.rdb_result_doubles.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
        L.rdb_result_int64.argtypes = [C.c_void_p, C.c_int64, C.c_int, C.POINTER(C.c_int64)]
        L.rdb_result_string.argtypes = [C.c_void_p, C.c_int64, C.c_int, C.c_char_p, C.c_int]
        _lib = L
    return _lib


class DuckError(Exception):
    pass


class Session:
    """One in-memory DuckDB database + connection."""

    def __init__(self, threads=0):
        self.L = lib()
        self.h = self.L.rdb_open(threads)
        if not self.h:
            raise DuckError("rdb_open failed")

    def close(self):
        if self.h:
            self.L.rdb_close(self.h)
            self.h = None

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def _fail(self):
        raise DuckError(self.L.rdb_error(self.h).decode(errors="replace"))

    def execute(self, sql):
        if self.L.rdb_exec(self.h, sql.encode()):
            self._fail()

    def create_table(self, name, columns):
        """columns: list of (name, kind, ndarray[, dictionary list][, not_null ndarray]) — the table is created and
        filled the way DuckDbQueryRunner::createTable does it (CREATE TABLE + Appender, one row at a time)."""
        ddl = ", ".join(f"{c[0]} {_SQL_TYPE[c[1]]}" for c in columns)
        self.execute(f"CREATE TABLE {name} ({ddl})")
        n = len(columns[0][2])
        k = len(columns)
        kinds = (C.c_int * k)(*[c[1] for c in columns])
        cols = (C.c_void_p * k)()
        dicts = (C.c_void_p * k)()
        nn = (C.c_void_p * k)()
        keep = []
        for i, c in enumerate(columns):
            a = np.ascontiguousarray(c[2], dtype=_NP_TYPE[c[1]])
            assert len(a) == n
            keep.append(a)
            cols[i] = a.ctypes.data
            d = c[3] if len(c) > 3 else None
            if d is not None:
                arr = (C.c_char_p * len(d))(*[s.encode() for s in d])
                keep.append(arr)
                dicts[i] = C.cast(arr, C.c_void_p)
            m = c[4] if len(c) > 4 else None
            if m is not None:
                mm = np.ascontiguousarray(m, dtype=np.uint8)
                keep.append(mm)
                nn[i] = mm.ctypes.data
        if self.L.rdb_append_columns(self.h, name.encode(), k, kinds, cols, dicts, nn, n):
            self._fail()

    def query(self, sql, strings=()):
        """Returns (values float64[rows, cols], valid bool[rows, cols], {col: [str, ...]} for the columns in `strings`)."""
        nc = C.c_int(0)
        rows = self.L.rdb_query(self.h, sql.encode(), C.byref(nc))
        if rows < 0:
            self._fail()
        vals = np.zeros((rows, nc.value), dtype=np.float64)
        valid = np.zeros((rows, nc.value), dtype=np.uint8)
        if rows and self.L.rdb_result_doubles(self.h, vals.ctypes.data, valid.ctypes.data):
            self._fail()
        strs = {}
        buf = C.create_string_buffer(256)
        for c in strings:
            out = []
            for r in range(rows):
                ln = self.L.rdb_result_string(self.h, r, c, buf, 256)
                out.append(None if ln < 0 else buf.value.decode())
            strs[c] = out
        return vals, valid.astype(bool), strs

    def cell_int64(self, row, col):
        v = C.c_int64(0)
        rc = self.L.rdb_result_int64(self.h, row, col, C.byref(v))
        if rc == 2:
            return None
        if rc:
            self._fail()
        return v.value
