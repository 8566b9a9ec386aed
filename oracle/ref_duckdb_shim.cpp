// TEST INFRASTRUCTURE (oracle) — not part of the product path.
//
// Thin C-ABI driver around the reference's own vendored DuckDB (velox/external/duckdb, compiled
// from the sources where they lie; see oracle/Makefile, target _ref/libduckdb_ref.so).  The
// reference's operator tests check every FilterProject / HashAggregation answer against this
// engine (velox/exec/tests/utils/QueryAssertions.cpp: DuckDbQueryRunner::createTable appends the
// input RowVectors row by row, ::execute runs the SQL and materialises the result), so running the
// same SQL here pins our restatement (oracle/velox_cpu.cpp, np_oracle.py) to the reference's own
// test oracle.  It is also the "second CPU baseline" of SURVEY.md section 8(d)(ii).
//
// The API is deliberately generic: create a table from column arrays, run SQL, read the result
// back as doubles (+ a validity byte per cell).
#include <cstdint>
#include <cstring>
#include <memory>
#include <string>
#include <vector>

#include "velox/external/duckdb/duckdb.hpp"

namespace {
struct Session {
  duckdb::DuckDB db;
  duckdb::Connection con;
  std::string error;
  std::unique_ptr<duckdb::MaterializedQueryResult> last;
  Session() : db(nullptr), con(db) {}
};
} // namespace

extern "C" {

// Column kinds understood by rdb_append_columns.
enum { RDB_INTEGER = 0, RDB_BIGINT = 1, RDB_DOUBLE = 2, RDB_DATE = 3, RDB_VARCHAR_CODES = 4 };

void* rdb_open(int threads) {
  try {
    auto* s = new Session();
    if (threads > 0) {
      s->con.Query("PRAGMA threads=" + std::to_string(threads));
    }
    return s;
  } catch (...) {
    return nullptr;
  }
}

void rdb_close(void* h) {
  delete static_cast<Session*>(h);
}

const char* rdb_error(void* h) {
  return static_cast<Session*>(h)->error.c_str();
}

// Runs a statement whose result is not needed (DDL, PRAGMA).  0 = ok.
int rdb_exec(void* h, const char* sql) {
  auto* s = static_cast<Session*>(h);
  try {
    auto r = s->con.Query(sql);
    if (!r->success) {
      s->error = r->error;
      return 1;
    }
    return 0;
  } catch (const std::exception& e) {
    s->error = e.what();
    return 1;
  }
}

// Appends nrows rows to an existing table.  kinds[c] selects how cols[c] is read:
//   INTEGER / DATE: int32_t[]; BIGINT: int64_t[]; DOUBLE: double[];
//   VARCHAR_CODES: int32_t[] of indices into dicts[c] (a const char*[]); a negative code is NULL.
// not_null[c] may be null (all valid) or a uint8_t[] (0 = NULL).
int rdb_append_columns(void* h, const char* table, int ncols, const int* kinds, const void* const* cols,
                       const void* const* dicts, const uint8_t* const* not_null, int64_t nrows) {
  auto* s = static_cast<Session*>(h);
  try {
    duckdb::Appender app(s->con, table);
    for (int64_t r = 0; r < nrows; ++r) {
      app.BeginRow();
      for (int c = 0; c < ncols; ++c) {
        if (not_null && not_null[c] && !not_null[c][r]) {
          app.Append<std::nullptr_t>(nullptr);
          continue;
        }
        switch (kinds[c]) {
          case RDB_INTEGER:
            app.Append<int32_t>(static_cast<const int32_t*>(cols[c])[r]);
            break;
          case RDB_BIGINT:
            app.Append<int64_t>(static_cast<const int64_t*>(cols[c])[r]);
            break;
          case RDB_DOUBLE:
            app.Append<double>(static_cast<const double*>(cols[c])[r]);
            break;
          case RDB_DATE:
            app.Append<duckdb::date_t>(duckdb::date_t(static_cast<const int32_t*>(cols[c])[r]));
            break;
          case RDB_VARCHAR_CODES: {
            int32_t code = static_cast<const int32_t*>(cols[c])[r];
            if (code < 0) {
              app.Append<std::nullptr_t>(nullptr);
            } else {
              app.Append<const char*>(static_cast<const char* const*>(dicts[c])[code]);
            }
            break;
          }
          default:
            s->error = "rdb_append_columns: unknown column kind";
            return 1;
        }
      }
      app.EndRow();
    }
    app.Close();
    return 0;
  } catch (const std::exception& e) {
    s->error = e.what();
    return 1;
  }
}

// Runs a query and keeps its materialised result.  Returns the row count (-1 on error) and the
// column count through *ncols.
int64_t rdb_query(void* h, const char* sql, int* ncols) {
  auto* s = static_cast<Session*>(h);
  try {
    s->last = s->con.Query(sql);
    if (!s->last->success) {
      s->error = s->last->error;
      s->last.reset();
      return -1;
    }
    *ncols = static_cast<int>(s->last->types.size());
    return static_cast<int64_t>(s->last->collection.Count());
  } catch (const std::exception& e) {
    s->error = e.what();
    return -1;
  }
}

// Copies the kept result out, row-major: values[r * ncols + c] as double (VARCHAR cells are
// returned through rdb_result_string instead and read 0 here), valid[...] = 0 for NULL.
int rdb_result_doubles(void* h, double* values, uint8_t* valid) {
  auto* s = static_cast<Session*>(h);
  if (!s->last) {
    s->error = "no result";
    return 1;
  }
  try {
    const auto rows = s->last->collection.Count();
    const auto ncols = s->last->types.size();
    for (duckdb::idx_t r = 0; r < rows; ++r) {
      for (duckdb::idx_t c = 0; c < ncols; ++c) {
        auto v = s->last->GetValue(c, r);
        const bool null = v.IsNull();
        valid[r * ncols + c] = null ? 0 : 1;
        double d = 0;
        if (!null && s->last->types[c].id() != duckdb::LogicalTypeId::VARCHAR) {
          if (s->last->types[c].id() == duckdb::LogicalTypeId::DATE) {
            d = static_cast<double>(v.GetValue<duckdb::date_t>().days);
          } else {
            d = v.GetValue<double>();
          }
        }
        values[r * ncols + c] = d;
      }
    }
    return 0;
  } catch (const std::exception& e) {
    s->error = e.what();
    return 1;
  }
}

// Exact 64-bit read of an integer cell (sums of BIGINT columns come back as HUGEINT; values that
// do not fit int64 make this fail).
int rdb_result_int64(void* h, int64_t row, int col, int64_t* out) {
  auto* s = static_cast<Session*>(h);
  try {
    auto v = s->last->GetValue(col, row);
    if (v.IsNull()) {
      return 2;
    }
    *out = v.GetValue<int64_t>();
    return 0;
  } catch (const std::exception& e) {
    s->error = e.what();
    return 1;
  }
}

// Copies a VARCHAR cell into buf (NUL terminated, truncated to cap - 1).  Returns its length, -1
// for NULL.
int rdb_result_string(void* h, int64_t row, int col, char* buf, int cap) {
  auto* s = static_cast<Session*>(h);
  try {
    auto v = s->last->GetValue(col, row);
    if (v.IsNull()) {
      return -1;
    }
    std::string str = v.ToString();
    int n = static_cast<int>(str.size());
    int m = n < cap - 1 ? n : cap - 1;
    memcpy(buf, str.data(), m);
    buf[m] = 0;
    return n;
  } catch (const std::exception& e) {
    s->error = e.what();
    return -2;
  }
}

} // extern "C"
