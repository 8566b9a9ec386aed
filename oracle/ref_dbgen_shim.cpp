// TEST INFRASTRUCTURE (oracle) — not part of the product path.
//
// Thin C-ABI driver around the reference's own vendored TPC-H dbgen, compiled from the sources
// where they lie under /root/reference/velox/tpch/gen/dbgen (see oracle/Makefile, target
// _ref/libdbgen_ref.so).  It restates only the *transposition* that
// velox/tpch/gen/TpchGen.cpp:401-489 (genTpchLineItem) performs around dbgen's mk_order:
//   * offsets / maxRows are counted in ORDERS (TpchGen.cpp:439-443),
//   * money / quantity / discount / tax = (double)int64 * 0.01        (TpchGen.cpp:79-81),
//   * dates = days since 1970-01-01 of the "YYYY-MM-DD" string          (TpchGen.cpp:83-85),
//   * iterator protocol sd_order + sd_line, row_start / mk_order / row_stop_h
//     (velox/tpch/gen/DBGenIterator.cpp:70-73, 100-104), scale factor < 1 clamps dbgen's
//     integer scale to 1 (DBGenIterator.cpp:55-59); row counts follow TpchGen.cpp:30-41,142-143.
// VARCHAR columns are returned as small integer codes (index into the dbgen distribution's
// lexicographically sorted value list) — the strings themselves are returned by
// ref_lineitem_dictionary().
#include <cstdint>
#include <cstring>
#include <cstdlib>
#include <mutex>
#include <string>

#include "dbgen/dbgen_gunk.hpp"
#include "dbgen/dss.h"
#include "dbgen/dsstypes.h"

namespace {
std::once_flag gInit;
void initOnce() {
  std::call_once(gInit, [] {
    DBGenContext ctx;
    load_dists(10 * 1024 * 1024, &ctx); // same text-pool size as DBGenIterator.cpp:40-41
  });
}

// Days from civil (proleptic Gregorian), 1970-01-01 == 0.
int32_t daysFromCivil(int y, unsigned m, unsigned d) {
  y -= m <= 2;
  const int era = (y >= 0 ? y : y - 399) / 400;
  const unsigned yoe = static_cast<unsigned>(y - era * 400);
  const unsigned doy = (153 * (m + (m > 2 ? -3 : 9)) + 2) / 5 + d - 1;
  const unsigned doe = yoe * 365 + yoe / 4 - yoe / 100 + doy;
  return era * 146097 + static_cast<int32_t>(doe) - 719468;
}
int32_t toDate(const char* s) { // "YYYY-MM-DD"
  int y = atoi(s);
  int m = atoi(s + 5);
  int d = atoi(s + 8);
  return daysFromCivil(y, m, d);
}

const char* kShipInstruct[] = {"COLLECT COD", "DELIVER IN PERSON", "NONE", "TAKE BACK RETURN"};
const char* kShipMode[] = {"AIR", "FOB", "MAIL", "RAIL", "REG AIR", "SHIP", "TRUCK"};
const char* kReturnFlag[] = {"A", "N", "R"};
const char* kLineStatus[] = {"F", "O"};

template <size_t N>
int32_t codeOf(const char* (&dict)[N], const char* s) {
  for (size_t i = 0; i < N; ++i) {
    if (strcmp(dict[i], s) == 0) return (int32_t)i;
  }
  return -1;
}
} // namespace

extern "C" {

// Number of orders at a scale factor (TpchGen.cpp:142-143: 1'500'000 * sf).
int64_t ref_order_count(double sf) {
  return (int64_t)(1500000 * sf);
}

struct RefLineitemColumns {
  int64_t* l_orderkey;
  int64_t* l_partkey;
  int64_t* l_suppkey;
  int32_t* l_linenumber;
  double* l_quantity;
  double* l_extendedprice;
  double* l_discount;
  double* l_tax;
  int32_t* l_returnflag; // code into {"A","N","R"}
  int32_t* l_linestatus; // code into {"F","O"}
  int32_t* l_shipdate;
  int32_t* l_commitdate;
  int32_t* l_receiptdate;
  int32_t* l_shipinstruct; // code
  int32_t* l_shipmode; // code
};

// Generates the lineitems of orders [orderOffset, orderOffset + maxOrders) (clamped to the table)
// into caller-provided arrays with room for 7 * maxOrders rows; any column pointer may be null.
// Returns the number of lineitem rows produced.
int64_t ref_gen_lineitem(double sf, int64_t orderOffset, int64_t maxOrders, RefLineitemColumns* out) {
  initOnce();
  DBGenContext ctx;
  if (sf < MIN_SCALE && sf > 0) {
    ctx.scale_factor = 1;
  } else {
    ctx.scale_factor = (long)sf;
  }
  int64_t total = ref_order_count(sf);
  if (orderOffset >= total) return 0;
  int64_t n = std::min<int64_t>(total - orderOffset, maxOrders);
  sd_order(ORDER, orderOffset, &ctx);
  sd_line(LINE, orderOffset, &ctx);
  order_t order;
  int64_t rows = 0;
  for (int64_t i = 0; i < n; ++i) {
    row_start(ORDER, &ctx);
    mk_order(i + orderOffset + 1, &order, &ctx, 0);
    row_stop_h(ORDER, &ctx);
    for (long l = 0; l < order.lines; ++l) {
      const line_t& line = order.l[l];
      int64_t r = rows + l;
      if (out->l_orderkey) out->l_orderkey[r] = line.okey;
      if (out->l_partkey) out->l_partkey[r] = line.partkey;
      if (out->l_suppkey) out->l_suppkey[r] = line.suppkey;
      if (out->l_linenumber) out->l_linenumber[r] = (int32_t)line.lcnt;
      if (out->l_quantity) out->l_quantity[r] = (double)line.quantity * 0.01;
      if (out->l_extendedprice) out->l_extendedprice[r] = (double)line.eprice * 0.01;
      if (out->l_discount) out->l_discount[r] = (double)line.discount * 0.01;
      if (out->l_tax) out->l_tax[r] = (double)line.tax * 0.01;
      if (out->l_returnflag) {
        char s[2] = {line.rflag[0], 0};
        out->l_returnflag[r] = codeOf(kReturnFlag, s);
      }
      if (out->l_linestatus) {
        char s[2] = {line.lstatus[0], 0};
        out->l_linestatus[r] = codeOf(kLineStatus, s);
      }
      if (out->l_shipdate) out->l_shipdate[r] = toDate(line.sdate);
      if (out->l_commitdate) out->l_commitdate[r] = toDate(line.cdate);
      if (out->l_receiptdate) out->l_receiptdate[r] = toDate(line.rdate);
      if (out->l_shipinstruct) out->l_shipinstruct[r] = codeOf(kShipInstruct, line.shipinstruct);
      if (out->l_shipmode) out->l_shipmode[r] = codeOf(kShipMode, line.shipmode);
    }
    rows += order.lines;
  }
  return rows;
}

// which: 0 returnflag, 1 linestatus, 2 shipinstruct, 3 shipmode.  Returns count; entry i in out[i].
int32_t ref_lineitem_dictionary(int32_t which, const char** out, int32_t cap) {
  const char** d = nullptr;
  int32_t n = 0;
  switch (which) {
    case 0: d = kReturnFlag; n = 3; break;
    case 1: d = kLineStatus; n = 2; break;
    case 2: d = kShipInstruct; n = 4; break;
    case 3: d = kShipMode; n = 7; break;
    default: return -1;
  }
  for (int32_t i = 0; i < n && i < cap; ++i) out[i] = d[i];
  return n;
}
} // extern "C"
