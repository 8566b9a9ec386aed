#!/usr/bin/env python
"""Generates tests/golden/*.json from the reference's own dbgen (oracle/_ref/libdbgen_ref.so, built from
/root/reference/velox/tpch/gen/dbgen by oracle/Makefile).  Run in the container where /root/reference exists:
    python tests/golden/make_golden.py
Everything written is exact integer arithmetic (cents, percent, days), so it pins the oracle and the CUDA path
independently of floating-point summation order.  The values are the reference's goldens where it has them
(TpchGenTest.cpp:311-351,394-398, TpchConnectorTest.cpp:178-192) and our anchors otherwise."""
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
import lineitem  # noqa: E402


def cents(x):
    return np.rint(np.asarray(x) * 100.0).astype(np.int64)


def main():
    out = {}
    first = lineitem.gen_range(1.0, 0, 200)
    out["first_200_orders_sf1"] = {k: np.asarray(v).tolist() if v.dtype.kind in "iu" else cents(v).tolist() for k, v in first.items()}
    out["first_200_orders_sf1"]["_units"] = "money / quantity / discount / tax columns are value*100 rounded (cents, raw units, percent)"
    t = lineitem.lineitem(0.01, cache=False)
    n = len(t["l_orderkey"])
    out["sf0.01"] = {"rows": int(n), "distinct_orderkeys": int(len(np.unique(t["l_orderkey"]))), "max_orderkey": int(t["l_orderkey"].max())}
    ep, disc, qty, tax = cents(t["l_extendedprice"]), cents(t["l_discount"]), cents(t["l_quantity"]), cents(t["l_tax"])
    q1 = {}
    sel = t["l_shipdate"] <= 10471
    for rf in range(3):
        for ls in range(2):
            m = sel & (t["l_returnflag"] == rf) & (t["l_linestatus"] == ls)
            if not m.any():
                continue
            q1[f"{lineitem.DICTS['l_returnflag'][rf]},{lineitem.DICTS['l_linestatus'][ls]}"] = {
                "count": int(m.sum()), "sum_qty_raw": int(qty[m].sum()), "sum_eprice_cents": int(ep[m].sum()),
                "sum_disc_pct": int(disc[m].sum()),
                "sum_disc_price_cents_x100": int((ep[m] * (100 - disc[m])).sum()),
                "sum_charge_cents_x10000": int((ep[m].astype(object) * (100 - disc[m]) * (100 + tax[m])).sum())}
    out["sf0.01"]["q1"] = q1
    m6 = (t["l_shipdate"] >= 8766) & (t["l_shipdate"] < 9131) & (disc >= 5) & (disc <= 7)
    out["sf0.01"]["q6"] = {"rows": int(m6.sum()), "sum_eprice_disc_cents_pct": int((ep[m6] * disc[m6]).sum())}
    keys, inv = np.unique(t["l_orderkey"], return_inverse=True)
    per = np.zeros(len(keys), dtype=np.int64)
    np.add.at(per, inv, ep)
    out["sf0.01"]["orderkey_groupby"] = {"groups": int(len(keys)), "checksum_key_times_cents": int((keys.astype(object) * per).sum() % (1 << 61)),
                                         "sum_cents": int(per.sum())}
    pairs = {}
    for mo in range(7):
        for ins in range(4):
            m = (t["l_shipmode"] == mo) & (t["l_shipinstruct"] == ins)
            pairs[f"{lineitem.DICTS['l_shipmode'][mo]}|{lineitem.DICTS['l_shipinstruct'][ins]}"] = {
                "count": int(m.sum()), "min_eprice_cents": int(ep[m].min()), "max_eprice_cents": int(ep[m].max()),
                "min_shipdate": int(t["l_shipdate"][m].min()), "max_shipdate": int(t["l_shipdate"][m].max())}
    out["sf0.01"]["shipmode_shipinstruct"] = pairs
    with open(os.path.join(HERE, "lineitem_golden.json"), "w") as f:
        json.dump(out, f, indent=0, sort_keys=True)
    print("wrote", os.path.join(HERE, "lineitem_golden.json"), os.path.getsize(os.path.join(HERE, "lineitem_golden.json")), "bytes")


if __name__ == "__main__":
    main()
