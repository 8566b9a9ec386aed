"""-m gpu: the device-side lineitem generator (csrc/tpchgen.cu) against the reference's OWN
dbgen (oracle/_ref/libdbgen_ref.so, compiled from /root/reference by oracle/Makefile): every column bit-identical —
integers exact, DOUBLE columns bit-equal — on whole SF0.01, on windows of SF1 / SF10 (skip-ahead to an order offset,
velox/tpch/gen/DBGenIterator.cpp:70-73) and at a fractional scale factor; plus the reference's goldens (60,175 rows at
SF0.01, velox/connectors/tpch/tests/TpchConnectorTest.cpp:178-192; first orders of velox/tpch/gen/tests/TpchGenTest.cpp:311-351)."""
import numpy as np
import pytest

import lineitem as ref
from helpers import abi

pytestmark = pytest.mark.gpu
ALL = [c for c, _ in ref.COLUMNS]


def compare(engine, sf, offset, orders):
    want = ref.gen_range(sf, offset, orders)
    got = engine.tpch_lineitem(sf, ALL, order_offset=offset, max_orders=orders)
    n = len(want["l_orderkey"])
    assert engine.tpch_lineitem_rows(sf, offset, orders) == n
    for name, dt in ref.COLUMNS:
        g = got[name].cpu().numpy()
        assert g.shape[0] == n, (name, g.shape[0], n)
        w = want[name]
        if dt == np.float64:
            assert np.array_equal(g.view(np.uint64), w.view(np.uint64)), name
        else:
            assert np.array_equal(g, w.astype(g.dtype)), name
    return got


def test_sf001_whole_table_and_goldens(engine):
    got = compare(engine, 0.01, 0, 15000)
    assert len(got["l_orderkey"]) == 60175
    k = got["l_orderkey"].cpu().numpy()
    sd = got["l_shipdate"].cpu().numpy()
    assert k[0] == 1 and (k == 1).sum() == 6 and (k == 2).sum() == 1           # TpchGenTest.cpp:394-398
    assert sd[0] == (np.datetime64("1996-03-13") - np.datetime64("1970-01-01")).astype(int)  # TpchGenTest.cpp:311-351


@pytest.mark.parametrize("sf,offset,orders", [(1, 0, 20000), (1, 1_234_567, 10000), (1, 1_499_000, 5000),
                                               (10, 7_000_001, 10000), (10, 14_995_000, 10000), (10, 123, 10000),
                                               (100, 149_990_000, 10000), (0.5, 100_000, 3000)])
def test_windows_match_the_reference_dbgen(engine, sf, offset, orders):
    compare(engine, sf, offset, orders)


def test_split_past_the_table_is_empty_and_subset_of_columns(engine):
    assert engine.tpch_lineitem_rows(1, 1_500_000, 100) == 0
    a = engine.tpch_lineitem(1, ["l_extendedprice", "l_returnflag"], order_offset=5000, max_orders=4000)
    b = engine.tpch_lineitem(1, ALL, order_offset=5000, max_orders=4000)
    for c in a:
        assert (a[c] == b[c]).all()


def test_sf1_q1_anchor_counts(engine):
    """BASELINE.md section 2 anchors (exact integers, computed with the reference's dbgen): Q1 counts at SF1."""
    d = engine.tpch_lineitem(1, ["l_returnflag", "l_linestatus", "l_shipdate"])
    assert len(d["l_shipdate"]) == 6_001_215
    keep = d["l_shipdate"] <= 10471
    g = (d["l_returnflag"] * 2 + d["l_linestatus"])[keep]
    counts = np.bincount(g.cpu().numpy(), minlength=6)
    assert counts[0] == 1_478_493 and counts[2] == 38_854 and counts[3] == 2_920_374 and counts[4] == 1_478_870  # AF NF NO RF
