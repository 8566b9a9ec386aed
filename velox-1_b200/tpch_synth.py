"""Lineitem-shaped synthetic columns generated directly in HBM (torch as the allocator / RNG).

Same schema, value domains and derivation rules as the reference's generator
(velox/tpch/gen/TpchGen.cpp:401-489 over dbgen's mk_order, velox/tpch/gen/dbgen/build.cpp:110-223):
quantity 1..50, discount 0..10, tax 0..8 (all stored as integer * 0.01, TpchGen.cpp:79-81),
extendedprice = retail price(partkey) * quantity, shipdate = orderdate + 1..121,
receiptdate = shipdate + 1..30, returnflag = R/A when receiptdate <= 1995-06-17 else N,
linestatus = F when shipdate <= 1995-06-17 else O, 1..7 lines per order, rows sorted by a sparse
l_orderkey.  The random streams are torch's, NOT dbgen's: bench.py uses this for throughput;
parity tests use the reference's own dbgen (oracle/_ref).
"""
import torch

ROWS = {1: 6_001_215, 10: 59_986_052, 100: 600_037_902}  # velox/tpch/gen/TpchGen.cpp:30-41
STARTDATE, CURRENTDATE, ENDDATE = 8035, 9298, 10591      # 1992-01-01, 1995-06-17, 1998-12-31


def row_count(sf: float) -> int:
    return ROWS.get(int(sf), int(6_001_215 * sf)) if sf >= 1 else int(6_001_215 * sf)


def lineitem(sf: float, columns, rows: int = None, order_offset: int = 0, seed: int = 0, device="cuda"):
    """dict name -> tensor for the requested lineitem columns; `rows` defaults to the table size."""
    g = torch.Generator(device=device)
    g.manual_seed(seed + 7919 * order_offset)
    n = row_count(sf) if rows is None else rows
    orders = n // 4 + n // 50 + 16  # mean 4 lines per order, plus slack
    lines = torch.randint(1, 8, (orders,), generator=g, device=device)
    ends = torch.cumsum(lines, 0)
    assert int(ends[-1]) >= n, "not enough orders drawn"
    order_of_row = torch.searchsorted(ends, torch.arange(n, device=device), right=True)
    out = {}
    odate = torch.randint(STARTDATE, ENDDATE - 151 + 1, (orders,), generator=g, device=device)
    od = odate[order_of_row]
    ship = od + torch.randint(1, 122, (n,), generator=g, device=device)
    if "l_orderkey" in columns:
        idx = order_of_row + order_offset
        out["l_orderkey"] = ((idx >> 3) << 5) + (idx & 7) + 1  # mk_sparse: 8 keys used out of every 32
    if "l_linenumber" in columns:
        starts = ends - lines
        out["l_linenumber"] = (torch.arange(n, device=device) - starts[order_of_row] + 1).to(torch.int32)
    qty = torch.randint(1, 51, (n,), generator=g, device=device)
    if "l_quantity" in columns:
        out["l_quantity"] = qty.to(torch.float64) * 0.01
    if "l_extendedprice" in columns or "l_partkey" in columns:
        part = torch.randint(1, int(200_000 * max(sf, 1)) + 1, (n,), generator=g, device=device)
        if "l_partkey" in columns:
            out["l_partkey"] = part
        rprice = 90000 + ((part // 10) % 20001) + 100 * (part % 1000)
        if "l_extendedprice" in columns:
            out["l_extendedprice"] = (rprice * qty).to(torch.float64) * 0.01
        del part, rprice
    if "l_discount" in columns:
        out["l_discount"] = torch.randint(0, 11, (n,), generator=g, device=device).to(torch.float64) * 0.01
    if "l_tax" in columns:
        out["l_tax"] = torch.randint(0, 9, (n,), generator=g, device=device).to(torch.float64) * 0.01
    if "l_shipdate" in columns:
        out["l_shipdate"] = ship.to(torch.int32)
    if "l_commitdate" in columns:
        out["l_commitdate"] = (od + torch.randint(30, 91, (n,), generator=g, device=device)).to(torch.int32)
    if "l_receiptdate" in columns or "l_returnflag" in columns:
        receipt = ship + torch.randint(1, 31, (n,), generator=g, device=device)
        if "l_receiptdate" in columns:
            out["l_receiptdate"] = receipt.to(torch.int32)
        if "l_returnflag" in columns:  # ids into {"A","N","R"}
            ra = torch.randint(0, 2, (n,), generator=g, device=device) * 2
            out["l_returnflag"] = torch.where(receipt <= CURRENTDATE, ra, torch.ones_like(ra)).to(torch.int32)
    if "l_linestatus" in columns:      # ids into {"F","O"}
        out["l_linestatus"] = (ship > CURRENTDATE).to(torch.int32)
    if "l_shipinstruct" in columns:
        out["l_shipinstruct"] = torch.randint(0, 4, (n,), generator=g, device=device).to(torch.int32)
    if "l_shipmode" in columns:
        out["l_shipmode"] = torch.randint(0, 7, (n,), generator=g, device=device).to(torch.int32)
    return out
