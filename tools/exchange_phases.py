#!/usr/bin/env python
"""Single-GPU timing of the device phases of the HASH exchange (no NCCL): extract packed rows, partition
into 8 parts, merge the rows into a fresh table.  C3 shape (SF10, 15 M groups)."""
import importlib, json, os, sys, time
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
pkg = importlib.import_module("velox-1_b200"); abi, ex = pkg.abi, pkg.expr
abi.check(abi.lib().vgpu_init(0))
engine = importlib.import_module("velox-1_b200.engine"); tpch = importlib.import_module("velox-1_b200.tpch_synth")
n = tpch.row_count(10.0)
data = tpch.lineitem(10.0, ["l_orderkey", "l_extendedprice"])
I64, D = abi.BIGINT, abi.DOUBLE
cols = [engine.DeviceColumn(I64, abi.FLAT, data["l_orderkey"], length=n), engine.DeviceColumn(D, abi.FLAT, data["l_extendedprice"], length=n)]
empty = [engine.DeviceColumn(I64, abi.FLAT, data["l_orderkey"][:0], length=0), engine.DeviceColumn(D, abi.FLAT, data["l_extendedprice"][:0], length=0)]
prog = engine.Program(None, [ex.col(0, I64), ex.col(1, D)])
groups = int(torch.unique_consecutive(data["l_orderkey"]).numel())
accs = [(abi.ACC_SUM_F64, 1)]
partial = engine.Aggregation(abi.AGG_HASH, [(0, 64, 0, 0)], accs, expected_groups=groups)
final = engine.Aggregation(abi.AGG_HASH, [(0, 64, 0, 0)], accs, expected_groups=groups)
def wall(fn):
    torch.cuda.synchronize(); t = time.perf_counter(); r = fn(); torch.cuda.synchronize(); return (time.perf_counter() - t) * 1e3, r
res = {}
for it in range(3):
    partial.reset(); final.reset(); final.add_input(prog, empty, 0)
    res["add_input"], _ = wall(lambda: partial.add_input(prog, cols, n))
    res["extract(num_groups)"], ng = wall(partial.num_groups)
    res["partition(8)"], (rows, counts) = wall(lambda: partial.partition(8))
    res["merge_rows"], _ = wall(lambda: final.merge_rows(rows))
    res["final.num_groups"], nf = wall(final.num_groups)
    res["final.reset"], _ = wall(final.reset)
print(json.dumps({k: round(v, 3) for k, v in res.items()}), ng, nf, counts[:3])
