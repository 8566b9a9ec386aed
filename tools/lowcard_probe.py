#!/usr/bin/env python
"""HASH aggregation over unsorted keys of low cardinality (the case the run folding does not help): time per aggregation."""
import importlib, json, os, sys
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
pkg = importlib.import_module("velox-1_b200"); abi, ex = pkg.abi, pkg.expr
abi.check(abi.lib().vgpu_init(0))
engine = importlib.import_module("velox-1_b200.engine")
n = int(sys.argv[1]) if len(sys.argv) > 1 else 59_986_052
I64, D = abi.BIGINT, abi.DOUBLE
g = torch.Generator(device="cuda").manual_seed(1)
vals = torch.rand(n, device="cuda", dtype=torch.float64, generator=g)
prog = engine.Program(None, [ex.col(0, I64), ex.col(1, D)])
for card in (4, 1000, 100_000):
    keys = torch.randint(0, card, (n,), device="cuda", dtype=torch.int64, generator=g)
    cols = [engine.DeviceColumn(I64, abi.FLAT, keys, length=n), engine.DeviceColumn(D, abi.FLAT, vals, length=n)]
    agg = engine.Aggregation(abi.AGG_HASH, [(0, 64, 0, 0)], [(abi.ACC_SUM_F64, 1)])
    ts = []
    for it in range(4):
        agg.reset(); torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); agg.add_input(prog, cols, n); agg.finalize(); e1.record(); torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    ng = agg.num_groups()
    rows = torch.from_numpy(agg.extract()).cuda()
    ok = ng == card and abs(float(rows[:, 3].contiguous().view(torch.float64).sum()) - float(vals.sum())) <= 1e-9 * float(vals.sum())
    print(json.dumps({"cardinality": card, "rows": n, "ms": round(min(ts[1:]), 3), "groups": ng, "check": "ok" if ok else "FAILED"}))
