#!/usr/bin/env python
"""One GPU: duration of the C5 scan with the exchange fused in, all `world` owners' windows being local HBM
(vgpu_debug_local_comms) — the kernel's own cost without the NVLink.  Usage: python tools/fused_scan_probe.py [world] [iters]"""
import importlib, json, os, sys
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
pkg = importlib.import_module("velox-1_b200"); abi, ex = pkg.abi, pkg.expr
abi.check(abi.lib().vgpu_init(0))
engine = importlib.import_module("velox-1_b200.engine")
world = int(sys.argv[1]) if len(sys.argv) > 1 else 8
iters = int(sys.argv[2]) if len(sys.argv) > 2 else 10
data = engine.tpch_lineitem(10.0, ["l_orderkey", "l_extendedprice"])
n = len(data["l_orderkey"])
I64, D = abi.BIGINT, abi.DOUBLE
cols = [engine.DeviceColumn(I64, abi.FLAT, data["l_orderkey"], length=n), engine.DeviceColumn(D, abi.FLAT, data["l_extendedprice"], length=n)]
prog = engine.Program(None, [ex.col(0, I64), ex.col(1, D)])
accs = [(abi.ACC_SUM_F64, 1)]
res = {"world": world, "rows": n}
for fused in (False, True):
    agg = engine.Aggregation(abi.AGG_HASH, [(0, 64, 0, 0)], accs)
    group = None
    if fused:
        group = engine.LocalExchangeGroup(world, int(3 * 1.3 * n * 16) + (64 << 20))
        group.attach(0, agg)
    ts = []
    for it in range(iters + 3):
        agg.reset()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); agg.add_input(prog, cols, n); e1.record()
        torch.cuda.synchronize()
        if it >= 3:
            ts.append(e0.elapsed_time(e1))
    engine.check_errors()
    res["ms_scan_fused" if fused else "ms_scan_local"] = round(sorted(ts)[len(ts) // 2], 4)
    del agg
    if group:
        group.close()
print(json.dumps(res))
