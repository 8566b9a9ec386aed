#!/usr/bin/env python
"""Kernel-level timings of the four single-GPU configurations with
device-resident input: CUDA events around vgpu_agg_add_input on the launching stream, median of
`--iters` runs after warm-up.  Prints one JSON line per case with algorithmic GB/s and the fraction
of the measured HBM peak.  Used for tuning and as the ncu target (`--only q1 --iters 1`).
"""
import argparse
import importlib
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "tests"), os.path.join(ROOT, "oracle")):
    if p not in sys.path:
        sys.path.insert(0, p)

import numpy as np  # noqa: E402
import torch  # noqa: E402


def peak():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        return float(json.load(open(path))["hbm_gbs"])
    return 6650.0


def bits_every(n, k, device):
    """int64 words of a not-null bitmap where row i is NULL iff i % k == 0."""
    i = torch.arange(((n + 63) // 64) * 64, device=device)
    nn = ((i % k) != 0) & (i < n)
    w = nn.view(-1, 64).to(torch.int64)
    sh = torch.arange(64, device=device, dtype=torch.int64)
    return (w << sh).sum(dim=1)  # wraps mod 2^64: bit 63 lands in the sign bit


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--sf", type=float, default=10.0)
    ap.add_argument("--iters", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--only", default="q1,q6,q6fp,orderkey,randkey,c4")
    ap.add_argument("--engines", default="log,table", help="HASH engines to time for orderkey / randkey")
    args = ap.parse_args()
    pkg = importlib.import_module("velox-1_b200")
    abi = pkg.abi
    abi.check(abi.lib().vgpu_init(0))
    engine = importlib.import_module("velox-1_b200.engine")
    tpch = importlib.import_module("velox-1_b200.tpch_synth")
    import helpers as H
    ex = pkg.expr
    I32, I64, D = abi.INTEGER, abi.BIGINT, abi.DOUBLE
    n = tpch.row_count(args.sf)
    want = args.only.split(",")
    names = ["l_orderkey", "l_returnflag", "l_linestatus", "l_quantity", "l_extendedprice", "l_discount", "l_tax",
             "l_shipdate", "l_shipmode", "l_shipinstruct"]
    data = tpch.lineitem(args.sf, names)
    typ = {"l_orderkey": I64, "l_returnflag": I32, "l_linestatus": I32, "l_quantity": D, "l_extendedprice": D,
           "l_discount": D, "l_tax": D, "l_shipdate": I32, "l_shipmode": I32, "l_shipinstruct": I32}

    def flat(c, nulls=None):
        return engine.DeviceColumn(typ[c], abi.FLAT, data[c], nulls=nulls, length=n)

    def timed(fn, reset):
        """GPU time of fn() per iteration: iterations are queued back to back (no host sync in
        between), so the events bracket device work only, not Python / planning overhead."""
        evs = []
        for i in range(args.warmup + args.iters):
            reset()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            fn()
            e1.record()
            if i < args.warmup:
                torch.cuda.synchronize()
            else:
                evs.append((e0, e1))
        torch.cuda.synchronize()
        ms = [a.elapsed_time(b) for a, b in evs]
        engine.check_errors()
        return float(np.median(ms)), float(np.min(ms))

    def report(name, bytes_per_row, med, best, extra=None):
        gbs = n * bytes_per_row / (med * 1e-3) / 1e9
        line = {"case": name, "rows": n, "bytes_per_row": bytes_per_row, "ms_median": round(med, 4),
                "ms_best": round(best, 4), "rows_per_s": n / (med * 1e-3), "alg_gbs": round(gbs, 1),
                "frac_of_measured_peak": round(gbs / peak(), 4)}
        line.update(extra or {})
        print(json.dumps(line), flush=True)

    if "q1" in want:
        f, outs = H.q1_exprs()
        cols = [flat(c) for c in H.Q1_COLS]
        prog = engine.Program(f, outs)
        agg = engine.Aggregation(abi.AGG_ARRAY, H.Q1_KEYS, H.Q1_ACCS)
        med, best = timed(lambda: agg.add_input(prog, cols, n), agg.reset)
        report("C2 Q1 (ARRAY sink, 4 groups)", 44, med, best, {"groups": agg.num_groups()})
    if "q6" in want:
        f, outs = H.q6_exprs()
        cols = [flat(c) for c in H.Q6_COLS]
        prog = engine.Program(f, outs)
        accs = [(abi.ACC_SUM_F64, 0), (abi.ACC_COUNT, -1)]
        agg = engine.Aggregation(abi.AGG_GLOBAL, [], accs)
        med, best = timed(lambda: agg.add_input(prog, cols, n), agg.reset)
        report("C1 Q6 (GLOBAL sink)", 28, med, best)
    if "q6fp" in want:
        f, outs = H.q6_exprs()
        cols = [flat(c) for c in H.Q6_COLS]
        prog = engine.Program(f, outs)
        med, best = timed(lambda: engine.filter_project(prog, cols, n, want_nulls=False, sync=False), lambda: None)
        # reads 28 B/row, writes the projection (8 B/row), the bitmap and the selected row numbers
        report("C1 Q6 FilterProject (selection + projection)", 28 + 8 + 0.125, med, best)
    if "orderkey" in want or "randkey" in want:
        outs = [ex.col(0, I64), ex.col(1, D)]
        prog = engine.Program(None, outs)
        accs = [(abi.ACC_SUM_F64, 1)]
        groups = int(torch.unique_consecutive(data["l_orderkey"]).numel())
        cases = []
        if "orderkey" in want:
            cases.append(("C3 GROUP BY l_orderkey (sorted runs of 1-7)", data["l_orderkey"]))
        if "randkey" in want:  # the same keys in random order: no runs to fold, nothing sorted
            perm = torch.randperm(n, device="cuda", generator=torch.Generator(device="cuda").manual_seed(5))
            cases.append(("high-cardinality GROUP BY, same keys shuffled", data["l_orderkey"][perm].contiguous()))
            del perm
        for label, keycol in cases:
            cols = [engine.DeviceColumn(I64, abi.FLAT, keycol, length=n), flat("l_extendedprice")]
            for eng in args.engines.split(","):
                os.environ["VGPU_HASH_ENGINE"] = eng
                agg = engine.Aggregation(abi.AGG_HASH, [(0, 64, 0, 0)], accs, expected_groups=groups)

                def step():
                    agg.add_input(prog, cols, n)
                    agg.finalize()

                # the table engine's reset wipes the whole table (init_table_kernel): time it inside the step as well
                med_in, best_in = timed(lambda: (agg.reset(), step()), lambda: None)
                med, best = timed(step, agg.reset)
                med_scan, _ = timed(lambda: agg.add_input(prog, cols, n), agg.reset)  # the scan alone (no finalize)
                agg.reset(); step()
                ng = agg.num_groups()
                info = agg.info()
                gbs = (n * 16 + ng * 16) / (med_in * 1e-3) / 1e9
                report(f"{label} [{eng} engine]", 16, med_in, best_in,
                       {"groups": ng, "ms_without_reset": round(med, 4), "ms_scan_only": round(med_scan, 4), "alg_gbs_incl_output": round(gbs, 1),
                        "frac_incl_output": round(gbs / peak(), 4), "engine_info": info,
                        "note": "ms_median includes reset (table init for the table engine) + scan + finalize"})
                del agg
            os.environ.pop("VGPU_HASH_ENGINE", None)
    if "c4" in want:
        mode_base = torch.tensor([6, 5, 4, 3, 2, 1, 0, 3], dtype=torch.int32, device="cuda")
        mode_idx = (6 - data["l_shipmode"]).to(torch.int32)
        instr_base = torch.tensor([2, 0, 3, 1], dtype=torch.int32, device="cuda")
        instr_idx = torch.tensor([1, 3, 0, 2], dtype=torch.int32, device="cuda")[data["l_shipinstruct"].long()]
        cols = [
            engine.DeviceColumn(I32, abi.DICTIONARY, mode_base, mode_idx, nulls=bits_every(n, 11, "cuda"), length=n),
            engine.DeviceColumn(I32, abi.DICTIONARY, instr_base, instr_idx, nulls=bits_every(n, 13, "cuda"), length=n),
            flat("l_extendedprice"), flat("l_shipdate"), flat("l_quantity", nulls=bits_every(n, 7, "cuda")),
        ]
        outs = [ex.col(0, I32), ex.col(1, I32), ex.col(2, D), ex.col(3, I32), ex.col(4, D)]
        keys = [(0, 0, 0, 6), (1, 0, 0, 3)]
        accs = [(abi.ACC_MIN, 2), (abi.ACC_MAX, 2), (abi.ACC_MIN, 3), (abi.ACC_MAX, 3), (abi.ACC_COUNT, 4),
                (abi.ACC_COUNT, -1)]
        prog = engine.Program(None, outs)
        agg = engine.Aggregation(abi.AGG_ARRAY, keys, accs)
        med, best = timed(lambda: agg.add_input(prog, cols, n), agg.reset)
        # bytes actually read: 2 x int32 indices + price 8 + date 4 + three null bitmaps; count(l_quantity) needs
        # only l_quantity's null bits, so its 8-byte values are never loaded (ncu: 1.22 GB = 20.4 B/row)
        report("C4 dictionary keys with nulls (ARRAY sink, 40 groups)", 20.375, med, best,
               {"groups": agg.num_groups(), "bytes_per_row_if_quantity_values_were_read": 28.375})


if __name__ == "__main__":
    main()
