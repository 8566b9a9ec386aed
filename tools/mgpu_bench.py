#!/usr/bin/env python
"""Config C5's group-by leg: lineitem SF10 per GPU, GROUP BY l_orderkey
sum(l_extendedprice), partial stage per GPU -> hash partition + peer stores over NVLink (or NCCL) -> final
aggregation per GPU.  Launch: python -m torch.distributed.run --nnodes=1 --nproc-per-node N
--master-addr 127.0.0.1 tools/mgpu_bench.py [--sf 10].  Prints one JSON line from rank 0: whole-job
rows/s (CUDA events, max over ranks), the time of the partial phase alone, exchanged bytes, the
NVLink rate they imply, and a result check (group count, sum of keys, sum of sums against torch reductions)."""
import argparse
import importlib
import json
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def groupby_leg(engine, abi, ex, exchange, dist, rank, world, keys_t, vals_t, steps, warmup, window=True):
    """Times the C5 group-by leg over this rank's (keys, values) and checks the result.  Returns a dict (rank 0) or None."""
    n = int(keys_t.numel())
    I64, D = abi.BIGINT, abi.DOUBLE
    cols = [engine.DeviceColumn(I64, abi.FLAT, keys_t, length=n), engine.DeviceColumn(D, abi.FLAT, vals_t, length=n)]
    empty = [engine.DeviceColumn(I64, abi.FLAT, keys_t[:0], length=0), engine.DeviceColumn(D, abi.FLAT, vals_t[:0], length=0)]
    prog = engine.Program(None, [ex.col(0, I64), ex.col(1, D)])
    uniq = torch.unique_consecutive(keys_t)
    groups = int(uniq.numel())
    accs = [(abi.ACC_SUM_F64, 1)]
    partial = engine.Aggregation(abi.AGG_HASH, [(0, 64, 0, 0)], accs, expected_groups=groups)
    final = engine.Aggregation(abi.AGG_HASH, [(0, 64, 0, 0)], accs, expected_groups=groups + groups // 8) if world > 1 else None
    mapped = False
    if world > 1 and window:
        # three parts x world regions; a region must hold this rank's share of any sender's entries (hash-uniform)
        per_rank = torch.tensor([n], dtype=torch.int64, device="cuda")
        dist.all_reduce(per_rank, op=dist.ReduceOp.MAX)
        mapped = exchange.reserve_window(int(3 * 1.3 * int(per_rank[0]) * 16) + (64 << 20))

    def sync():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize()

    def step(with_exchange):
        partial.reset()
        partial.add_input(prog, cols, n)
        if with_exchange and world > 1:
            final.reset()
            final.add_input(prog, empty, 0)
            exchange.shuffle(partial, final)   # leaves `final` finalized (asynchronously)
        else:
            partial.finalize()

    def timed(with_exchange):
        for _ in range(warmup):
            step(with_exchange)
        sync()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(steps):
            step(with_exchange)
        if with_exchange and world > 1:
            final.finalize()  # the last exchange's tail runs on the communicator's side stream: the timed stream waits for it
        e1.record()
        sync()
        t = torch.tensor([e0.elapsed_time(e1) / steps], dtype=torch.float64, device="cuda")
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t[0])

    ms_partial = timed(False)
    ms_classic = timed(True) if world > 1 else ms_partial     # scan, then partition + peer stores, then final
    fused = bool(world > 1 and mapped and world <= 8 and os.environ.get("VGPU_EXCHANGE") != "nccl"
                 and os.environ.get("VGPU_FUSED_EXCHANGE", "1") != "0")
    if fused:
        exchange.attach(partial)                               # the scan itself stores the entries into the owners' windows
    ms_full = timed(True) if fused else ms_classic
    scan_ms = None
    if os.environ.get("VGPU_BENCH_PHASES"):  # diagnostic: the scan kernel's own duration inside the full step (events around it)
        import ctypes as C
        lib = abi.lib()
        lib.vgpu_debug_kernel_timing(1)
        for _ in range(5):
            step(True)
        sync()
        tot, cnt = C.c_double(), C.c_int64()
        lib.vgpu_debug_kernel_time(C.byref(tot), C.byref(cnt))
        lib.vgpu_debug_kernel_timing(0)
        t = torch.tensor([tot.value / max(cnt.value, 1)], dtype=torch.float64, device="cuda")
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        scan_ms = float(t[0])
    engine.check_errors()
    result = final if world > 1 else partial
    rows = torch.from_numpy(result.extract()).cuda()          # [groups here, 5]: key, flags, rows, sum, nn
    mine = torch.zeros(4, dtype=torch.float64, device="cuda")
    mine[0] = rows.shape[0]
    mine[1] = rows[:, 0].double().sum()                       # sum of keys (exact below 2^53: checked as float64)
    mine[2] = rows[:, 3].contiguous().view(torch.float64).sum()
    want = torch.tensor([groups, float(uniq.double().sum()), float(vals_t.sum())], dtype=torch.float64, device="cuda")
    sent = 0
    if world > 1:
        if mapped and os.environ.get("VGPU_EXCHANGE") != "nccl":
            m = exchange.last_counts()
            sent = int(m[rank].sum() - m[rank][rank]) * 16
        else:
            sent = max(exchange.bytes_sent, 0)
    mine[3] = sent
    if world > 1:
        dist.all_reduce(mine)
        dist.all_reduce(want)
    if rank != 0:
        return None
    ok = (int(mine[0]) == int(want[0]) and abs(float(mine[1]) - float(want[1])) <= 1e-12 * abs(float(want[1]))
          and abs(float(mine[2]) - float(want[2])) <= 1e-9 * abs(float(want[2])))
    ex_ms = max(ms_full - ms_partial, 1e-6)
    info = partial.info()
    return {
        "case": "C5 group-by leg: GROUP BY l_orderkey sum(l_extendedprice), partial -> exchange -> final", "n_gpus": world,
        "rows_per_gpu": n, "groups_total": int(mine[0]), "ms_per_step": ms_full, "ms_partial_phase": ms_partial,
        "ms_exchange_and_final": ex_ms if world > 1 else 0.0, "rows_per_s": n * world / (ms_full * 1e-3),
        "exchanged_bytes_per_step": int(mine[3]),
        "ms_per_step_unfused": ms_classic,
        "ms_scan_kernel_in_step": scan_ms,
        "nvlink_gbs_per_gpu_out": (float(mine[3]) / world / ((ms_full if fused else ex_ms) * 1e-3) / 1e9) if world > 1 else 0.0,
        "exchange_path": ("window, fused into the scan: the LOG-sink kernel stores each entry into its owner's window over NVLink while it scans; counts and stream facts in one ncclAllGather" if fused
                          else "window: partition + NVLink peer stores in one kernel, counts in one ncclAllGather" if mapped and os.environ.get("VGPU_EXCHANGE") != "nccl"
                          else ("nccl: partition, counts all-gather, grouped ncclSend/ncclRecv, merge" if world > 1 else "none")),
        "partial_engine": info,
        "result_check": (f"ok: {int(mine[0])} groups, sum of keys and sum of sums match torch reductions over all ranks" if ok
                         else f"FAILED: got {mine.tolist()} want {want.tolist()}"),
        "note": "partial phase = reset + scan (run aggregation + entry log) + its own deduplication; full step = reset + scan + exchange + final deduplication; ms_per_step_unfused = the same step with scan and exchange as separate kernels; nvlink_gbs is over the whole step when fused (the transfer overlaps the scan)",
    }


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--sf", type=float, default=10.0)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    args = ap.parse_args()
    rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    pkg = importlib.import_module("velox-1_b200")
    abi, ex = pkg.abi, pkg.expr
    abi.check(abi.lib().vgpu_init(local))
    engine = importlib.import_module("velox-1_b200.engine")
    tpch = importlib.import_module("velox-1_b200.tpch_synth")
    exchange = None
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
        ids = [engine.Exchange.make_unique_id() if rank == 0 else None]
        dist.broadcast_object_list(ids, src=0)
        exchange = engine.Exchange(rank, world, ids[0])
    n = tpch.row_count(args.sf)
    orders = n // 4 + n // 50 + 16
    data = tpch.lineitem(args.sf, ["l_orderkey", "l_extendedprice"], order_offset=rank * ((orders + 7) // 8 * 8), seed=rank)
    out = groupby_leg(engine, abi, ex, exchange, dist, rank, world, data["l_orderkey"], data["l_extendedprice"], args.steps, args.warmup)
    if rank == 0:
        out["sf_per_gpu"] = args.sf
        print(json.dumps(out), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
