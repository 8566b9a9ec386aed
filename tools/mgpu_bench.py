#!/usr/bin/env python
"""Config C5's group-by leg: lineitem SF10 per GPU, GROUP BY l_orderkey
sum(l_extendedprice), partial aggregation per GPU -> hash partition -> NCCL all-to-all -> final
aggregation per GPU.  Launch: python -m torch.distributed.run --nnodes=1 --nproc-per-node N
--master-addr 127.0.0.1 tools/mgpu_bench.py [--sf 10].  Prints one JSON line from rank 0: whole-job
rows/s (CUDA events, max over ranks), the time of the partial phase alone, exchanged bytes and the
NVLink rate they imply."""
import argparse
import importlib
import json
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--sf", type=float, default=10.0)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    args = ap.parse_args()
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    pkg = importlib.import_module("velox-1_b200")
    abi, ex = pkg.abi, pkg.expr
    abi.check(abi.lib().vgpu_init(local))
    engine = importlib.import_module("velox-1_b200.engine")
    tpch = importlib.import_module("velox-1_b200.tpch_synth")
    ids = [engine.Exchange.make_unique_id() if rank == 0 else None]
    dist.broadcast_object_list(ids, src=0)
    exchange = engine.Exchange(rank, world, ids[0])
    n = tpch.row_count(args.sf)
    orders = n // 4 + n // 50 + 16
    data = tpch.lineitem(args.sf, ["l_orderkey", "l_extendedprice"], order_offset=rank * ((orders + 7) // 8 * 8), seed=rank)
    I64, D = abi.BIGINT, abi.DOUBLE
    cols = [engine.DeviceColumn(I64, abi.FLAT, data["l_orderkey"], length=n), engine.DeviceColumn(D, abi.FLAT, data["l_extendedprice"], length=n)]
    empty = [engine.DeviceColumn(I64, abi.FLAT, data["l_orderkey"][:0], length=0), engine.DeviceColumn(D, abi.FLAT, data["l_extendedprice"][:0], length=0)]
    prog = engine.Program(None, [ex.col(0, I64), ex.col(1, D)])
    groups = int(torch.unique_consecutive(data["l_orderkey"]).numel())
    accs = [(abi.ACC_SUM_F64, 1)]
    partial = engine.Aggregation(abi.AGG_HASH, [(0, 64, 0, 0)], accs, expected_groups=groups)
    final = engine.Aggregation(abi.AGG_HASH, [(0, 64, 0, 0)], accs, expected_groups=groups + groups // 8)

    def sync():
        torch.cuda.synchronize()
        dist.barrier()
        torch.cuda.synchronize()

    def step(with_exchange):
        partial.reset()
        partial.add_input(prog, cols, n)
        if with_exchange:
            final.reset()
            final.add_input(prog, empty, 0)
            exchange.shuffle(partial, final)

    def timed(with_exchange):
        for _ in range(args.warmup):
            step(with_exchange)
        sync()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(args.steps):
            step(with_exchange)
        e1.record()
        sync()
        t = torch.tensor([e0.elapsed_time(e1) / args.steps], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t[0])

    ms_partial = timed(False)
    ms_full = timed(True)
    engine.check_errors()
    mine = torch.tensor([final.num_groups(), exchange.bytes_sent], dtype=torch.int64, device="cuda")
    dist.all_reduce(mine)
    total_groups, total_sent = int(mine[0]), int(mine[1])
    if rank == 0:
        ex_ms = max(ms_full - ms_partial, 1e-6)
        print(json.dumps({
            "case": "C5 group-by leg: GROUP BY l_orderkey sum(l_extendedprice), partial -> all-to-all -> final", "n_gpus": world,
            "rows_per_gpu": n, "sf_per_gpu": args.sf, "groups_total": total_groups, "expected_groups_total": groups * world if world > 0 else groups,
            "ms_per_step": ms_full, "ms_partial_phase": ms_partial, "ms_exchange_and_final": ex_ms,
            "rows_per_s": n * world / (ms_full * 1e-3), "exchanged_bytes_per_step": total_sent,
            "nvlink_gbs_per_gpu_out": total_sent / world / (ex_ms * 1e-3) / 1e9,
            "note": "reset of both tables (init kernels) is inside the step; exchange = partition + counts all-gather + grouped ncclSend/ncclRecv + merge"}), flush=True)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
