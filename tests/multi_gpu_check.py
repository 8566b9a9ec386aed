"""Parity of the multi-GPU path (SURVEY.md section 8(e), config C5's shape at test size), run as
  python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 tests/multi_gpu_check.py
One process per GPU: scan + partial aggregation of this rank's order range, vgpu_agg_partition,
NCCL all-to-all of the packed partial groups (vgpu_comm_alltoallv), vgpu_agg_merge_rows into the final
aggregation; rank 0 compares the union of all ranks' final groups with the numpy oracle over the whole
table.  Q1 (ARRAY sink, groups present on every rank) and GROUP BY l_partkey (HASH sink, groups spread
over ranks) and GROUP BY l_orderkey (HASH, disjoint per rank)."""
import importlib
import os
import pickle
import sys

import numpy as np
import torch
import torch.distributed as dist

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
for p in (ROOT, os.path.join(ROOT, "oracle"), HERE):
    if p not in sys.path:
        sys.path.insert(0, p)


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    pkg = importlib.import_module("velox-1_b200")
    abi, ex = pkg.abi, pkg.expr
    abi.check(abi.lib().vgpu_init(local))
    engine = importlib.import_module("velox-1_b200.engine")
    import helpers as H
    import lineitem
    ids = [engine.Exchange.make_unique_id() if rank == 0 else None]
    dist.broadcast_object_list(ids, src=0)
    exchange = engine.Exchange(rank, world, ids[0])
    mapped = exchange.reserve_window(512 << 20)   # peer-mapped window: the fused partition + NVLink store path
    if os.environ.get("VGPU_EXCHANGE") != "nccl":
        assert mapped, "CUDA IPC window not mapped: the P2P exchange path was not exercised"

    sf = 0.1
    orders = 150000
    per = (orders + world - 1) // world
    mine = lineitem.gen_range(sf, rank * per, min(per, orders - rank * per))
    n = len(mine["l_orderkey"])
    scratch = f"/tmp/vgpu_mgpu_{os.environ.get('MASTER_PORT', '0')}"
    os.makedirs(scratch, exist_ok=True)
    results = {}

    streaming = mapped and world <= 8 and os.environ.get("VGPU_FUSED_EXCHANGE", "1") != "0"

    def run(name, mode, filt, outs, col_names, types, keys, accs, fused=False):
        cols = [engine.flat_column(t, mine[c]) for c, t in zip(col_names, types)]
        prog = engine.Program(filt, outs)
        empty = [engine.flat_column(t, mine[c][:0]) for c, t in zip(col_names, types)]
        partial = engine.Aggregation(mode, keys, accs)
        final = engine.Aggregation(mode, keys, accs)
        if fused:
            exchange.attach(partial)                           # the scan stores its entries into the owners' windows
            half = (n // 2 + 63) // 64 * 64                    # two batches
            partial.add_input(prog, [engine.flat_column(t, mine[c][:half]) for c, t in zip(col_names, types)], half)
            partial.add_input(prog, [engine.flat_column(t, mine[c][half:]) for c, t in zip(col_names, types)], n - half)
        else:
            partial.add_input(prog, cols, n)
        final.add_input(prog, empty, 0)
        exchange.shuffle(partial, final)                       # vgpu_comm_exchange_groups
        results[name] = H.decode_rows(final.extract(), keys, accs, [o.type for o in outs], mode)
        if fused:
            # two-word entries (one accumulator) must have streamed; wider entries may exceed the streaming kernel's
            # shared-memory buffers at 8 ranks and then take the partition-after-the-scan path (same result)
            partial.finalize()
            if len(accs) == 1:
                assert partial.num_groups() == 0, "fused exchange: entries stayed behind in the partial aggregation"
            return
        if mode == abi.AGG_GLOBAL:
            return  # vgpu_agg_merge_rows has no notion of the owning rank of the single global row
        again = engine.Aggregation(mode, keys, accs)
        again.add_input(prog, empty, 0)
        exchange.shuffle_stepwise(partial, again)              # partition / all-to-all / merge, step by step
        stepwise = H.decode_rows(again.extract(), keys, accs, [o.type for o in outs], mode)
        assert stepwise.keys() == results[name].keys(), f"{name}: the two exchange paths own different groups"
        H.assert_groups_equal(stepwise, results[name], accs, rtol=1e-12)

    f, outs = H.q1_exprs()
    I32, I64, D = abi.INTEGER, abi.BIGINT, abi.DOUBLE
    for rep in range(3):  # repeated exchanges alternate the window halves
        run("q1", abi.AGG_ARRAY, f, outs, H.Q1_COLS, [I32, I32, D, D, D, D, I32], H.Q1_KEYS, H.Q1_ACCS)
    f6, outs6 = H.q6_exprs()
    q6_accs = [(abi.ACC_SUM_F64, 0), (abi.ACC_COUNT, -1)]
    run("q6", abi.AGG_GLOBAL, f6, outs6, H.Q6_COLS, [I32, D, D, D], [], q6_accs)
    gb_accs = [(abi.ACC_SUM_F64, 1), (abi.ACC_COUNT, -1), (abi.ACC_MIN, 1), (abi.ACC_MAX, 1)]
    for key in ("l_partkey", "l_orderkey", "l_orderkey"):
        run(key, abi.AGG_HASH, None, [ex.col(0, I64), ex.col(1, D)], [key, "l_extendedprice"], [I64, D], [(0, 64, 0, 0)], gb_accs)
    if streaming:
        for rep in range(2):
            run("l_orderkey fused", abi.AGG_HASH, None, [ex.col(0, I64), ex.col(1, D)], ["l_orderkey", "l_extendedprice"], [I64, D], [(0, 64, 0, 0)], gb_accs[:1], fused=True)
        run("l_partkey fused", abi.AGG_HASH, None, [ex.col(0, I64), ex.col(1, D)], ["l_partkey", "l_extendedprice"], [I64, D], [(0, 64, 0, 0)], gb_accs, fused=True)
    # NULL keys and the key equal to the tables' empty marker travel in the regions' dedicated slots
    mine["k_special"] = np.where(mine["l_orderkey"] % 7 == 0, -1, mine["l_orderkey"] % 1000).astype(np.int64)
    special_nn = (mine["l_orderkey"] % 5) != 0

    def run_special(fused=False):
        cols = [engine.flat_column(I64, mine["k_special"], special_nn), engine.flat_column(D, mine["l_extendedprice"])]
        outs_s = [ex.col(0, I64), ex.col(1, D)]
        prog = engine.Program(None, outs_s)
        empty = [engine.flat_column(I64, mine["k_special"][:0]), engine.flat_column(D, mine["l_extendedprice"][:0])]
        partial = engine.Aggregation(abi.AGG_HASH, [(0, 64, 0, 0)], gb_accs)
        final = engine.Aggregation(abi.AGG_HASH, [(0, 64, 0, 0)], gb_accs)
        if fused:
            exchange.attach(partial)
        partial.add_input(prog, cols, n)
        final.add_input(prog, empty, 0)
        exchange.shuffle(partial, final)
        results["special fused" if fused else "special"] = H.decode_rows(final.extract(), [(0, 64, 0, 0)], gb_accs, [I64, D], abi.AGG_HASH)
    run_special()
    if streaming:
        run_special(fused=True)
    with open(os.path.join(scratch, f"rank{rank}.pkl"), "wb") as fh:
        pickle.dump(results, fh)
    dist.barrier()
    ok = True
    if rank == 0:
        whole = lineitem.gen_range(sf, 0, orders)
        total = len(whole["l_orderkey"])
        parts = [pickle.load(open(os.path.join(scratch, f"rank{r}.pkl"), "rb")) for r in range(world)]
        for name in results:
            union = {}
            for p in parts:
                assert not (union.keys() & p[name].keys()), f"{name}: a group is final on two ranks"
                union.update(p[name])
            if name == "q1":
                want = H.run_oracle_agg(f, outs, [(whole[c], None) for c in H.Q1_COLS], total, H.Q1_KEYS, H.Q1_ACCS)
                H.assert_groups_equal(union, want, H.Q1_ACCS)
            elif name == "q6":
                want = H.run_oracle_agg(f6, outs6, [(whole[c], None) for c in H.Q6_COLS], total, [], q6_accs)
                assert sum(len(p[name]) for p in parts) == 1, "a global aggregation reports one row over all ranks"
                H.assert_groups_equal(union, want, q6_accs)
            elif name.startswith("special"):
                ks = np.where(whole["l_orderkey"] % 7 == 0, -1, whole["l_orderkey"] % 1000).astype(np.int64)
                knn = (whole["l_orderkey"] % 5) != 0
                want = H.run_oracle_agg(None, [ex.col(0, I64), ex.col(1, D)], [(ks, knn), (whole["l_extendedprice"], None)], total,
                                        [(0, 64, 0, 0)], gb_accs)
                assert (None,) in union and (-1,) in union
                H.assert_groups_equal(union, want, gb_accs)
            else:
                accs_n = gb_accs[:1] if name == "l_orderkey fused" else gb_accs
                want = H.run_oracle_agg(None, [ex.col(0, I64), ex.col(1, D)], [(whole[name.split(" ")[0]], None), (whole["l_extendedprice"], None)], total,
                                        [(0, 64, 0, 0)], accs_n)
                H.assert_groups_equal(union, want, accs_n)
            spread = [len(p[name]) for p in parts]
            print(f"multi-gpu parity ok: {name}: {len(union)} groups over {total} rows, final groups per rank {spread}", flush=True)
    dist.barrier()
    dist.destroy_process_group()
    return 0 if ok else 1


if __name__ == "__main__":
    sys.exit(main())
