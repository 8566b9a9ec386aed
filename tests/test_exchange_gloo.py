"""World-size-2 test of the host logic of the multi-GPU exchange on CPU over
gloo: per-rank partial groups -> hash partition by key -> counts all-gather -> variable all-to-all laid
out by engine.plan_alltoall (the product's planner of send/receive segments) -> final merge, compared
with a single-process answer.  The device kernels are replaced by the numpy oracle here; the same
protocol with libvgpu's partition / merge kernels over NCCL runs in tests/multi_gpu_check.py."""
import importlib
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
WORLD = 2
ORDERS = 15000  # lineitem SF0.01


def mix64(k):
    """splitmix64 finaliser, the hash of vgpu_agg_partition (include/vgpu.h)."""
    k = np.asarray(k, dtype=np.uint64)
    k = (k ^ (k >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)
    k = (k ^ (k >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)
    return k ^ (k >> np.uint64(31))


def partial_rows(data):
    """Packed group rows [key, flags, rows, sum bits, count, nn(sum), nn(count)] of one rank's slice."""
    import np_oracle as orc
    n = len(data["l_partkey"])
    no = np.zeros(n, dtype=bool)
    res = orc.aggregate([(data["l_partkey"], no)], [("sum_f64", (data["l_extendedprice"], no)), ("count", None)], np.ones(n, dtype=bool))
    rows = np.zeros((len(res), 7), dtype=np.int64)
    for i, (k, v) in enumerate(sorted(res.items())):
        rows[i] = [k[0], 0, v[1][0], np.float64(v[0][0]).view(np.int64), v[1][0], v[0][1], v[1][1]]
    return rows


def worker(rank, port, out_path):
    for p in (ROOT, os.path.join(ROOT, "oracle"), HERE):
        if p not in sys.path:
            sys.path.insert(0, p)
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=WORLD)
    engine = importlib.import_module("velox-1_b200.engine")
    import lineitem
    per = ORDERS // WORLD
    data = lineitem.gen_range(0.01, rank * per, per, ["l_partkey", "l_extendedprice"])
    rows = partial_rows(data)
    with np.errstate(over="ignore"):
        dest = (mix64(rows[:, 0].view(np.uint64) ^ (rows[:, 1].view(np.uint64) << np.uint64(63))) % np.uint64(WORLD)).astype(np.int64)
    order = np.argsort(dest, kind="stable")
    send = np.ascontiguousarray(rows[order])
    counts = np.bincount(dest, minlength=WORLD).tolist()
    gathered = [None] * WORLD
    dist.all_gather_object(gathered, counts)
    matrix = np.array(gathered)            # [src][dst]
    recv_counts = matrix[:, rank]
    row_bytes = rows.shape[1] * 8
    sc, so, rc, ro = engine.plan_alltoall(counts, recv_counts, row_bytes)
    assert sum(sc) == send.nbytes and sum(rc) == int(recv_counts.sum()) * row_bytes
    sbuf = torch.from_numpy(send.view(np.uint8).reshape(-1))
    rbuf = torch.zeros(max(1, sum(rc)), dtype=torch.uint8)
    reqs = []
    for peer in range(WORLD):
        if peer == rank:
            rbuf[ro[peer]:ro[peer] + rc[peer]] = sbuf[so[peer]:so[peer] + sc[peer]]
            continue
        if sc[peer]:
            reqs.append(dist.isend(sbuf[so[peer]:so[peer] + sc[peer]].clone(), peer))
        if rc[peer]:
            reqs.append(dist.irecv(rbuf[ro[peer]:ro[peer] + rc[peer]], peer))
    for r in reqs:
        r.wait()
    got = rbuf[:sum(rc)].numpy().view(np.int64).reshape(-1, rows.shape[1])
    final = {}
    for r in got:  # addIntermediateResults: sums and counts add
        assert int(mix64(np.uint64(r[0])) % np.uint64(WORLD)) == rank, "row routed to the wrong rank"
        s, c = final.get(int(r[0]), (0.0, 0))
        final[int(r[0])] = (s + float(np.int64(r[3]).view(np.float64)), c + int(r[4]))
    everyone = [None] * WORLD
    dist.all_gather_object(everyone, final)
    if rank == 0:
        merged = {}
        for part in everyone:
            assert not (merged.keys() & part.keys()), "a group ended up on two ranks"
            merged.update(part)
        np.save(out_path, np.array([[k, v[0], v[1]] for k, v in sorted(merged.items())], dtype=np.float64))
    dist.destroy_process_group()


def test_partition_exchange_merge_world2(tmp_path):
    sys.path.insert(0, HERE)
    import lineitem
    if not os.path.exists(lineitem.REF_SO):
        pytest.skip("oracle/_ref/libdbgen_ref.so not built")
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    out = str(tmp_path / "final.npy")
    mp.spawn(worker, args=(port, out), nprocs=WORLD, join=True)
    got = np.load(out)
    data = lineitem.gen_range(0.01, 0, ORDERS, ["l_partkey", "l_extendedprice"])
    keys, inv = np.unique(data["l_partkey"], return_inverse=True)
    want_sum = np.zeros(len(keys))
    np.add.at(want_sum, inv, data["l_extendedprice"])
    want_cnt = np.bincount(inv)
    assert np.array_equal(got[:, 0].astype(np.int64), keys)
    assert np.array_equal(got[:, 2].astype(np.int64), want_cnt)        # counts: exact
    assert np.allclose(got[:, 1], want_sum, rtol=1e-9, atol=0)         # FP sums: 1e-9 (north star)


def test_plan_alltoall_offsets():
    sys.path.insert(0, ROOT)
    engine = importlib.import_module("velox-1_b200.engine")
    sc, so, rc, ro = engine.plan_alltoall([3, 0, 5], [1, 2, 0], 56)
    assert sc == [168, 0, 280] and so == [0, 168, 168]
    assert rc == [56, 112, 0] and ro == [0, 56, 168]


def test_partition_of_matches_the_reference_hash():
    """One raw BIGINT key: partition = folly::hasher<int64_t>(key) % n = twang_mix64(key) % n, NULL -> kNullHash = 1
    (velox/exec/HashPartitionFunction.cpp:59-92, velox/exec/VectorHasher.cpp:68-108, velox/vector/BaseVector.h:74;
    twang_mix64 as restated in the reference tree at velox/experimental/gpu/tests/HashTableTest.cu:51-60)."""
    import ctypes as C
    import importlib
    import numpy as np
    abi = importlib.import_module("velox-1_b200").abi
    M = (1 << 64) - 1

    def twang(k):
        k &= M
        k = ((~k) + (k << 21)) & M
        k ^= k >> 24
        k = (k + (k << 3) + (k << 8)) & M
        k ^= k >> 14
        k = (k + (k << 2) + (k << 4)) & M
        k ^= k >> 28
        k = (k + (k << 31)) & M
        return k

    rng = np.random.default_rng(3)
    keys = [0, 1, -1, 2**63 - 1, -2**63, 6_000_000] + [int(x) for x in rng.integers(-2**62, 2**62, 200)]
    out = C.c_int32()
    for n in (1, 2, 3, 8, 31):
        for k in keys:
            abi.check(abi.lib().vgpu_partition_of(k, 0, n, 1, C.byref(out)))
            assert out.value == twang(k) % n, (k, n)
        abi.check(abi.lib().vgpu_partition_of(12345, 1, n, 1, C.byref(out)))
        assert out.value == 1 % n
    # the packed-key function is deterministic and spreads
    seen = set()
    for k in range(1000):
        abi.check(abi.lib().vgpu_partition_of(k, 0, 8, 0, C.byref(out)))
        seen.add(out.value)
    assert seen == set(range(8))
