"""Device-side helpers over the C-ABI: columns in HBM (torch tensors as the allocator), programs,
FilterProject and aggregation calls.  Mirrors what the C++ operators in host/ do with RowVectors.
"""
import ctypes as C
from dataclasses import dataclass
from typing import List, Optional, Sequence

import numpy as np
import torch

from . import abi, expr as ex

NP_OF = {abi.TINYINT: np.int8, abi.SMALLINT: np.int16, abi.INTEGER: np.int32, abi.BIGINT: np.int64,
         abi.REAL: np.float32, abi.DOUBLE: np.float64}


def pack_bits(mask: np.ndarray) -> np.ndarray:
    """bool[n] -> uint64 words, bit i of word i//64 = mask[i] (velox/common/base/Nulls.h layout)."""
    n = len(mask)
    words = (n + 63) // 64
    b = np.zeros(words * 64, dtype=np.uint8)
    b[:n] = mask
    return np.packbits(b, bitorder="little").view(np.uint64)


def unpack_bits(words: np.ndarray, n: int) -> np.ndarray:
    return np.unpackbits(words.view(np.uint8), bitorder="little")[:n].astype(bool)


def stream_ptr() -> int:
    return int(torch.cuda.current_stream().cuda_stream)


@dataclass
class DeviceColumn:
    """One column resident in HBM, described the way DecodedVector exposes it."""
    type: int
    encoding: int = abi.FLAT
    values: Optional[torch.Tensor] = None      # flat values / dictionary base values (uint8 view for BOOLEAN bits)
    indices: Optional[torch.Tensor] = None     # int32
    nulls: Optional[torch.Tensor] = None       # int64 words, bit=1 -> not null
    base_nulls: Optional[torch.Tensor] = None
    constant: object = None
    constant_is_null: bool = False
    length: int = 0

    def desc(self) -> abi.Column:
        c = abi.Column()
        c.type, c.encoding = self.type, self.encoding
        c.values = self.values.data_ptr() if self.values is not None else None
        c.indices = self.indices.data_ptr() if self.indices is not None else None
        c.nulls = self.nulls.data_ptr() if self.nulls is not None else None
        c.base_nulls = self.base_nulls.data_ptr() if self.base_nulls is not None else None
        c.base_length = len(self.values) if (self.encoding == abi.DICTIONARY and self.values is not None) else 0
        if self.encoding == abi.CONSTANT and not self.constant_is_null:
            if self.type in (abi.REAL, abi.DOUBLE):
                c.constant.d = float(self.constant)
            else:
                c.constant.i = int(self.constant)
        c.constant_is_null = 1 if self.constant_is_null else 0
        return c

    def nbytes(self) -> int:
        return sum(t.numel() * t.element_size() for t in (self.values, self.indices, self.nulls, self.base_nulls)
                   if t is not None)


def _dev(a: np.ndarray, device) -> torch.Tensor:
    return torch.from_numpy(np.ascontiguousarray(a)).to(device)


def _bits_dev(mask: Optional[np.ndarray], device):
    return None if mask is None else _dev(pack_bits(np.asarray(mask, dtype=bool)).view(np.int64), device)


def string_views(strings: Sequence[bytes]) -> np.ndarray:
    """The 16-byte StringViews of `strings` (velox/type/StringView.h: u32 size, 4 prefix bytes, then the next 8 bytes
    when the string has at most 12, else a pointer — left 0 here: the device never follows it) as an (n, 2) int64
    array: what a FlatVector<StringView>'s values buffer holds, and what a VGPU_VARCHAR column points at."""
    raw = np.zeros((len(strings), 16), dtype=np.uint8)
    for i, s in enumerate(strings):
        raw[i, 0:4] = np.frombuffer(np.uint32(len(s)).tobytes(), dtype=np.uint8)
        head = s[:12] if len(s) <= 12 else s[:4]
        raw[i, 4:4 + len(head)] = np.frombuffer(head, dtype=np.uint8)
    return raw.view(np.int64).reshape(len(strings), 2)


def number_of_string(s: bytes) -> int:
    """include/vgpu.h vgpu_number_of_string (velox/exec/VectorHasher.h:324-329 stringAsNumber)."""
    assert len(s) <= 7
    return 0 if not s else int.from_bytes(s, "little") + (1 << (8 * len(s)))


def string_of_number(number: int) -> bytes:
    """include/vgpu.h vgpu_string_of_number."""
    if number == 0:
        return b""
    size = (number.bit_length() - 1) // 8
    assert 1 <= size <= 7 and number >> (8 * size) == 1, number
    return (number - (1 << (8 * size))).to_bytes(size, "little")


def flat_column(type: int, values: np.ndarray, not_null: Optional[np.ndarray] = None, device="cuda") -> DeviceColumn:
    """Flat vector; `not_null` is a bool mask (True = present) or None.  VARCHAR: `values` = string_views(...)."""
    if type == abi.BOOLEAN:
        v = _dev(pack_bits(np.asarray(values, dtype=bool)).view(np.int64), device)
    elif type == abi.VARCHAR:
        v = _dev(np.asarray(values, dtype=np.int64).reshape(-1, 2), device)
    else:
        v = _dev(np.asarray(values, dtype=NP_OF[type]), device)
    return DeviceColumn(type, abi.FLAT, v, nulls=_bits_dev(not_null, device), length=len(values))


def dictionary_column(type: int, base: np.ndarray, indices: np.ndarray, not_null: Optional[np.ndarray] = None,
                      base_not_null: Optional[np.ndarray] = None, device="cuda") -> DeviceColumn:
    if type == abi.BOOLEAN:
        b = _dev(pack_bits(np.asarray(base, dtype=bool)).view(np.int64), device)
    elif type == abi.VARCHAR:
        b = _dev(np.asarray(base, dtype=np.int64).reshape(-1, 2), device)
    else:
        b = _dev(np.asarray(base, dtype=NP_OF[type]), device)
    return DeviceColumn(type, abi.DICTIONARY, b, _dev(np.asarray(indices, dtype=np.int32), device),
                        _bits_dev(not_null, device), _bits_dev(base_not_null, device), length=len(indices))


def constant_column(type: int, value, length: int) -> DeviceColumn:
    return DeviceColumn(type, abi.CONSTANT, constant=value, constant_is_null=value is None, length=length)


def _col_array(cols: Sequence[DeviceColumn]):
    arr = (abi.Column * max(1, len(cols)))(*[c.desc() for c in cols])
    return arr


class Program:
    """ExprSet analogue: a filter plus output expressions lowered once, reused per batch."""

    def __init__(self, filter: Optional[ex.Expr], outputs: Sequence[ex.Expr], n_columns: Optional[int] = None):
        self.lowered = ex.lower(filter, list(outputs))
        self.output_types = [o.type for o in outputs]
        desc, keep = ex.to_desc(self.lowered, n_columns)
        self._keep = keep
        self.handle = C.c_void_p()
        abi.check(abi.lib().vgpu_program_create(C.byref(desc), C.byref(self.handle)))

    def __del__(self):
        try:
            if self.handle:
                abi.lib().vgpu_program_destroy(self.handle)
        except Exception:
            pass


def check_errors():
    abi.check(abi.lib().vgpu_check_errors(stream_ptr()))


@dataclass
class FilterProjectResult:
    sel_bits: torch.Tensor           # int64 words
    sel_idx: Optional[torch.Tensor]  # int32, first num_selected entries valid
    num_selected: int
    values: List[Optional[torch.Tensor]]
    nulls: List[Optional[torch.Tensor]]


def filter_project(program: Program, cols: Sequence[DeviceColumn], nrows: int, in_rows: Optional[torch.Tensor] = None,
                   want_indices=True, want_nulls=True, sync=True) -> FilterProjectResult:
    """vgpu_filter_project over device-resident columns (FilterProject::getOutput for one batch)."""
    dev = "cuda"
    total = nrows if in_rows is None else len(in_rows)
    words = (total + 63) // 64
    sel_bits = torch.zeros(max(words, 1), dtype=torch.int64, device=dev)
    sel_idx = torch.empty(max(total, 1), dtype=torch.int32, device=dev) if want_indices else None
    nsel = torch.zeros(1, dtype=torch.int64, device=dev)
    outs = (abi.OutColumn * max(1, len(program.output_types)))()
    values, nulls = [], []
    for i, t in enumerate(program.output_types):
        if t == abi.BOOLEAN:
            v = torch.zeros(max(words, 1), dtype=torch.int64, device=dev)
        else:
            v = torch.empty(max(total, 1), dtype=getattr(torch, np.dtype(NP_OF[t]).name), device=dev)
        nb = torch.zeros(max(words, 1), dtype=torch.int64, device=dev) if want_nulls else None
        outs[i].values = v.data_ptr()
        outs[i].nulls = nb.data_ptr() if nb is not None else None
        values.append(v)
        nulls.append(nb)
    arr = _col_array(cols)
    abi.check(abi.lib().vgpu_filter_project(
        program.handle, arr, len(cols), nrows, in_rows.data_ptr() if in_rows is not None else None,
        0 if in_rows is None else len(in_rows), sel_bits.data_ptr(),
        sel_idx.data_ptr() if sel_idx is not None else None, nsel.data_ptr(), outs, stream_ptr()))
    n = -1
    if sync:
        check_errors()
        n = int(nsel.item())
    return FilterProjectResult(sel_bits, sel_idx, n, values, nulls)


class Aggregation:
    """GroupingSet analogue over the C-ABI."""

    def __init__(self, mode: int, keys: Sequence[tuple], accs: Sequence[tuple], ignore_null_keys=False,
                 expected_groups=0):
        """keys: (output index, bits, min, max); accs: (vgpu_acc_kind, output index or -1)."""
        self.n_accs = len(accs)
        self._keys = (abi.KeySpec * max(1, len(keys)))(*[abi.KeySpec(*k) for k in keys])
        self._accs = (abi.AccSpec * max(1, len(accs)))(*[abi.AccSpec(*a) for a in accs])
        d = abi.AggDesc()
        d.mode, d.n_keys, d.n_accs = mode, len(keys), len(accs)
        d.keys = C.cast(self._keys, C.POINTER(abi.KeySpec))
        d.accs = C.cast(self._accs, C.POINTER(abi.AccSpec))
        d.ignore_null_keys = 1 if ignore_null_keys else 0
        d.expected_groups = expected_groups
        self.handle = C.c_void_p()
        abi.check(abi.lib().vgpu_agg_create(C.byref(d), C.byref(self.handle)))
        self.row_words = 3 + 2 * self.n_accs

    def __del__(self):
        try:
            if self.handle:
                abi.lib().vgpu_agg_destroy(self.handle)
        except Exception:
            pass

    def add_input(self, program: Program, cols: Sequence[DeviceColumn], nrows: int,
                  in_rows: Optional[torch.Tensor] = None):
        arr = _col_array(cols)
        abi.check(abi.lib().vgpu_agg_add_input(
            self.handle, program.handle, arr, len(cols), nrows,
            in_rows.data_ptr() if in_rows is not None else None, 0 if in_rows is None else len(in_rows),
            stream_ptr()))

    def reset(self):
        abi.check(abi.lib().vgpu_agg_reset(self.handle, stream_ptr()))

    def finalize(self):
        """noMoreInput: completes the device-side aggregation (asynchronous; more input may still follow)."""
        abi.check(abi.lib().vgpu_agg_finalize(self.handle, stream_ptr()))

    def check_errors(self):
        abi.check(abi.lib().vgpu_agg_check_errors(self.handle, stream_ptr()))

    def info(self) -> dict:
        out = (C.c_int64 * 8)()
        abi.check(abi.lib().vgpu_debug_agg_info(self.handle, stream_ptr(), out))
        names = ["engine", "sorted", "front_entries", "deduped_entries", "deduped_groups", "partitions", "entry_words", "entry_bound"]
        return dict(zip(names, list(out)))

    def num_groups(self) -> int:
        n = C.c_int64()
        abi.check(abi.lib().vgpu_agg_num_groups(self.handle, stream_ptr(), C.byref(n)))
        return n.value

    def extract(self) -> np.ndarray:
        """All groups as an int64 array [n_groups, 3 + 2*n_accs] (key, flags, rows, values, counts)."""
        n = self.num_groups()
        buf = torch.empty((max(n, 1), self.row_words), dtype=torch.int64, device="cuda")
        got = C.c_int64()
        abi.check(abi.lib().vgpu_agg_extract(self.handle, 0, n, buf.data_ptr(), C.byref(got), stream_ptr()))
        check_errors()
        return buf[: got.value].cpu().numpy()

    def merge_rows(self, rows: torch.Tensor):
        assert rows.dtype == torch.int64 and rows.is_contiguous()
        n = rows.shape[0] if rows.dim() == 2 else rows.numel() // self.row_words
        abi.check(abi.lib().vgpu_agg_merge_rows(self.handle, rows.data_ptr(), n, stream_ptr()))

    def partition(self, n_parts: int):
        """Returns (rows tensor [n_groups, row_words] grouped by destination, counts list)."""
        n = self.num_groups()
        out = torch.empty((max(n, 1), self.row_words), dtype=torch.int64, device="cuda")
        counts = (C.c_int64 * n_parts)()
        abi.check(abi.lib().vgpu_agg_partition(self.handle, n_parts, out.data_ptr(), counts, stream_ptr()))
        return out[:n], list(counts)


class HostBatchRunner:
    """End-to-end path for HOST-resident batches: pinned host columns are copied to HBM in chunks on
    a copy stream (double-buffered) while the compute stream runs the fused pipeline on the previous
    chunk — what the C++ operator does with a host RowVector handed to addInput."""

    def __init__(self, types: Sequence[int], chunk_rows: int = 4 << 20, n_buffers: int = 2):
        assert chunk_rows % 1024 == 0
        self.types = list(types)
        self.chunk = chunk_rows
        self.copy_stream = torch.cuda.Stream()
        self.bufs = [[torch.empty(chunk_rows, dtype=getattr(torch, np.dtype(NP_OF[t]).name), device="cuda")
                      for t in self.types] for _ in range(n_buffers)]
        self.ready = [torch.cuda.Event() for _ in range(n_buffers)]
        self.free = [torch.cuda.Event() for _ in range(n_buffers)]
        self.h2d_bytes = 0

    def run(self, agg: "Aggregation", program: Program, host_cols: Sequence[torch.Tensor], n: int):
        compute = torch.cuda.current_stream()
        nb = len(self.bufs)
        self.h2d_bytes = 0
        for i, off in enumerate(range(0, n, self.chunk)):
            m = min(self.chunk, n - off)
            b = i % nb
            if i >= nb:
                self.copy_stream.wait_event(self.free[b])
            with torch.cuda.stream(self.copy_stream):
                for dst, src in zip(self.bufs[b], host_cols):
                    dst[:m].copy_(src[off:off + m], non_blocking=True)
                    self.h2d_bytes += m * src.element_size()
                self.ready[b].record(self.copy_stream)
            compute.wait_event(self.ready[b])
            cols = [DeviceColumn(t, abi.FLAT, buf, length=m) for t, buf in zip(self.types, self.bufs[b])]
            agg.add_input(program, cols, m)
            self.free[b].record(compute)


class Exchange:
    """Multi-GPU step between partial and final aggregation (one process per GPU): hash-partition the
    partial groups, all-to-all the packed rows over NCCL, fold them into the final aggregation
    (partialAggregation -> partitionedOutput -> exchange -> finalAggregation of the reference)."""

    def __init__(self, rank: int, world: int, unique_id: bytes):
        self.rank, self.world = rank, world
        idbuf = (C.c_uint8 * abi.COMM_ID_BYTES).from_buffer_copy(unique_id)
        self.handle = C.c_void_p()
        abi.check(abi.lib().vgpu_comm_create(idbuf, rank, world, C.byref(self.handle)))
        self.bytes_sent = 0

    def reserve_window(self, nbytes: int) -> bool:
        """Collective: peer-mapped exchange window of nbytes per rank (same value on every rank).  True when mapped."""
        ok = C.c_int32()
        abi.check(abi.lib().vgpu_comm_reserve_window(self.handle, nbytes, C.byref(ok)))
        return bool(ok.value)

    def last_counts(self) -> np.ndarray:
        m = (C.c_int64 * (self.world * self.world))()
        abi.check(abi.lib().vgpu_comm_last_counts(self.handle, m, stream_ptr()))
        return np.array(list(m), dtype=np.int64).reshape(self.world, self.world)

    def attach(self, partial: "Aggregation"):
        """vgpu_agg_attach_exchange: the partial aggregation's scans store their entries straight into the owners'
        windows (needs reserve_window); shuffle() then only exchanges counts and deduplicates."""
        abi.check(abi.lib().vgpu_agg_attach_exchange(partial.handle, self.handle))

    @staticmethod
    def detach(partial: "Aggregation"):
        abi.check(abi.lib().vgpu_agg_attach_exchange(partial.handle, None))

    @staticmethod
    def make_unique_id() -> bytes:
        idbuf = (C.c_uint8 * abi.COMM_ID_BYTES)()
        abi.check(abi.lib().vgpu_comm_unique_id(idbuf))
        return bytes(idbuf)

    def shuffle(self, partial: "Aggregation", final: "Aggregation"):
        """One call: vgpu_comm_exchange_groups (dense all-gather for ARRAY / GLOBAL aggregations, partition +
        counts + all-to-all for HASH ones).  `final` must have seen an (empty) add_input."""
        sent = C.c_int64()
        abi.check(abi.lib().vgpu_comm_exchange_groups(self.handle, partial.handle, final.handle, stream_ptr(), C.byref(sent)))
        self.bytes_sent = sent.value

    def shuffle_stepwise(self, partial: "Aggregation", final: "Aggregation"):
        """The same exchange spelled out with the individual entry points (partition, counts all-gather,
        all-to-all, merge); kept as the reference for vgpu_comm_exchange_groups in the multi-GPU test."""
        W = self.world
        rows, counts = partial.partition(W)
        row_bytes = partial.row_words * 8
        send_counts = torch.tensor(counts, dtype=torch.int64, device="cuda")
        all_counts = torch.empty(W * W, dtype=torch.int64, device="cuda")
        abi.check(abi.lib().vgpu_comm_allgather(self.handle, send_counts.data_ptr(), all_counts.data_ptr(), W * 8,
                                                stream_ptr()))
        matrix = all_counts.cpu().numpy().reshape(W, W)  # [src][dst]
        recv_counts = matrix[:, self.rank]
        sc, so, rc, ro = plan_alltoall(counts, recv_counts, row_bytes)
        recv = torch.empty((max(int(recv_counts.sum()), 1), partial.row_words), dtype=torch.int64, device="cuda")
        arr = lambda xs: (C.c_int64 * W)(*[int(x) for x in xs])
        abi.check(abi.lib().vgpu_comm_alltoallv(self.handle, rows.data_ptr(), arr(sc), arr(so), recv.data_ptr(), arr(rc),
                                                arr(ro), stream_ptr()))
        self.bytes_sent = int(sum(sc)) - int(sc[self.rank])
        final.merge_rows(recv[: int(recv_counts.sum())])


TORCH_OF = {"l_orderkey": torch.int64, "l_partkey": torch.int64, "l_suppkey": torch.int64, "l_linenumber": torch.int32,
            "l_quantity": torch.float64, "l_extendedprice": torch.float64, "l_discount": torch.float64, "l_tax": torch.float64,
            "l_returnflag": torch.int32, "l_linestatus": torch.int32, "l_shipdate": torch.int32, "l_commitdate": torch.int32,
            "l_receiptdate": torch.int32, "l_shipinstruct": torch.int32, "l_shipmode": torch.int32}


def tpch_lineitem_rows(sf: float, order_offset: int, max_orders: int) -> int:
    n = C.c_int64()
    abi.check(abi.lib().vgpu_tpch_lineitem_rows(float(sf), order_offset, max_orders, C.byref(n), stream_ptr()))
    return n.value


def tpch_lineitem(sf: float, columns, order_offset: int = 0, max_orders: int = None, rows: int = None):
    """TableScan over the TPC-H connector on the device: the lineitem rows of orders [order_offset, +max_orders) at
    scale factor sf, generated straight into HBM bit-identically to the reference's dbgen (vgpu_tpch_gen_lineitem).
    Returns dict name -> tensor (VARCHAR columns as int32 dictionary ids)."""
    total_orders = int(1500000 * sf)
    if max_orders is None:
        max_orders = max(total_orders - order_offset, 0)
    if rows is None:
        rows = tpch_lineitem_rows(sf, order_offset, max_orders)
    out = {c: torch.empty(max(rows, 1), dtype=TORCH_OF[c], device="cuda") for c in columns}
    desc = abi.LineitemColumns()
    for c, t in out.items():
        setattr(desc, c, t.data_ptr())
    got = C.c_int64()
    abi.check(abi.lib().vgpu_tpch_gen_lineitem(float(sf), order_offset, max_orders, C.byref(desc), rows, C.byref(got), stream_ptr()))
    check_errors()
    assert got.value == rows, (got.value, rows)
    return {c: t[:rows] for c, t in out.items()}


def exchange_local(partials, finals, window_bytes=256 << 20):
    """All ranks of a window exchange emulated on this one GPU (vgpu_debug_exchange_local): rank r owns
    partials[r] -> finals[r]."""
    w = len(partials)
    pa = (C.c_void_p * w)(*[p.handle for p in partials])
    fa = (C.c_void_p * w)(*[f.handle for f in finals])
    abi.check(abi.lib().vgpu_debug_exchange_local(pa, fa, w, window_bytes, stream_ptr()))


class LocalExchangeGroup:
    """`world` ranks of a window exchange on this one GPU (vgpu_debug_local_comms): for tests and smoke() of the exchange
    fused into the scan.  attach(r, partial) before the partial's add_input; shuffle(partials, finals) for all ranks."""

    def __init__(self, world: int, window_bytes=256 << 20):
        self.world = world
        self.handles = (C.c_void_p * world)()
        abi.check(abi.lib().vgpu_debug_local_comms(world, window_bytes, self.handles))

    def attach(self, rank: int, partial: "Aggregation"):
        abi.check(abi.lib().vgpu_agg_attach_exchange(partial.handle, self.handles[rank]))

    def shuffle(self, partials, finals):
        w = self.world
        pa = (C.c_void_p * w)(*[p.handle for p in partials])
        fa = (C.c_void_p * w)(*[f.handle for f in finals])
        abi.check(abi.lib().vgpu_debug_exchange_local_comms(self.handles, pa, fa, w, stream_ptr()))

    def close(self):
        if self.handles is not None:
            torch.cuda.synchronize()
            for h in self.handles:
                if h:
                    abi.lib().vgpu_comm_destroy(C.c_void_p(h))
            self.handles = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def plan_alltoall(send_rows, recv_rows, row_bytes):
    """Row counts per peer -> byte counts and byte offsets of the contiguous send / receive segments."""
    sc = [int(x) * row_bytes for x in send_rows]
    rc = [int(x) * row_bytes for x in recv_rows]
    so = [0] * len(sc)
    ro = [0] * len(rc)
    for i in range(1, len(sc)):
        so[i] = so[i - 1] + sc[i - 1]
        ro[i] = ro[i - 1] + rc[i - 1]
    return sc, so, rc, ro
