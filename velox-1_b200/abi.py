"""ctypes view of include/vgpu.h (the C-ABI of libvgpu.so).

Everything above this module talks to the CUDA side only through these prototypes; there is no
CPU fallback — a missing or unloadable library raises at import of the functions.
"""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libvgpu.so")

# ---- enums (mirror include/vgpu.h) -----------------------------------------------------------
OK, ERR_INVALID, ERR_CUDA, ERR_USER, ERR_UNSUPPORTED, ERR_NCCL = range(6)
BOOLEAN, TINYINT, SMALLINT, INTEGER, BIGINT, REAL, DOUBLE, VARCHAR = range(8)
FLAT, DICTIONARY, CONSTANT = range(3)
OPND_NONE, OPND_ACC, OPND_SLOT, OPND_COL, OPND_IMM, OPND_NULL = range(6)
OP_NOP, OP_ADD, OP_SUB, OP_MUL, OP_DIV, OP_NEG, OP_MOD = range(7)
OP_CAST, OP_MOV = 8, 9
OP_SELECT, OP_COALESCE = 10, 11
OP_EQ, OP_NE, OP_LT, OP_LE, OP_GT, OP_GE, OP_BETWEEN = range(16, 23)
OP_AND, OP_OR, OP_NOT, OP_IS_NULL = range(24, 28)
ACC_SUM_F64, ACC_SUM_I64, ACC_COUNT, ACC_MIN, ACC_MAX = range(5)
AGG_GLOBAL, AGG_ARRAY, AGG_HASH = range(3)
COMM_ID_BYTES = 128


class Scalar(C.Union):
    _fields_ = [("i", C.c_int64), ("d", C.c_double)]


class Column(C.Structure):
    _fields_ = [
        ("type", C.c_int32), ("encoding", C.c_int32),
        ("values", C.c_void_p), ("indices", C.c_void_p),
        ("nulls", C.c_void_p), ("base_nulls", C.c_void_p),
        ("base_length", C.c_int64), ("constant", Scalar),
        ("constant_is_null", C.c_int32), ("reserved", C.c_int32),
    ]


class Operand(C.Structure):
    _fields_ = [("kind", C.c_uint8), ("type", C.c_uint8), ("index", C.c_uint16)]


class Op(C.Structure):
    _fields_ = [
        ("opcode", C.c_uint8), ("type", C.c_uint8), ("dst_slot", C.c_int16),
        ("a", Operand), ("b", Operand), ("c", Operand),
        ("imm", Scalar), ("imm2", Scalar),
    ]


class ProgramDesc(C.Structure):
    _fields_ = [
        ("ops", C.POINTER(Op)), ("n_ops", C.c_int32), ("filter", Operand),
        ("outputs", C.POINTER(Operand)), ("n_outputs", C.c_int32), ("n_columns", C.c_int32),
    ]


class OutColumn(C.Structure):
    _fields_ = [("values", C.c_void_p), ("nulls", C.c_void_p)]


class AccSpec(C.Structure):
    _fields_ = [("kind", C.c_int32), ("input", C.c_int32)]


class KeySpec(C.Structure):
    _fields_ = [("output", C.c_int32), ("bits", C.c_int32), ("min", C.c_int64), ("max", C.c_int64)]


class AggDesc(C.Structure):
    _fields_ = [
        ("mode", C.c_int32), ("n_keys", C.c_int32), ("keys", C.POINTER(KeySpec)),
        ("n_accs", C.c_int32), ("accs", C.POINTER(AccSpec)),
        ("ignore_null_keys", C.c_int32), ("reserved", C.c_int32), ("expected_groups", C.c_int64),
    ]


LINEITEM_COLUMNS = ["l_orderkey", "l_partkey", "l_suppkey", "l_linenumber", "l_quantity", "l_extendedprice", "l_discount",
                    "l_tax", "l_returnflag", "l_linestatus", "l_shipdate", "l_commitdate", "l_receiptdate", "l_shipinstruct",
                    "l_shipmode"]


class LineitemColumns(C.Structure):
    _fields_ = [(name, C.c_void_p) for name in LINEITEM_COLUMNS]


class Stats(C.Structure):
    _fields_ = [("kernel_launches", C.c_int64), ("rows_in", C.c_int64),
                ("bytes_h2d", C.c_int64), ("bytes_d2h", C.c_int64)]


assert C.sizeof(Op) == 32 and C.sizeof(Column) == 64 and C.sizeof(Operand) == 4

# name -> argtypes; every function returns int (vgpu_status)
_P, _I32, _I64, _SZ = C.c_void_p, C.c_int32, C.c_int64, C.c_size_t
PROTOTYPES = {
    "vgpu_abi_version": [],
    "vgpu_last_error": [C.c_char_p, _SZ],
    "vgpu_init": [C.c_int],
    "vgpu_device_count": [C.POINTER(C.c_int)],
    "vgpu_shutdown": [],
    "vgpu_alloc": [_SZ, C.POINTER(_P)],
    "vgpu_free": [_P],
    "vgpu_alloc_host": [_SZ, C.POINTER(_P)],
    "vgpu_free_host": [_P],
    "vgpu_h2d_async": [_P, _P, _SZ, _P],
    "vgpu_d2h_async": [_P, _P, _SZ, _P],
    "vgpu_memset_async": [_P, C.c_int, _SZ, _P],
    "vgpu_stream_create": [C.POINTER(_P)],
    "vgpu_stream_destroy": [_P],
    "vgpu_stream_sync": [_P],
    "vgpu_event_create": [C.POINTER(_P)],
    "vgpu_event_destroy": [_P],
    "vgpu_event_record": [_P, _P],
    "vgpu_stream_wait_event": [_P, _P],
    "vgpu_event_sync": [_P],
    "vgpu_host_is_pinned": [_P, C.POINTER(_I32)],
    "vgpu_translate_i32": [_P, _P, _P, _I32, _I64, _P],
    "vgpu_program_create": [C.POINTER(ProgramDesc), C.POINTER(_P)],
    "vgpu_program_destroy": [_P],
    "vgpu_filter_project": [_P, C.POINTER(Column), _I32, _I64, _P, _I64, _P, _P, _P,
                            C.POINTER(OutColumn), _P],
    "vgpu_selection_indices_noalloc": [_P, _I64, _P, _P, _P],
    "vgpu_check_errors": [_P],
    "vgpu_agg_create": [C.POINTER(AggDesc), C.POINTER(_P)],
    "vgpu_agg_destroy": [_P],
    "vgpu_agg_reset": [_P, _P],
    "vgpu_agg_add_input": [_P, _P, C.POINTER(Column), _I32, _I64, _P, _I64, _P],
    "vgpu_agg_finalize": [_P, _P],
    "vgpu_agg_check_errors": [_P, _P],
    "vgpu_agg_memory_bytes": [_P, C.POINTER(_I64)],
    "vgpu_agg_num_groups": [_P, _P, C.POINTER(_I64)],
    "vgpu_agg_row_bytes": [_P, C.POINTER(_I32)],
    "vgpu_agg_extract": [_P, _I64, _I64, _P, C.POINTER(_I64), _P],
    "vgpu_agg_merge_rows": [_P, _P, _I64, _P],
    "vgpu_agg_partition": [_P, _I32, _P, C.POINTER(_I64), _P],
    "vgpu_partition_of": [_I64, _I32, _I32, _I32, C.POINTER(_I32)],
    "vgpu_comm_unique_id": [C.POINTER(C.c_uint8)],
    "vgpu_comm_create": [C.POINTER(C.c_uint8), _I32, _I32, C.POINTER(_P)],
    "vgpu_comm_destroy": [_P],
    "vgpu_comm_alltoallv": [_P, _P, C.POINTER(_I64), C.POINTER(_I64), _P, C.POINTER(_I64),
                            C.POINTER(_I64), _P],
    "vgpu_comm_allgather": [_P, _P, _P, _SZ, _P],
    "vgpu_comm_exchange_groups": [_P, _P, _P, _P, C.POINTER(_I64)],
    "vgpu_comm_reserve_window": [_P, _SZ, C.POINTER(_I32)],
    "vgpu_comm_last_counts": [_P, C.POINTER(_I64), _P],
    "vgpu_tpch_lineitem_rows": [C.c_double, _I64, _I64, C.POINTER(_I64), _P],
    "vgpu_tpch_gen_lineitem": [C.c_double, _I64, _I64, C.POINTER(LineitemColumns), _I64, C.POINTER(_I64), _P],
    "vgpu_get_stats": [C.POINTER(Stats)],
    "vgpu_reset_stats": [],
    "vgpu_debug_set_dry_run": [C.c_int],
    "vgpu_debug_last_kernel_source": [C.c_char_p, _SZ, C.POINTER(C.c_uint64)],
    "vgpu_debug_last_kernel_note": [C.c_char_p, _SZ],
    "vgpu_debug_agg_info": [_P, _P, C.POINTER(_I64)],
    "vgpu_debug_exchange_local": [C.POINTER(_P), C.POINTER(_P), _I32, _SZ, _P],
    "vgpu_agg_attach_exchange": [_P, _P],
    "vgpu_debug_local_comms": [_I32, _SZ, C.POINTER(_P)],
    "vgpu_debug_exchange_local_comms": [C.POINTER(_P), C.POINTER(_P), C.POINTER(_P), _I32, _P],
    "vgpu_debug_kernel_timing": [C.c_int],
    "vgpu_debug_kernel_time": [C.POINTER(C.c_double), C.POINTER(_I64)],
}

_lib = None


class VgpuError(RuntimeError):
    def __init__(self, status, message):
        super().__init__(f"vgpu status {status}: {message}")
        self.status = status
        self.message = message


def lib():
    """Loads libvgpu.so (once).  Raises if it has not been built: there is no fallback."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise ImportError(f"{LIB_PATH} is missing: run `python -c 'import __graft_entry__ as g; g.build()'`")
        handle = C.CDLL(LIB_PATH, mode=C.RTLD_GLOBAL)
        for name, argtypes in PROTOTYPES.items():
            fn = getattr(handle, name)  # AttributeError if the symbol is not exported
            fn.argtypes = argtypes
            fn.restype = C.c_int
        _lib = handle
    return _lib


def last_error():
    buf = C.create_string_buffer(512)
    lib().vgpu_last_error(buf, 512)
    return buf.value.decode(errors="replace")


def last_kernel_note():
    buf = C.create_string_buffer(4096)
    lib().vgpu_debug_last_kernel_note(buf, 4096)
    return buf.value.decode(errors="replace")


def last_kernel_source():
    """(source body, cubin hash) of the last plan-specialised kernel generated on this thread."""
    buf = C.create_string_buffer(1 << 20)
    h = C.c_uint64()
    lib().vgpu_debug_last_kernel_source(buf, 1 << 20, C.byref(h))
    return buf.value.decode(), h.value


def check(status):
    if status != OK:
        raise VgpuError(status, last_error())
