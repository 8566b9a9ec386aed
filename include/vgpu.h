/*
 * vgpu.h — C ABI of the B200 (sm_100a) filter-and-aggregate engine.
 *
 * This is the drop-in boundary: everything the reference's operators would bind through FFI to
 * run FilterProject -> HashAggregation on the GPU.  Plain pointers and sizes only; no C++ or
 * torch types.  Every call returns a vgpu_status (0 = OK); the message of the last failure on the
 * calling thread is available from vgpu_last_error().  No exception crosses this boundary — the
 * C++ operators (velox-1_b200/host) translate a non-zero status into VELOX_USER_FAIL / VELOX_FAIL
 * the way velox/common/base/Exceptions.h does for the CPU operators.
 *
 * Reference interfaces each group of entry points replaces (paths under the reference tree):
 *   vgpu_column            <- velox/vector/DecodedVector.h:76-140 (base values + indices + nulls)
 *   vgpu_op / vgpu_program <- velox/expression/Expr.h (ExprSet, compiled Expr tree),
 *                             velox/expression/ExprCompiler.cpp:302-335
 *   vgpu_filter_project    <- velox/exec/FilterProject.cpp:125-191 (getOutput/filter/project),
 *                             velox/exec/OperatorUtils.cpp:177-255 (processFilterResults)
 *   vgpu_agg_*             <- velox/exec/HashAggregation.cpp:235-274,388-464,
 *                             velox/exec/GroupingSet.cpp:177-307,513-559,644-712,
 *                             velox/exec/HashTable.cpp:426-562, velox/exec/VectorHasher.cpp:68-332,
 *                             velox/functions/lib/aggregates/{SumAggregateBase,AverageAggregateBase}.h,
 *                             velox/functions/prestosql/aggregates/{Count,MinMax}Aggregate*.cpp
 *   vgpu_agg_partition / vgpu_agg_merge_rows / vgpu_comm_*
 *                          <- velox/exec/HashPartitionFunction.cpp:59-92 + PartitionedOutput/Exchange
 *                             (partial -> repartition by key -> final,
 *                             velox/exec/tests/MultiFragmentTest.cpp:180-212)
 *   (input data: velox/tpch/gen/TpchGen.cpp:401-489 is restated for tests in oracle/ref_dbgen_shim.cpp,
 *    not part of this ABI)
 *
 * Bitmaps follow velox/common/base/Nulls.h:26-35: uint64 little-endian words, bit i of a nulls
 * bitmap is 1 when row i is NOT null; a NULL bitmap pointer means "no nulls".
 * All device pointers must be valid on the device the call runs on (the current CUDA device);
 * `stream` is a cudaStream_t passed as void* (NULL = the legacy default stream).
 */
#ifndef VGPU_H_
#define VGPU_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define VGPU_ABI_VERSION 1

typedef enum {
  VGPU_OK = 0,
  VGPU_ERR_INVALID = 1,     /* bad argument / malformed program (VELOX_CHECK analogue)          */
  VGPU_ERR_CUDA = 2,        /* CUDA runtime failure                                             */
  VGPU_ERR_USER = 3,        /* data error raised by a kernel: integer overflow, division by zero,
                               cast out of range (VeloxUserError analogue)                      */
  VGPU_ERR_UNSUPPORTED = 4, /* valid Velox plan shape this build does not cover                 */
  VGPU_ERR_NCCL = 5
} vgpu_status;

/* Device value types.  Numbering follows velox/type/Type.h:62-84 (TypeKind).  DATE is INTEGER (days since epoch).
 * VARCHAR reaches the device in one of two ways:
 *  - as int32 value ids: the host maps strings to ids (VectorHasher::makeValueIds analogue) and keeps the dictionary;
 *  - as VGPU_VARCHAR: `values` is the vector's own array of 16-byte StringViews (velox/type/StringView.h: u32 size,
 *    4 prefix bytes, then 8 inline bytes or a pointer), 16-byte aligned, FLAT or as the base of a DICTIONARY.  The
 *    kernels read a value as the BIGINT the reference uses for short string keys (velox/exec/VectorHasher.h:324-329
 *    stringAsNumber: the bytes as a little-endian word + 1 << (8 * size); "" is 0), so programs read the column through
 *    BIGINT operands and group keys come back as that number (vgpu_string_of_number turns it back into bytes).  Only
 *    strings of at most 7 bytes have a number: a longer one that a passing row uses latches an error that
 *    vgpu_check_errors reports as VGPU_ERR_UNSUPPORTED — the host checks lengths first and maps batches with long
 *    strings to numbers >= 2^60 of its own.  A CONSTANT VGPU_VARCHAR column carries the number in `constant.i`.
 *    VGPU_VARCHAR columns need the plan-specialised kernels (NVRTC or an AOT plan). */
typedef enum {
  VGPU_BOOLEAN = 0, /* bit-packed, 1 bit per row */
  VGPU_TINYINT = 1,
  VGPU_SMALLINT = 2,
  VGPU_INTEGER = 3,
  VGPU_BIGINT = 4,
  VGPU_REAL = 5,
  VGPU_DOUBLE = 6,
  VGPU_VARCHAR = 7 /* 16-byte StringViews, read as their stringAsNumber value */
} vgpu_type;

/* Host-side helpers for VGPU_VARCHAR keys: the number of a string of at most 7 bytes (-1 when longer), and back
 * (returns the size; `out` receives the bytes, no terminator; -1 when `number` is not such a number). */
static inline int64_t vgpu_number_of_string(const char* data, uint32_t size) {
  uint64_t w = 0;
  uint32_t i;
  if (size > 7u) return -1;
  for (i = 0; i < size; ++i) w |= (uint64_t)(uint8_t)data[i] << (8u * i);
  return size == 0u ? 0 : (int64_t)(w + ((uint64_t)1 << (8u * size)));
}
static inline int32_t vgpu_string_of_number(int64_t number, char out[8]) {
  const uint64_t w = (uint64_t)number;
  int32_t size, i;
  if (w == 0) return 0;
  for (size = 1; size <= 7; ++size) {
    if ((w >> (8 * size)) == 1u) { /* the marker bit of a string of `size` bytes */
      for (i = 0; i < size; ++i) out[i] = (char)(w >> (8 * i));
      return size;
    }
  }
  return -1;
}

typedef enum { VGPU_FLAT = 0, VGPU_DICTIONARY = 1, VGPU_CONSTANT = 2 } vgpu_encoding;

typedef union {
  int64_t i;
  double d;
} vgpu_scalar;

/* One input column as the kernels see it (the DecodedVector triple). */
typedef struct {
  int32_t type;               /* vgpu_type of the (base) values                                   */
  int32_t encoding;           /* vgpu_encoding                                                    */
  const void* values;         /* FLAT: `nrows` values.  DICTIONARY: `base_length` base values.     */
  const int32_t* indices;     /* DICTIONARY: `nrows` indices into values; else NULL               */
  const uint64_t* nulls;      /* per OUTER row; NULL = none                                       */
  const uint64_t* base_nulls; /* DICTIONARY: per base value; NULL = none                          */
  int64_t base_length;        /* DICTIONARY only                                                  */
  vgpu_scalar constant;       /* CONSTANT only                                                    */
  int32_t constant_is_null;   /* CONSTANT only                                                    */
  int32_t reserved;
} vgpu_column;

/* ---- expression programs ------------------------------------------------------------------ */

/* Operand kinds of an op. */
typedef enum {
  VGPU_OPND_NONE = 0,
  VGPU_OPND_ACC = 1,  /* result of the previous value op (kept in registers)                   */
  VGPU_OPND_SLOT = 2, /* scratch value slot `index` (shared memory)                            */
  VGPU_OPND_COL = 3,  /* input column `index`                                                  */
  VGPU_OPND_IMM = 4,  /* the op's immediate (typed like the op); never null                    */
  VGPU_OPND_NULL = 5  /* typed NULL constant                                                   */
} vgpu_operand_kind;

typedef struct {
  uint8_t kind;  /* vgpu_operand_kind */
  uint8_t type;  /* vgpu_type of the value read through this operand                            */
  uint16_t index;
} vgpu_operand;

/* Opcodes.  Semantics cite the reference function each one restates.
 * "default null": result is NULL when any argument is NULL (velox/expression/Expr.cpp:371-441). */
typedef enum {
  VGPU_OP_NOP = 0,
  /* arithmetic, default null.  DOUBLE/REAL: IEEE (velox/functions/prestosql/Arithmetic.h:45-86);
   * integers: checked, raising VGPU_ERR_USER on overflow or division by zero
   * (velox/functions/prestosql/registration/CheckedArithmeticRegistration.cpp:23-27,
   *  velox/common/base/CheckedArithmetic.h). */
  VGPU_OP_ADD = 1,
  VGPU_OP_SUB = 2,
  VGPU_OP_MUL = 3,
  VGPU_OP_DIV = 4,
  VGPU_OP_NEG = 5,
  VGPU_OP_MOD = 6,
  /* cast a -> op.type, default null; FP->integer rounds half away from zero and range-checks,
   * NaN raises (velox/type/Conversions.h:176-302); integer->FP is static_cast (:304-374);
   * integer narrowing range-checks. */
  VGPU_OP_CAST = 8,
  /* move a value operand into ACC (and optionally a slot) — FieldReference / ConstantExpr. */
  VGPU_OP_MOV = 9,
  /* special forms over values (not default null).  Like the reference, a branch is only "evaluated"
   * on the rows that take it: an error latched in the branch a row does not take is dropped.
   *   SELECT  : if(a, b, c) — a is a BOOLEAN operand; rows where a is TRUE take b, rows where a is
   *             FALSE or NULL take c (c of kind VGPU_OPND_NULL = no ELSE branch); an error in a is
   *             kept (velox/expression/SwitchExpr.cpp:64-160; CASE WHEN chains nest in c).
   *   COALESCE: coalesce(a, b) — a where it is not NULL, else b; b's errors only count on rows
   *             where a is NULL (velox/expression/CoalesceExpr.cpp:39-81; n-ary nests in b).
   * b is read with `imm`, c with `imm2`.  Result type = op.type, never BOOLEAN. */
  VGPU_OP_SELECT = 10,
  VGPU_OP_COALESCE = 11,
  /* comparisons a ? b on op.type, default null, boolean result
   * (velox/functions/prestosql/Comparisons.h:43-80; NaN compares false except NE). */
  VGPU_OP_EQ = 16,
  VGPU_OP_NE = 17,
  VGPU_OP_LT = 18,
  VGPU_OP_LE = 19,
  VGPU_OP_GT = 20,
  VGPU_OP_GE = 21,
  /* a BETWEEN b AND c, inclusive, default null (Comparisons.h:175-185). */
  VGPU_OP_BETWEEN = 22,
  /* boolean special forms with SQL three-valued logic; an error or NULL in one conjunct is
   * dropped when another conjunct decides the row (velox/expression/ConjunctExpr.cpp:24-174). */
  VGPU_OP_AND = 24,
  VGPU_OP_OR = 25,
  VGPU_OP_NOT = 26,     /* default null (velox/functions/prestosql/Not.cpp)                    */
  VGPU_OP_IS_NULL = 27, /* never null (velox/functions/lib/IsNull.cpp)                          */
  /* a IN (set of up to 8 immediates packed in the following VGPU_OP_NOP ops' imm fields) is
   * compiled by the host into OR of EQ; not a device opcode. */
  VGPU_OP_MAX_ = 32
} vgpu_opcode;

/* One instruction.  The result always lands in ACC; when dst_slot >= 0 it is also stored to that
 * scratch slot.  Boolean results live in the boolean ACC / boolean slots (separate numbering). */
typedef struct {
  uint8_t opcode;   /* vgpu_opcode */
  uint8_t type;     /* operand type the op computes in (result type for CAST / MOV)            */
  int16_t dst_slot; /* -1 = ACC only                                                            */
  vgpu_operand a, b, c;
  vgpu_scalar imm;  /* value read by VGPU_OPND_IMM operands of this op                          */
  vgpu_scalar imm2; /* second immediate (BETWEEN upper bound when c is IMM)                     */
} vgpu_op;

/* A compiled pipeline program: ops, an optional filter and the output list.
 *   filter : boolean operand (ACC/SLOT/COL) or kind NONE.  A row passes when the filter is TRUE
 *            and not NULL (velox/exec/OperatorUtils.cpp:187-210).
 *   outputs: value operands evaluated for passing rows — the projections of FilterProject, or the
 *            grouping keys followed by aggregate inputs when the program feeds an aggregation. */
typedef struct {
  const vgpu_op* ops;
  int32_t n_ops;
  vgpu_operand filter;
  const vgpu_operand* outputs;
  int32_t n_outputs;
  int32_t n_columns; /* number of input columns the program refers to (for validation)          */
} vgpu_program_desc;

typedef struct vgpu_program_s* vgpu_program_t;

/* ---- library / device -------------------------------------------------------------------- */

int vgpu_abi_version(void);
/* Copies the calling thread's last error message (NUL-terminated, truncated to cap). */
int vgpu_last_error(char* buf, size_t cap);
/* Selects `device` for the calling thread and checks it is an sm_100 part. */
int vgpu_init(int device);
int vgpu_device_count(int* count);
int vgpu_shutdown(void);

/* Device / pinned-host memory and copies (the operators' pool()->allocate analogue). */
int vgpu_alloc(size_t bytes, void** dptr);
int vgpu_free(void* dptr);
int vgpu_alloc_host(size_t bytes, void** hptr); /* pinned */
int vgpu_free_host(void* hptr);
int vgpu_h2d_async(void* dst, const void* src, size_t bytes, void* stream);
int vgpu_d2h_async(void* dst, const void* src, size_t bytes, void* stream);
int vgpu_memset_async(void* dst, int value, size_t bytes, void* stream);
int vgpu_stream_create(void** stream);
int vgpu_stream_destroy(void* stream);
int vgpu_stream_sync(void* stream);
/* Events order a copy stream against a compute stream without blocking the host (staging rings of the operators). */
int vgpu_event_create(void** event);
int vgpu_event_destroy(void* event);
int vgpu_event_record(void* event, void* stream);
int vgpu_stream_wait_event(void* stream, void* event);
int vgpu_event_sync(void* event);
/* *pinned = 1 when `hptr` is page-locked host memory (vgpu_alloc_host or registered): an async copy from it does
 * not stage through the driver. */
int vgpu_host_is_pinned(const void* hptr, int32_t* pinned);

/* dst[i] = table[src[i]] for i < n, all device arrays: turns the int32 dictionary indices of a VARCHAR key batch into
 * the operator-wide value ids of the batch's base strings (VectorHasher::makeValueIds per BASE value on the host,
 * velox/exec/VectorHasher.cpp:241-289, the per-row step here).  An index outside [0, table_len) latches an error. */
int vgpu_translate_i32(int32_t* dst, const int32_t* src, const int32_t* table, int32_t table_len, int64_t n, void* stream);

/* ---- FilterProject ----------------------------------------------------------------------- */

/* Validates and uploads a program (ExprSet construction; velox/exec/FilterProject.cpp:59-103). */
int vgpu_program_create(const vgpu_program_desc* desc, vgpu_program_t* out);
int vgpu_program_destroy(vgpu_program_t program);

/* One projected output column, written at the ORIGINAL row positions (rows that do not pass the
 * filter are left undefined) — Velox wraps the full-length result in a dictionary over
 * selectedIndices (velox/exec/Operator.cpp:195-228), so no value compaction is needed. */
typedef struct {
  void* values;    /* device, nrows values of the output operand's type (BOOLEAN: bit-packed)   */
  uint64_t* nulls; /* device bitmap, may be NULL when the caller knows the result has no nulls  */
} vgpu_out_column;

/* Evaluates the program over `nrows` rows.
 *   in_rows/n_in_rows : optional ascending input selection (the SelectivityVector of the batch as
 *                       an index list); NULL = all rows [0, nrows).
 *   sel_bits          : out, device bitmap of nrows bits — selectedBits of FilterEvalCtx.
 *   sel_idx           : out, device int32[>= nrows] ascending passing row numbers, may be NULL.
 *   num_selected      : out, device int64.
 *   outs              : n_outputs entries (may be NULL when the program has no outputs).
 * Asynchronous on `stream`; data errors are latched and reported by vgpu_check_errors(). */
int vgpu_filter_project(vgpu_program_t program, const vgpu_column* cols, int32_t ncols,
                        int64_t nrows, const int32_t* in_rows, int64_t n_in_rows,
                        uint64_t* sel_bits, int32_t* sel_idx, int64_t* num_selected,
                        const vgpu_out_column* outs, void* stream);

/* The last step of vgpu_filter_project on its own, with NO allocation: selection bitmap -> ascending row numbers.
 * (No counterpart in the reference: added for callers that cannot call
 * cudaMalloc — code under CUDA graph capture, tight serving loops, harnesses that forbid it.  vgpu_filter_project
 * keeps the per-tile counts / offsets this step needs in a grow-only cudaMalloc'ed buffer; this variant keeps them in
 * a fixed array in the library's device image instead.)
 *   sel_bits     : in, device bitmap of nrows bits (32-bit aligned); bits past nrows are ignored.
 *   nrows        : at most VGPU_NOALLOC_MAX_ROWS; more returns VGPU_ERR_UNSUPPORTED (split the batch).
 *   sel_idx      : out, device int32[>= number of set bits] ascending row numbers of the set bits.
 *   num_selected : out, device int64.
 * Asynchronous on `stream`.  Calls no cudaMalloc, cudaFree or other allocating routine, ever.  The scratch (192 KiB) is
 * held by every GPU that loads the library until the process exits, is shared by all calls on a device and keeps the
 * previous call's contents (each call rewrites what it reads; nothing needs clearing): calls that may overlap on
 * the GPU (different streams without an ordering between them) race on it and must be ordered by the caller. */
#define VGPU_NOALLOC_MAX_ROWS (16ll * 1024 * 1024)
int vgpu_selection_indices_noalloc(const uint64_t* sel_bits, int64_t nrows, int32_t* sel_idx, int64_t* num_selected,
                                   void* stream);

/* Synchronises `stream` and returns VGPU_ERR_USER (with the Velox-style message) when a kernel
 * launched through this library on the current device latched a data error since the last check
 * (FilterProject calls and every aggregation; see vgpu_agg_check_errors for one aggregation). */
int vgpu_check_errors(void* stream);

/* ---- HashAggregation ---------------------------------------------------------------------- */

/* Device accumulator kinds.  Raw-input and intermediate-input steps of every Presto aggregate on
 * the path decompose into these (AVG raw = SUM_F64 + COUNT; AVG/COUNT final = SUM_F64 + SUM_I64
 * over the intermediate columns; velox/functions/lib/aggregates/AverageAggregateBase.h:65-390,
 * velox/functions/prestosql/aggregates/CountAggregate.cpp:81-90). */
typedef enum {
  VGPU_ACC_SUM_F64 = 0,  /* double += (double)x, nulls ignored                                  */
  VGPU_ACC_SUM_I64 = 1,  /* checked int64 += x (SumAggregateBase.h:150-158)                     */
  VGPU_ACC_COUNT = 2,    /* += 1 per passing row whose input is not null; input NONE = count(*) */
  VGPU_ACC_MIN = 3,      /* typed by the input operand (integers, double, or BOOLEAN as 0 / 1: bool_and) */
  VGPU_ACC_MAX = 4
} vgpu_acc_kind;

typedef struct {
  int32_t kind;  /* vgpu_acc_kind */
  int32_t input; /* index into the program's outputs, or -1 (count(*))                          */
} vgpu_acc_spec;

/* How one grouping key becomes bits of the 64-bit normalized key
 * (VectorHasher value ids / kNormalizedKey, velox/exec/VectorHasher.h:118-578,
 *  velox/exec/HashTable.cpp:1376-1461).  id = value - min + 1, 0 = NULL; `bits` wide. */
typedef struct {
  int32_t output; /* index into the program's outputs                                          */
  int32_t bits;   /* width of the id field; 64 = use the raw value (single BIGINT key)          */
  int64_t min;    /* subtracted before packing (range mode)                                    */
  int64_t max;    /* values outside [min,max] latch "key out of range" for a host redo.
                     bits == 64: optional range statistics of the key (min < max; 0,0 = unknown,
                     the VectorHasher min/max of the reference).  A hint: results never depend on
                     it.  With VGPU_RANGED_KEYS=1 keys inside the range are placed monotonically
                     in the HBM table (experimental; hashing is the default and measured faster). */
} vgpu_key_spec;

typedef enum {
  VGPU_AGG_GLOBAL = 0, /* no keys: one group, present even with no input                       */
  VGPU_AGG_ARRAY = 1,  /* product of key cardinalities small: per-thread shared-memory partial
                          tables, merged per block (kArray analogue)                           */
  VGPU_AGG_HASH = 2    /* open-addressing table in HBM keyed by the normalized key             */
} vgpu_agg_mode;

typedef struct {
  int32_t mode; /* vgpu_agg_mode */
  int32_t n_keys;
  const vgpu_key_spec* keys;
  int32_t n_accs;
  const vgpu_acc_spec* accs;
  int32_t ignore_null_keys;   /* drop rows with a NULL key (velox/exec/OperatorUtils.cpp:141-157) */
  int32_t reserved;
  int64_t expected_groups;    /* HASH: initial capacity hint (0 = default)                       */
} vgpu_agg_desc;

typedef struct vgpu_agg_s* vgpu_agg_t;

int vgpu_agg_create(const vgpu_agg_desc* desc, vgpu_agg_t* out);
int vgpu_agg_destroy(vgpu_agg_t agg);
/* Drops all groups, keeps the configuration (partial-aggregation flush). */
int vgpu_agg_reset(vgpu_agg_t agg, void* stream);

/* Fused filter -> project -> accumulate over one batch: evaluates `program` (filter + outputs)
 * and folds passing rows into the groups (GroupingSet::addInput).  Never materialises the
 * selection or the projections. */
int vgpu_agg_add_input(vgpu_agg_t agg, vgpu_program_t program, const vgpu_column* cols,
                       int32_t ncols, int64_t nrows, const int32_t* in_rows, int64_t n_in_rows,
                       void* stream);

/* End of input (HashAggregation::noMoreInput, velox/exec/HashAggregation.cpp:466-472): completes the device-side
 * aggregation asynchronously on `stream`.  HASH aggregations log run-aggregated {key, accumulators} entries while
 * scanning and deduplicate them here (radix partition + shared-memory tables); the other modes have nothing left to
 * do.  Implied by vgpu_agg_num_groups / vgpu_agg_extract / vgpu_agg_partition; more input may follow. */
int vgpu_agg_finalize(vgpu_agg_t agg, void* stream);

/* Like vgpu_check_errors, for the data errors (sum overflow, checked arithmetic in fused projections, key domain)
 * latched by this aggregation's kernels only: with N drivers per device every operator instance raises its own
 * errors, as exceptions thrown out of HashAggregation::addInput do (velox/exec/Driver.cpp:536-586). */
int vgpu_agg_check_errors(vgpu_agg_t agg, void* stream);

/* Device bytes the aggregation's group state holds right now (a host-side upper bound: no synchronisation) — what a
 * partial HashAggregation compares with max_partial_aggregation_memory before it flushes
 * (velox/exec/HashAggregation.cpp:326-348). */
int vgpu_agg_memory_bytes(vgpu_agg_t agg, int64_t* bytes);

/* Number of groups so far (synchronises the stream). */
int vgpu_agg_num_groups(vgpu_agg_t agg, void* stream, int64_t* num_groups);

/* Row width of vgpu_agg_extract output: 8 bytes normalized key, 8 bytes flags (bit 0: NULL key), 8 bytes rows, then
 * 8 bytes per accumulator (SUM_F64 as double; everything else int64, MIN/MAX of doubles as double), then 8 bytes per
 * accumulator of "non-null input count" — a group's SUM/MIN/MAX is NULL when that count is 0
 * (velox/exec/Aggregate.h:305-337).  Only zero / non-zero of that count (and of `rows`) is specified: the engines
 * count exactly where a COUNT accumulator needs the number and otherwise may report 1 for "some". */
int vgpu_agg_row_bytes(vgpu_agg_t agg, int32_t* bytes);

/* Writes up to max_groups packed group rows into device memory `rows`, starting at group number
 * `offset` of an arbitrary but stable enumeration; *n_written (host) receives the count.
 * Synchronises the stream.  (GroupingSet::getOutput / RowContainer::listRows.) */
int vgpu_agg_extract(vgpu_agg_t agg, int64_t offset, int64_t max_groups, void* rows,
                     int64_t* n_written, void* stream);

/* Folds packed group rows (same layout, e.g. received from another GPU) into the groups:
 * sums and counts add, MIN/MAX combine (Aggregate::addIntermediateResults). */
int vgpu_agg_merge_rows(vgpu_agg_t agg, const void* rows, int64_t n_rows, void* stream);

/* Hash-partitions the current groups into n_parts contiguous segments of packed rows:
 * part = mix64(normalized key) % n_parts.  rows_out must hold num_groups rows; counts_out is a
 * HOST array of n_parts int64 (segment lengths, in order).  Synchronises the stream. */
int vgpu_agg_partition(vgpu_agg_t agg, int32_t n_parts, void* rows_out, int64_t* counts_out,
                       void* stream);

/* ---- multi-GPU exchange (one process per GPU) ------------------------------------------- */

/* Partition (destination rank) of a group between the partial and the final aggregation: hash % n_parts
 * (velox/exec/HashPartitionFunction.cpp:59-92).  raw_bigint_key != 0 — the grouping key is ONE BIGINT column (bits = 64):
 * the hash is VectorHasher::hash's, folly::hasher<int64_t> = twang_mix64 per value and kNullHash = 1 for NULL
 * (velox/exec/VectorHasher.cpp:68-108, velox/vector/BaseVector.h:74), so a GPU partial stage routes every group where a
 * CPU Velox PartitionedOutput would.  Otherwise `key` is a packed normalized key / dense group number, which the
 * reference never hashes as such; a fixed splitmix64 is used.  Host function, no device needed. */
int vgpu_partition_of(int64_t key, int32_t is_null, int32_t n_parts, int32_t raw_bigint_key, int32_t* part);

#define VGPU_COMM_ID_BYTES 128
typedef struct vgpu_comm_s* vgpu_comm_t;
int vgpu_comm_unique_id(uint8_t id[VGPU_COMM_ID_BYTES]); /* rank 0; broadcast out of band */
int vgpu_comm_create(const uint8_t id[VGPU_COMM_ID_BYTES], int32_t rank, int32_t world,
                     vgpu_comm_t* out);
int vgpu_comm_destroy(vgpu_comm_t comm);
/* Variable-size all-to-all of bytes between device buffers over NCCL (grouped send/recv).
 * send_counts/recv_counts/send_offsets/recv_offsets: HOST arrays of `world` byte counts. */
int vgpu_comm_alltoallv(vgpu_comm_t comm, const void* send, const int64_t* send_counts,
                        const int64_t* send_offsets, void* recv, const int64_t* recv_counts,
                        const int64_t* recv_offsets, void* stream);
/* All-gather of `bytes` per rank between device buffers (used for segment counts). */
int vgpu_comm_allgather(vgpu_comm_t comm, const void* send, void* recv, size_t bytes,
                        void* stream);

/* Peer-mapped window (collective; every rank passes the same size): each rank allocates `bytes` of device memory,
 * publishes a CUDA IPC handle and maps every peer's window, so that exchange kernels store their output straight into
 * the destination GPU's memory over NVLink / NVSwitch.  Three parts serve successive exchanges in turn (an exchange's
 * tail — count all-gather, fold, deduplication — runs on a side stream while the next scan already stores into the next
 * part); each part holds one region per source rank.  For a HASH exchange of E entries of w words per rank,
 * 3 * 1.3 * E * w * 8 bytes is ample.
 * *mapped (may be NULL) = 1 when every rank mapped every peer; otherwise the window is dropped on all ranks and
 * vgpu_comm_exchange_groups moves the payload with NCCL (the same result). */
int vgpu_comm_reserve_window(vgpu_comm_t comm, size_t bytes, int32_t* mapped);

/* Count matrix of the last window exchange of a HASH aggregation: matrix[src * world + dst] = entries src stored into
 * dst's window (HOST array of world * world int64).  Synchronises the stream. */
int vgpu_comm_last_counts(vgpu_comm_t comm, int64_t* matrix, void* stream);

/* The whole exchange step between a partial and a final aggregation of the same description
 * (partialAggregation -> partitionedOutput -> exchange -> finalAggregation,
 * velox/exec/tests/MultiFragmentTest.cpp:180-212): group g goes to rank mix64(key) % world.
 * With a mapped window (vgpu_comm_reserve_window) and no host synchronisation anywhere:
 *   GLOBAL / ARRAY aggregations (small key domains): one kernel stores this rank's dense group state into every
 *     peer's window, raises an epoch flag, waits for the peers' flags and folds the groups this rank owns.
 *   HASH aggregations (log engine): one kernel hash-partitions the partial stage's logged {key, accumulators}
 *     entries and stores each destination's share into that peer's window; the per-destination counts travel in one
 *     small ncclAllGather (the only inter-rank synchronisation); the final aggregation deduplicates what it received
 *     (and is finalized on return, asynchronously).
 * Without a window: dense ncclAllGather / vgpu_agg_partition + counts all-gather + vgpu_comm_alltoallv +
 * vgpu_agg_merge_rows (one host synchronisation for the counts).
 * `final_agg` must have seen an (empty) vgpu_agg_add_input and be reset before the exchange after next.
 * *bytes_sent (host, may be NULL) receives the payload bytes this rank sent to other ranks, or -1 when only the
 * device knows (HASH through the window: see vgpu_comm_last_counts). */
int vgpu_comm_exchange_groups(vgpu_comm_t comm, vgpu_agg_t partial_agg, vgpu_agg_t final_agg, void* stream,
                              int64_t* bytes_sent);

/* ---- TableScan over the TPC-H connector, on the device -------------------------------------- */

/* lineitem columns in HBM (device pointers; any may be NULL = not wanted).  Widths are Velox's: BIGINT / DOUBLE 8 bytes,
 * INTEGER / DATE 4 bytes.  The four VARCHAR columns are produced as int32 ids into the sorted dictionaries that
 * vgpu_tpch_dictionary returns (what a DictionaryVector<StringView> over those bases carries). */
typedef struct {
  int64_t* l_orderkey;
  int64_t* l_partkey;
  int64_t* l_suppkey;
  int32_t* l_linenumber;
  double* l_quantity;
  double* l_extendedprice;
  double* l_discount;
  double* l_tax;
  int32_t* l_returnflag;   /* ids of {"A","N","R"} */
  int32_t* l_linestatus;   /* ids of {"F","O"} */
  int32_t* l_shipdate;
  int32_t* l_commitdate;
  int32_t* l_receiptdate;
  int32_t* l_shipinstruct; /* ids of {"COLLECT COD","DELIVER IN PERSON","NONE","TAKE BACK RETURN"} */
  int32_t* l_shipmode;     /* ids of {"AIR","FOB","MAIL","RAIL","REG AIR","SHIP","TRUCK"} */
} vgpu_lineitem_columns;

/* Lineitem rows of the orders [order_offset, order_offset + max_orders) at scale factor sf — the split arithmetic of
 * TpchDataSource::next / genTpchLineItem (velox/connectors/tpch/TpchConnector.cpp:139-163,
 * velox/tpch/gen/TpchGen.cpp:401-489: offsets and sizes count ORDERS).  Synchronises the stream. */
int vgpu_tpch_lineitem_rows(double sf, int64_t order_offset, int64_t max_orders, int64_t* rows, void* stream);

/* Generates those rows straight into HBM, bit-identical to the reference's dbgen (velox/tpch/gen/dbgen/build.cpp:110-223
 * mk_order with the seeds of dss.h:491-540; DOUBLE columns are integer * 0.01 like TpchGen.cpp:79-81; dates are days
 * since the epoch).  `capacity_rows`: rows the output arrays hold (7 * max_orders always suffices).  *rows (host, may
 * be NULL) receives the row count; passing it synchronises the stream, NULL keeps the call asynchronous. */
int vgpu_tpch_gen_lineitem(double sf, int64_t order_offset, int64_t max_orders, const vgpu_lineitem_columns* out,
                           int64_t capacity_rows, int64_t* rows, void* stream);

/* Dictionary of a VARCHAR column: which = 0 l_returnflag, 1 l_linestatus, 2 l_shipinstruct, 3 l_shipmode.
 * Returns the number of entries (-1: bad argument) and stores up to `cap` static strings in out (may be NULL). */
int vgpu_tpch_dictionary(int32_t which, const char** out, int32_t cap);

/* ---- statistics -------------------------------------------------------------------------- */

typedef struct {
  int64_t kernel_launches; /* kernels launched by this library on the calling thread's device   */
  int64_t rows_in;
  int64_t bytes_h2d;
  int64_t bytes_d2h;
} vgpu_stats;
int vgpu_get_stats(vgpu_stats* out);
int vgpu_reset_stats(void);

/* ---- diagnostics of the plan-specialised kernels ---------------------------------------------- */

/* Dry run (no device needed): vgpu_agg_add_input / vgpu_filter_project only generate the source
 * of the plan-specialised kernel for the shape they are given.  build() uses this to prebuild the
 * cubins of velox-1_b200/aot_plans.py. */
int vgpu_debug_set_dry_run(int on);
/* Source body of the last kernel generated on this thread and the hash that names its cubin.
 * Returns VGPU_ERR_INVALID when `cap` is too small (the text is truncated). */
int vgpu_debug_last_kernel_source(char* buf, size_t cap, uint64_t* hash);
/* How the last shape was handled on this thread: "built from aot|disk|nvrtc", "declined: ...",
 * "nvrtc failed: ...". */
int vgpu_debug_last_kernel_note(char* buf, size_t cap);

/* State of a HASH aggregation's engine, for tests and bench lines: info[0] engine (0 undecided, 1 log, 2 table),
 * and for a finalized log engine [1] sorted input detected, [2] front entries (whole groups when sorted),
 * [3] entries deduplicated, [4] groups the deduplication produced, [5] hash partitions, [6] 8-byte words per entry,
 * [7] host-side upper bound of the logged entries.  Synchronises the stream. */
int vgpu_debug_agg_info(vgpu_agg_t agg, void* stream, int64_t info[8]);

/* Every rank of a `world`-rank window exchange emulated on the CURRENT device (rank r: partials[r] -> finals[r]); for
 * boxes with fewer GPUs than ranks.  Runs the kernels of vgpu_comm_exchange_groups' window path (partition + store,
 * dedicated-group slots, receiver-side deduplication; dense store / fold for GLOBAL and ARRAY) with local buffers as
 * windows and a device copy in place of the count all-gather.  The finals hold extracted groups on return. */
int vgpu_debug_exchange_local(vgpu_agg_t* partials, vgpu_agg_t* finals, int32_t world, size_t window_bytes, void* stream);

/* Fuses the exchange into the partial aggregation's scan (HASH aggregations; replaces the Exchange between the
 * partial and final HashAggregation of velox/exec/HashPartitionFunction.cpp:59-92 + LocalPartition / PartitionedOutput
 * for this path): after this call the kernels of vgpu_agg_add_input(partial, ...) store every group entry straight into
 * the window of the rank that owns its key while they scan, so the NVLink transfer overlaps the scan; the following
 * vgpu_comm_exchange_groups(comm, partial, final) only exchanges the counts and deduplicates.  Needs a mapped window
 * (vgpu_comm_reserve_window) and at most 8 ranks; where that does not hold, or for batches whose shape the streaming
 * kernel does not cover, the entries are logged locally and partitioned by vgpu_comm_exchange_groups as before.
 * Entries that may continue in another warp's rows or another batch (two per warp and batch at most) are stored one by
 * one into a reserve of 256 entries per sending CTA and destination; an aggregation fed by hundreds of small batches
 * before one exchange can exhaust it (reported by vgpu_agg_check_errors as a full window region) and should not attach.
 * No other exchange may run through `comm` between the partial aggregation's first batch and its exchange.
 * comm == NULL detaches.  The attachment survives vgpu_agg_reset. */
int vgpu_agg_attach_exchange(vgpu_agg_t partial_agg, vgpu_comm_t comm);

/* Test support: `world` communicators whose ranks all live on the CURRENT device (windows are plain local
 * allocations, no NCCL), for vgpu_agg_attach_exchange + vgpu_debug_exchange_local_comms on boxes with fewer GPUs than
 * ranks.  Destroy each with vgpu_comm_destroy. */
int vgpu_debug_local_comms(int32_t world, size_t window_bytes, vgpu_comm_t* out);
/* vgpu_comm_exchange_groups for all ranks of such a local group at once (the count all-gather is a device copy). */
int vgpu_debug_exchange_local_comms(vgpu_comm_t* comms, vgpu_agg_t* partials, vgpu_agg_t* finals, int32_t world, void* stream);

/* Timing of the scan kernels (the fused / interpreter kernel of vgpu_agg_add_input and
 * vgpu_filter_project, not the small bookkeeping kernels): while on, CUDA events are recorded on
 * the launching stream right before and after each such kernel.  vgpu_debug_kernel_time waits for
 * them, returns the summed duration and the number of launches, and clears the list. */
int vgpu_debug_kernel_timing(int on);
int vgpu_debug_kernel_time(double* ms_total, int64_t* launches);

#ifdef __cplusplus
}
#endif
#endif /* VGPU_H_ */
