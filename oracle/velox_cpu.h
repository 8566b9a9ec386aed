/* TEST INFRASTRUCTURE — C ABI of the CPU restatement of the reference's FilterProject ->
 * HashAggregation path (oracle/velox_cpu.cpp).  It is the timed CPU baseline (bench.py
 * --impl reference / cpu_baseline) and a second checker beside oracle/np_oracle.py.  Nothing under
 * velox-1_b200/ includes or links this. */
#ifndef VELOX_CPU_ORACLE_H_
#define VELOX_CPU_ORACLE_H_
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

/* Type codes = velox TypeKind (velox/type/Type.h:62-84): 0 BOOLEAN(1 byte per row here),
 * 3 INTEGER, 4 BIGINT, 6 DOUBLE. */
typedef struct {
  int32_t type;
  int32_t pad;
  const void* values;     /* flat values */
  const uint8_t* nulls;   /* optional byte per row, 1 = NULL (already decoded) */
} vcpu_column;

/* kind: 0 sum(double) 1 sum(bigint, checked) 2 count 3 min 4 max; input = output index or -1 */
typedef struct {
  int32_t kind;
  int32_t input;
} vcpu_acc;

/* Runs filter -> project -> partial aggregation on `threads` driver threads (contiguous row
 * ranges, batches of `batch_rows`), then one final merge
 * (partialAggregation -> localPartition -> finalAggregation, TpchQueryBuilder.cpp:220-232).
 * Expressions are s-expressions: (fn arg...) | cN:T | literal:T | null:T with T in b,i32,i64,f64
 * and fn in plus minus multiply divide eq neq lt lte gt gte between and or not is_null.
 * Result rows (int64 words): n_keys key values, n_keys null flags, then per accumulator
 * {value bits, non-null count}.  Returns the number of groups, or -1 (message in vcpu_error()).
 * *seconds receives the wall time of the operator pipeline (excluding parsing). */
int64_t vcpu_run(const char* filter, const char* const* outputs, int32_t n_outputs,
                 const vcpu_column* cols, int32_t n_cols, int64_t n_rows,
                 const int32_t* key_outputs, int32_t n_keys, const vcpu_acc* accs, int32_t n_accs,
                 int32_t ignore_null_keys, int32_t threads, int32_t batch_rows,
                 int64_t* out_rows, int64_t out_cap_rows, double* seconds);
const char* vcpu_error(void);

#ifdef __cplusplus
}
#endif
#endif
