// Plan-specialised kernels: generation (fused.cu) and compile / cache / launch (fused_rt.cu).
#pragma once
#include <string>

#include "internal.h"

namespace vg {

struct FusedPlan {
  int threads = 512;
  int smem_bytes = 0; // dynamic shared memory of the kernel
};

// Writes the kernel body for the pipeline described by `p` (after plan_pipeline and the sink
// fields are filled).  Returns false, with the reason, for shapes the generator declines.
bool fused_gen_source(const PipelineParams& p, FusedPlan& plan, std::string& source, std::string& why);

struct FusedKernel;

// Finds or builds the kernel for `p`: in-memory cache, then a prebuilt cubin (aot/ next to the
// library, then the user cache directory), then NVRTC.  Returns nullptr when the shape is declined
// or no compiler is available (the caller then runs the interpreter); *grid receives the launch
// width (CTAs), fixed per device and kernel so results are run-to-run deterministic.
FusedKernel* fused_prepare(const PipelineParams& p, int* grid);
int fused_launch(FusedKernel* k, const PipelineParams& p, int grid, cudaStream_t stream);
int fused_kernel_threads(const FusedKernel* k);

// Dry run (no device): vgpu_agg_add_input / vgpu_filter_project only generate the source, which
// fused_last_source() returns.  Used by build() to prebuild the cubins of aot_plans.py.
void fused_set_dry_run(bool on);
bool fused_dry_run();
// Small dictionaries (<= 32 base values) looked up with a warp shuffle instead of a gather (VGPU_SMALL_DICT=0 disables).
bool fused_small_dict_enabled();
// True once after a kernel failed to compile under VGPU_STRICT=1 (the error text is in vgpu_last_error).
bool fused_strict_failure();
const std::string& fused_last_source();
const std::string& fused_last_note(); // why the last shape was declined / how it was built
uint64_t fused_source_hash(const std::string& source);

} // namespace vg
