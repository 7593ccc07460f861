// Host-side step driver: ONE native call per host-fed training step instead of ~10 Python-level CUDA calls.
//
// A step through the public API (Trainer.train_packed) is: pinned H2D copy of the batch into an input slot on the copy
// stream -> the compute stream waits for it -> replay of the slot's CUDA graph -> D2H of the step's loss scalars. Issued
// from Python (stream contexts, Event objects, Tensor.copy_, graph.replay) that is 33 us of host time per step; issued
// here it is 15-17 us (bench.py `host_issue_us_per_step`, one B200 box), which matters when 8 ranks and their loader
// threads share the host cores (this is the tf.data / Estimator C++ run-loop role, N4). It does NOT change the GPU-side
// rate: the 10 us/step that end-to-end lost against the device-timed loop was the D2H copy op between two graph launches
// (engine hand-offs), removed by letting the optimizer kernel store its result into a pinned host ring
// (dfm_flat_opt_step_pub): e2e 9.98 -> 10.77 M samples/s.
//
// Slot protocol: input slots are double-buffered. `done[s]` marks the end of the last step that READ slot s (the next
// copy into s waits for it on the copy stream); `copied[s]` marks the end of the last copy INTO slot s (the step waits for
// it on the compute stream, and the caller may reuse its pinned source buffer once it has completed).
#include <cuda_runtime.h>

#include "common.cuh"

namespace {

constexpr int kMaxSlots = 8;

struct StepDriver {
  int nslots = 0;
  cudaEvent_t copied[kMaxSlots] = {};
  cudaEvent_t done[kMaxSlots] = {};
  bool have_done[kMaxSlots] = {};
  bool have_copy[kMaxSlots] = {};
  unsigned long long steps = 0;
};

}  // namespace

extern "C" {

int dfm_driver_create(void** handle, int nslots) {
  DFM_REQUIRE(handle != nullptr, "driver_create: null handle pointer");
  DFM_REQUIRE(nslots >= 1 && nslots <= kMaxSlots, "driver_create: nslots must be in [1, %d]", kMaxSlots);
  StepDriver* d = new StepDriver();
  d->nslots = nslots;
  for (int s = 0; s < nslots; ++s) {
    cudaError_t e = cudaEventCreateWithFlags(&d->copied[s], cudaEventDisableTiming);
    if (e == cudaSuccess) e = cudaEventCreateWithFlags(&d->done[s], cudaEventDisableTiming);
    if (e != cudaSuccess) {
      cudaGetLastError();
      dfm_set_error("driver_create: cudaEventCreate failed: %s", cudaGetErrorString(e));
      delete d;  // events created so far are released with the context; creation failing here means no usable context
      return DFM_ERR_CUDA;
    }
  }
  *handle = d;
  return DFM_OK;
}

int dfm_driver_destroy(void* handle) {
  StepDriver* d = (StepDriver*)handle;
  if (!d) return DFM_OK;
  for (int s = 0; s < d->nslots; ++s) {
    if (d->copied[s]) cudaEventDestroy(d->copied[s]);
    if (d->done[s]) cudaEventDestroy(d->done[s]);
  }
  cudaGetLastError();
  delete d;
  return DFM_OK;
}

// One host-fed step. host_src must be page-locked (the copy is asynchronous only then); graph_exec is the
// cudaGraphExec_t of the step captured for `slot` (it reads dev_dst); dev_scalars -> host_scalars is the step's result
// read-back (scalar_bytes may be 0). Nothing here blocks the calling thread.
int dfm_driver_step(void* handle, int slot, const void* host_src, void* dev_dst, unsigned long long bytes,
                    void* graph_exec, void* copy_stream, void* compute_stream, const void* dev_scalars,
                    void* host_scalars, unsigned long long scalar_bytes) {
  StepDriver* d = (StepDriver*)handle;
  DFM_REQUIRE(d != nullptr && slot >= 0 && slot < d->nslots, "driver_step: bad handle or slot %d", slot);
  DFM_REQUIRE(host_src && dev_dst && graph_exec && bytes > 0, "driver_step: null argument");
  DFM_REQUIRE(copy_stream != compute_stream, "driver_step: copy and compute streams must differ");
  cudaStream_t cs = (cudaStream_t)copy_stream, ms = (cudaStream_t)compute_stream;
  if (d->have_done[slot]) DFM_CHECK(cudaStreamWaitEvent(cs, d->done[slot], 0));  // last reader of the slot finished
  DFM_CHECK(cudaMemcpyAsync(dev_dst, host_src, bytes, cudaMemcpyHostToDevice, cs));
  DFM_CHECK(cudaEventRecord(d->copied[slot], cs));
  d->have_copy[slot] = true;
  DFM_CHECK(cudaStreamWaitEvent(ms, d->copied[slot], 0));
  DFM_CHECK(cudaGraphLaunch((cudaGraphExec_t)graph_exec, ms));
  DFM_CHECK(cudaEventRecord(d->done[slot], ms));
  d->have_done[slot] = true;
  if (scalar_bytes)
    DFM_CHECK(cudaMemcpyAsync(host_scalars, dev_scalars, scalar_bytes, cudaMemcpyDeviceToHost, ms));
  ++d->steps;
  return DFM_OK;
}

// 1 = the last copy into `slot` has completed (its pinned source buffer may be reused), 0 = still in flight, <0 = error.
int dfm_driver_copy_done(void* handle, int slot) {
  StepDriver* d = (StepDriver*)handle;
  if (!d || slot < 0 || slot >= d->nslots) return -1;
  if (!d->have_copy[slot]) return 1;
  cudaError_t e = cudaEventQuery(d->copied[slot]);
  if (e == cudaSuccess) return 1;
  if (e == cudaErrorNotReady) { cudaGetLastError(); return 0; }
  dfm_set_error("driver_copy_done: %s", cudaGetErrorString(e));
  return -1;
}

int dfm_driver_copy_wait(void* handle, int slot) {
  StepDriver* d = (StepDriver*)handle;
  DFM_REQUIRE(d != nullptr && slot >= 0 && slot < d->nslots, "driver_copy_wait: bad handle or slot %d", slot);
  if (d->have_copy[slot]) DFM_CHECK(cudaEventSynchronize(d->copied[slot]));
  return DFM_OK;
}

unsigned long long dfm_driver_steps(void* handle) { return handle ? ((StepDriver*)handle)->steps : 0ull; }

}  // extern "C"
