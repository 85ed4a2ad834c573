// noalloc.cu -- skrnn_clip_adam_noalloc: the global-norm clip + Adam update (reference
// model.py:182-184) for callers that own the parameter, gradient and Adam-slot buffers themselves
// and must not trigger a device allocation.
//
// Who needs this.  skrnn_apply_clip_adam works on a handle, and creating a handle allocates the
// weights, the optimiser slots and the activation workspaces with cudaMalloc.  A trainer whose
// tensors live in a framework memory pool (which typically has already claimed the whole GPU, so a
// raw cudaMalloc from a library either fails or fragments the pool), or an update step that runs
// in a context where allocating is forbidden outright, cannot afford that.  This entry point takes
// plain device pointers for everything it reads and writes and keeps its scratch memory -- the
// per-chunk partial sums of the squared gradient and the grid-barrier counter -- in one fixed
// array that is part of the loaded library image.
//
// Limits and sharing.  The scratch array is SKRNN_NOALLOC_SCRATCH_BYTES (64 KiB + 256 B) and every
// GPU that loads the library holds one copy until the process exits.  It stores one double per
// 4096-element chunk of the gradient, so at most SKRNN_NOALLOC_MAX_PARAMS = 8192 * 4096 parameters
// can be updated per call; larger counts are rejected.  All calls on a device share the array and
// it keeps whatever the previous call left: the barrier counter is cleared (stream-ordered) at the
// start of every call and every partial sum that is read is rewritten first, but two calls that
// overlap on the GPU (different streams, or different host threads) race on it -- serialise them,
// the first call included (its set-up below is plain unsynchronised bookkeeping).
//
// The first call on each device looks up the device address of the array (cudaGetSymbolAddress)
// and queries the device; make that first call before starting any stream capture.  The device is
// the one that owns params_dev, not whichever happens to be current.
#include "common.cuh"

namespace {

constexpr int NA_CHUNK = 4096;
constexpr int NA_MAX_CHUNKS = 8192;
constexpr int NA_HEADER = 256;  // barrier counter lives in the first 256 bytes
constexpr size_t NA_SCRATCH_BYTES = NA_HEADER + (size_t)NA_MAX_CHUNKS * sizeof(double);
constexpr int NA_MAX_DEVICES = 64;

// In host code this name is only a handle for cudaGetSymbolAddress, never a device pointer.
static __device__ __align__(256) unsigned char g_noalloc_scratch[NA_SCRATCH_BYTES];

struct NaDevice {
  bool ready = false;
  void *scratch = nullptr;
  int num_sms = 0;
  int blocks_per_sm = 0;
};
NaDevice g_na_dev[NA_MAX_DEVICES];

__global__ void __launch_bounds__(256)
clip_adam_noalloc_kernel(int64_t n, const float *__restrict__ g, float *__restrict__ p, float *__restrict__ m,
                         float *__restrict__ v, float alpha, float clip, float one_m_b1, float one_m_b2, float eps,
                         unsigned char *__restrict__ scratch, float *__restrict__ gnorm_out) {
  __shared__ double red[8];
  __shared__ double total_sh;
  unsigned int *ctr = reinterpret_cast<unsigned int *>(scratch);
  double *partial = reinterpret_cast<double *>(scratch + NA_HEADER);
  const int nchunks = (int)((n + NA_CHUNK - 1) / NA_CHUNK);
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;

  // phase 1: one partial sum of squares per fixed 4096-element chunk (fixed order => reproducible)
  for (int c = blockIdx.x; c < nchunks; c += gridDim.x) {
    const int64_t beg = (int64_t)c * NA_CHUNK;
    const int64_t end = beg + NA_CHUNK < n ? beg + NA_CHUNK : n;
    float ss = 0.f;
    for (int64_t i = beg + threadIdx.x; i < end; i += blockDim.x) {
      float x = g[i];
      ss = fmaf(x, x, ss);
    }
    double d = (double)ss;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) d += __shfl_xor_sync(0xffffffffu, d, o);
    if (lane == 0) red[wid] = d;
    __syncthreads();
    if (threadIdx.x == 0) {
      double t = 0.0;
      for (int i = 0; i < 8; ++i) t += red[i];
      partial[c] = t;
    }
    __syncthreads();
  }
  // grid barrier: every block is resident (cooperative launch); the counter was zeroed by the host
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence();
    atomicAdd(ctr, 1u);
    while (*((volatile unsigned int *)ctr) < gridDim.x) { }
    __threadfence();
  }
  __syncthreads();
  // phase 2: every block adds the partials in the same order, then updates its share
  {
    double d = 0.0;
    for (int c = threadIdx.x; c < nchunks; c += blockDim.x) d += __ldcg(partial + c);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) d += __shfl_xor_sync(0xffffffffu, d, o);
    if (lane == 0) red[wid] = d;
    __syncthreads();
    if (threadIdx.x == 0) {
      double t = 0.0;
      for (int i = 0; i < 8; ++i) t += red[i];
      total_sh = t;
    }
    __syncthreads();
  }
  const float gn = (float)sqrt(total_sh);
  const float scale = clip * fminf(1.0f / gn, 1.0f / clip);  // tf.clip_by_global_norm
  if (blockIdx.x == 0 && threadIdx.x == 0 && gnorm_out != nullptr) *gnorm_out = gn;
  const int64_t tid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const int64_t nth = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = tid; i < n; i += nth) {  // TF ApplyAdam
    float gi = g[i] * scale;
    float mi = m[i], vi = v[i];
    mi = mi + (gi - mi) * one_m_b1;
    vi = vi + (gi * gi - vi) * one_m_b2;
    m[i] = mi;
    v[i] = vi;
    p[i] = p[i] - alpha * mi / (sqrtf(vi) + eps);
  }
}

}  // namespace

extern "C" int skrnn_clip_adam_noalloc(int64_t n, const float *grads_dev, float *params_dev, float *m_dev,
                                       float *v_dev, float lr, float grad_clip, double beta1_power,
                                       double beta2_power, float *gnorm_out_dev, void *stream) {
  SK_CHECK(grads_dev && params_dev && m_dev && v_dev, "clip_adam_noalloc: null buffer");
  SK_CHECK(n > 0, "clip_adam_noalloc: parameter count %lld must be positive", (long long)n);
  SK_CHECK(n <= (int64_t)NA_MAX_CHUNKS * NA_CHUNK,
           "clip_adam_noalloc: %lld parameters exceed the fixed scratch capacity of %lld (SKRNN_NOALLOC_MAX_PARAMS); "
           "use skrnn_apply_clip_adam on a handle instead",
           (long long)n, (long long)NA_MAX_CHUNKS * NA_CHUNK);
  SK_CHECK(grad_clip > 0.f, "clip_adam_noalloc: grad_clip must be positive");
  SK_CHECK(beta1_power > 0.0 && beta1_power < 1.0 && beta2_power > 0.0 && beta2_power < 1.0,
           "clip_adam_noalloc: beta powers must lie in (0,1) (0.9 and 0.999 before the first step)");
  // run on the GPU that owns the buffers, whatever device happens to be current (handle entry points and other
  // libraries may have switched it); the caller's device is restored on return
  cudaPointerAttributes pat;
  SK_CUDA(cudaPointerGetAttributes(&pat, params_dev));
  SK_CHECK(pat.type == cudaMemoryTypeDevice || pat.type == cudaMemoryTypeManaged,
           "clip_adam_noalloc: params_dev is not a device pointer");
  const int dev = pat.device;
  SK_CHECK(dev >= 0 && dev < NA_MAX_DEVICES, "clip_adam_noalloc: device ordinal %d unsupported", dev);
  SkDeviceGuard guard(dev);
  NaDevice &d = g_na_dev[dev];
  if (!d.ready) {
    // once per device, with that device current: resolve the device address of the scratch array
    void *ptr = nullptr;
    SK_CUDA(cudaGetSymbolAddress(&ptr, g_noalloc_scratch));
    SK_CHECK(ptr != nullptr, "clip_adam_noalloc: scratch symbol has no device address");
    SK_CUDA(cudaDeviceGetAttribute(&d.num_sms, cudaDevAttrMultiProcessorCount, dev));
    SK_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&d.blocks_per_sm, clip_adam_noalloc_kernel, 256, 0));
    SK_CHECK(d.blocks_per_sm >= 1, "clip_adam_noalloc: kernel does not fit on device %d", dev);
    d.scratch = ptr;
    d.ready = true;
  }
  cudaStream_t st = (cudaStream_t)stream;
  // the array keeps the last call's contents: the barrier counter must start at zero
  SK_CUDA(cudaMemsetAsync(d.scratch, 0, NA_HEADER, st));
  int grid = d.num_sms * (d.blocks_per_sm < 4 ? d.blocks_per_sm : 4);
  const int64_t nchunks = (n + NA_CHUNK - 1) / NA_CHUNK;
  if (nchunks < grid) grid = (int)nchunks;
  float alpha = (float)((double)lr * sqrt(1.0 - beta2_power) / (1.0 - beta1_power));
  float omb1 = 1.0f - 0.9f, omb2 = 1.0f - 0.999f, eps = 1e-3f;
  unsigned char *scratch = static_cast<unsigned char *>(d.scratch);
  void *args[] = {&n, &grads_dev, &params_dev, &m_dev, &v_dev, &alpha, &grad_clip, &omb1, &omb2, &eps, &scratch,
                  &gnorm_out_dev};
  SK_CUDA(cudaLaunchCooperativeKernel((void *)clip_adam_noalloc_kernel, dim3(grid), dim3(256), args, 0, st));
  return 0;
}
