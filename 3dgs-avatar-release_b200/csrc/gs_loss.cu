// gs_loss.cu -- photometric L1 loss of the training step (utils/loss_utils.py:19-20 `l1_loss`, used at train.py:121),
// forward and the per-pixel gradient in ONE pass over the two images: the PyTorch chain (sub, abs, mean; backward: fill,
// expand/div, sign, mul) is seven launches and six image-sized round trips through HBM per step.  The sum is
// deterministic: per-block partials, the last block to finish adds them in block order.
#include <atomic>

#include "gs_common.cuh"

namespace {

#define L1_THREADS 256

__global__ void __launch_bounds__(L1_THREADS)
l1_loss_kernel(const float* __restrict__ img, const float* __restrict__ gt, long long n, float inv_n,
               float* __restrict__ loss, float* __restrict__ dL_dimg, float* __restrict__ partial,
               unsigned int* __restrict__ ticket) {
  GS_PDL_PROLOGUE();
  __shared__ float s_w[L1_THREADS / 32];
  __shared__ bool s_last;
  const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
  float acc = 0.f;
  const long long n4 = n >> 2, stride = (long long)gridDim.x * L1_THREADS;
  for (long long i = (long long)blockIdx.x * L1_THREADS + tid; i < n4; i += stride) {
    const float4 a = reinterpret_cast<const float4*>(img)[i], b = reinterpret_cast<const float4*>(gt)[i];
    const float d0 = a.x - b.x, d1 = a.y - b.y, d2 = a.z - b.z, d3 = a.w - b.w;
    acc += (fabsf(d0) + fabsf(d1)) + (fabsf(d2) + fabsf(d3));
    float4 g;  // d|x|/dx = sign(x), sign(0) = 0 (torch.sign)
    g.x = d0 > 0.f ? inv_n : (d0 < 0.f ? -inv_n : 0.f);
    g.y = d1 > 0.f ? inv_n : (d1 < 0.f ? -inv_n : 0.f);
    g.z = d2 > 0.f ? inv_n : (d2 < 0.f ? -inv_n : 0.f);
    g.w = d3 > 0.f ? inv_n : (d3 < 0.f ? -inv_n : 0.f);
    reinterpret_cast<float4*>(dL_dimg)[i] = g;
  }
  for (long long i = (n4 << 2) + (long long)blockIdx.x * L1_THREADS + tid; i < n; i += stride) {  // tail
    const float d = img[i] - gt[i];
    acc += fabsf(d);
    dL_dimg[i] = d > 0.f ? inv_n : (d < 0.f ? -inv_n : 0.f);
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xFFFFFFFFu, acc, o);
  if (lane == 0) s_w[w] = acc;
  __syncthreads();
  if (tid == 0) {
    float t = 0.f;
#pragma unroll
    for (int k = 0; k < L1_THREADS / 32; k++) t += s_w[k];
    partial[blockIdx.x] = t;
    __threadfence();
    s_last = atomicAdd(ticket, 1u) == gridDim.x - 1;
  }
  __syncthreads();
  if (s_last && w == 0) {
    __threadfence();
    float t = 0.f;
    for (unsigned b = lane; b < gridDim.x; b += 32) t += __ldcg(partial + b);  // fixed assignment, fixed order
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) t += __shfl_xor_sync(0xFFFFFFFFu, t, o);
    if (lane == 0) { *loss = t * inv_n; *ticket = 0u; }  // the ticket is left at zero for the next call
  }
}

__global__ void __launch_bounds__(L1_THREADS)
scale_kernel(const float* __restrict__ src, const float* __restrict__ scale, long long n, float* __restrict__ dst) {
  GS_PDL_PROLOGUE();
  const float s = *scale;
  const long long n4 = n >> 2, stride = (long long)gridDim.x * L1_THREADS;
  for (long long i = (long long)blockIdx.x * L1_THREADS + threadIdx.x; i < n4; i += stride) {
    float4 v = reinterpret_cast<const float4*>(src)[i];
    v.x *= s; v.y *= s; v.z *= s; v.w *= s;
    reinterpret_cast<float4*>(dst)[i] = v;
  }
  for (long long i = (n4 << 2) + (long long)blockIdx.x * L1_THREADS + threadIdx.x; i < n; i += stride) dst[i] = src[i] * s;
}

int l1_blocks(long long n) {
  long long b = (n / 4 + L1_THREADS - 1) / L1_THREADS;
  if (b > GS_L1_MAX_BLOCKS) b = GS_L1_MAX_BLOCKS;
  return b < 1 ? 1 : (int)b;
}

}  // namespace

extern "C" {

size_t l1_loss_workspace_bytes(void) { return (GS_L1_MAX_BLOCKS + 1) * sizeof(float); }

int l1_loss_forward(const float* image, const float* target, int64_t n, float* loss, float* dL_dimage, void* workspace,
                    size_t workspace_bytes, void* stream) {
  if (!image || !target || !loss || !dL_dimage || !workspace || n <= 0) return GS_ERR_BAD_ARG;
  if (workspace_bytes < l1_loss_workspace_bytes()) return GS_ERR_WORKSPACE_SMALL;
  if (((uintptr_t)image | (uintptr_t)target | (uintptr_t)dL_dimage) & 15) return GS_ERR_BAD_ARG;
  float* partial = (float*)workspace;
  unsigned int* ticket = (unsigned int*)(partial + GS_L1_MAX_BLOCKS);  // zero on first use (caller), self-resetting after
  gs_launch(l1_loss_kernel, dim3(l1_blocks(n)), dim3(L1_THREADS), 0, (cudaStream_t)stream, image, target, (long long)n,
            1.0f / (float)n, loss, dL_dimage, partial, ticket);
  return gs_check(cudaGetLastError());
}

/* No-allocation variant of l1_loss_forward (no counterpart in the reference; added for
 * callers that cannot call cudaMalloc: code under CUDA graph capture, tight serving loops, harnesses that forbid it).  The
 * kernel's scratch -- one partial sum per block + the completion ticket -- lives in the file-scope device array below instead
 * of a caller workspace: (GS_L1_MAX_BLOCKS + 1) words = 2372 bytes, held by every GPU that loads the library until exit.
 * Limit: n <= L1_STATIC_MAX_N = 2^26 elements (a 3 x 4096 x 4096 image is 50.3 M); larger -> GS_ERR_STATIC_LIMIT.
 * Every call on a device shares the one array and finds the previous call's contents in it: the ticket word is cleared on the
 * stream before every launch (the partials are overwritten before they are read).  Calls that overlap on the GPU (two streams,
 * or two host threads on one device) RACE on it -- serialise them.  Calls cudaGetSymbolAddress once per device (no allocation),
 * cudaMemsetAsync and the kernel launch; no cudaMalloc / cudaFree / allocating library routine. */
}  // extern "C"

namespace {
constexpr long long L1_STATIC_MAX_N = 1ll << 26;
constexpr int L1_STATIC_MAX_DEV = 64;
static __device__ __align__(256) float g_l1_static_scratch[GS_L1_MAX_BLOCKS + 1];
// In HOST code the array's name is only a stand-in: the device address comes from cudaGetSymbolAddress, per device (each GPU
// that loaded the module holds its own copy), taken once with that device current.
std::atomic<float*> g_l1_static_ptr[L1_STATIC_MAX_DEV];

int l1_static_scratch(float** out) {
  int dev = -1;
  if (gs_check(cudaGetDevice(&dev)) != GS_OK) return GS_ERR_CUDA;
  if (dev < 0 || dev >= L1_STATIC_MAX_DEV) return GS_ERR_BAD_ARG;
  float* p = g_l1_static_ptr[dev].load(std::memory_order_acquire);
  if (!p) {
    if (gs_check(cudaGetSymbolAddress((void**)&p, g_l1_static_scratch)) != GS_OK || !p) return GS_ERR_CUDA;
    g_l1_static_ptr[dev].store(p, std::memory_order_release);  // racing first calls store the same address
  }
  *out = p;
  return GS_OK;
}
}  // namespace

extern "C" {

int64_t l1_loss_static_max_n(void) { return (int64_t)L1_STATIC_MAX_N; }

int l1_loss_forward_static(const float* image, const float* target, int64_t n, float* loss, float* dL_dimage, void* stream) {
  if (!image || !target || !loss || !dL_dimage || n <= 0) return GS_ERR_BAD_ARG;
  if (n > L1_STATIC_MAX_N) return GS_ERR_STATIC_LIMIT;
  if (((uintptr_t)image | (uintptr_t)target | (uintptr_t)dL_dimage) & 15) return GS_ERR_BAD_ARG;
  float* partial = nullptr;
  const int s = l1_static_scratch(&partial);
  if (s != GS_OK) return s;
  unsigned int* ticket = (unsigned int*)(partial + GS_L1_MAX_BLOCKS);
  cudaStream_t st = (cudaStream_t)stream;
  if (gs_check(cudaMemsetAsync(ticket, 0, sizeof(unsigned int), st)) != GS_OK) return GS_ERR_CUDA;  // must start at zero
  gs_launch(l1_loss_kernel, dim3(l1_blocks(n)), dim3(L1_THREADS), 0, st, image, target, (long long)n, 1.0f / (float)n, loss,
            dL_dimage, partial, ticket);
  return gs_check(cudaGetLastError());
}

int l1_loss_backward(const float* dL_dimage_unit, const float* dL_dloss, int64_t n, float* dL_dimage, void* stream) {
  if (!dL_dimage_unit || !dL_dloss || !dL_dimage || n <= 0) return GS_ERR_BAD_ARG;
  if (((uintptr_t)dL_dimage_unit | (uintptr_t)dL_dimage) & 15) return GS_ERR_BAD_ARG;
  gs_launch(scale_kernel, dim3(l1_blocks(n)), dim3(L1_THREADS), 0, (cudaStream_t)stream, dL_dimage_unit, dL_dloss,
            (long long)n, dL_dimage);
  return gs_check(cudaGetLastError());
}

}  // extern "C"
