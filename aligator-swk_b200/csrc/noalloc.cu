// noalloc.cu — garb200_kkt_error_noalloc: lqrComputeKktError without a solver handle and without
// ANY allocation (no cudaMalloc / cudaFree / cudaMallocHost, no allocating library call), for callers
// that cannot allocate: code under CUDA graph capture, tight serving loops, harnesses that forbid it.
// This entry point has no counterpart in the
// reference.
//
// It runs the SAME kernel as garb200_kkt_error (kkt_error_kernel, forward.cu).  What that kernel
// needs besides the caller's inputs is scratch memory which the handle-based path cudaMalloc's in
// garb200_create: the per-knot dimension arrays nu_t / nc_t [T+1] and the per-problem result staging
// [batch][3].  Here that scratch lives in ONE file-scope static __device__ array of
// GARB200_NOALLOC_SCRATCH_BYTES (1 MiB): every GPU that loads the library holds it until exit.
//
// In host code `g_scratch` is only a host-side stand-in for the device array, NOT a device address:
// its device address is looked up once per device with cudaGetSymbolAddress (that device current),
// and only that pointer is handed to kernels.
//
// Every call on a device shares the array and it keeps the previous call's contents: the regions a
// call uses are (re)written by that call before the kernel reads them, and the result staging is
// zeroed first.  Calls that overlap on the GPU (two streams, two host threads on one device) RACE on
// it — serialise them.
#include <cuda_runtime.h>
#include <cstdio>
#include <mutex>
#include <string>
#include "kernels.cuh"

namespace garb200 {
int set_error(int code, const std::string &msg);   // capi.cu: garb200_last_error() text of this thread

static __device__ __align__(256) unsigned char g_scratch[GARB200_NOALLOC_SCRATCH_BYTES];

// nu_t / nc_t of uniform knots written on the device (no host buffer involved): knots t < T get
// (nu, nc), the terminal knot (nu_term, nc_term).
__global__ void noalloc_fill_dims_kernel(int *nu_t, int *nc_t, int T, int nu, int nc, int nu_term, int nc_term) {
  for (int t = blockIdx.x * blockDim.x + threadIdx.x; t <= T; t += gridDim.x * blockDim.x) {
    nu_t[t] = t < T ? nu : nu_term;
    nc_t[t] = t < T ? nc : nc_term;
  }
}

static constexpr int kMaxDevices = 64;
static std::mutex g_mu;
static unsigned char *g_dev_ptr[kMaxDevices] = {};

// Device address of g_scratch on `device` (which the caller has made current); looked up once.
static int scratch_address(int device, unsigned char **out) {
  std::lock_guard<std::mutex> lk(g_mu);
  if (!g_dev_ptr[device]) {
    void *ptr = nullptr;
    const cudaError_t e = cudaGetSymbolAddress(&ptr, g_scratch);
    if (e != cudaSuccess || !ptr)
      return set_error(GARB200_ERR_CUDA, std::string("cudaGetSymbolAddress(g_scratch): ") + cudaGetErrorString(e));
    g_dev_ptr[device] = static_cast<unsigned char *>(ptr);
  }
  *out = g_dev_ptr[device];
  return GARB200_OK;
}

static size_t round256(size_t x) { return (x + 255) & ~(size_t)255; }

} // namespace garb200

using namespace garb200;

#define NA_CK(call)                                                                              \
  do {                                                                                           \
    cudaError_t e_ = (call);                                                                     \
    if (e_ != cudaSuccess)                                                                       \
      return set_error(GARB200_ERR_CUDA, std::string(#call) + ": " + cudaGetErrorString(e_));    \
  } while (0)

extern "C" int garb200_kkt_error_noalloc(const garb200_config *dims, const double *d_knots,
                                         const double *d_G0, const double *d_g0, const double *d_xs,
                                         const double *d_us, const double *d_vs,
                                         const double *d_lbdas, const double *d_theta, double mudyn,
                                         double mueq, void *cuda_stream, double *out) {
  if (!dims || !d_knots || !d_xs || !d_lbdas || !out)
    return set_error(GARB200_ERR_INVALID, "null argument");
  const garb200_config &c = *dims;
  if (c.nx < 1 || c.nu < 0 || c.nc < 0 || c.nth < 0 || c.horizon < 0 || c.batch < 1 || c.nc0 < 0)
    return set_error(GARB200_ERR_INVALID, "negative or zero dimension");
  if (c.nc0 > c.nx) return set_error(GARB200_ERR_INVALID, "nc0 > nx is not supported");
  if ((c.nc0 > 0 && (!d_G0 || !d_g0)) || (c.nu > 0 && !d_us) || (c.nc > 0 && !d_vs))
    return set_error(GARB200_ERR_INVALID, "null argument for a non-empty block");
  const int T = c.horizon;
  if ((c.nu_t == nullptr) != (c.nc_t == nullptr))
    return set_error(GARB200_ERR_INVALID, "give both nu_t and nc_t, or neither");
  if (c.nu_t)
    for (int t = 0; t <= T; t++)
      if (c.nu_t[t] < 0 || c.nu_t[t] > c.nu || c.nc_t[t] < 0 || c.nc_t[t] > c.nc)
        return set_error(GARB200_ERR_INVALID, "per-knot nu/nc outside [0, max]");
  if (!c.nu_t && (c.nu_term < 0 || c.nu_term > c.nu || c.nc_term < 0 || c.nc_term > c.nc))
    return set_error(GARB200_ERR_INVALID, "terminal nu/nc outside [0, max]");
  // scratch plan: int nu_t[T+1] | int nc_t[T+1] | (256-byte aligned) double out[batch][3]
  const size_t dims_bytes = round256((size_t)2 * (T + 1) * sizeof(int));
  const size_t out_bytes = (size_t)c.batch * 3 * sizeof(double);
  const size_t need = dims_bytes + out_bytes;
  if (need > (size_t)GARB200_NOALLOC_SCRATCH_BYTES) {
    char buf[256];
    snprintf(buf, sizeof buf, "garb200_kkt_error_noalloc: horizon %d, batch %d need %zu bytes of scratch, the static "
             "array holds GARB200_NOALLOC_SCRATCH_BYTES = %u (use garb200_kkt_error for larger problems)",
             T, c.batch, need, (unsigned)GARB200_NOALLOC_SCRATCH_BYTES);
    return set_error(GARB200_ERR_INVALID, buf);
  }
  int ndev = 0;
  NA_CK(cudaGetDeviceCount(&ndev));
  if (c.device < 0 || c.device >= ndev || c.device >= kMaxDevices)
    return set_error(GARB200_ERR_CUDA, "no such CUDA device");
  NA_CK(cudaSetDevice(c.device));
  unsigned char *base = nullptr;
  { const int rc = scratch_address(c.device, &base); if (rc != GARB200_OK) return rc; }
  cudaStream_t st = static_cast<cudaStream_t>(cuda_stream);
  int *s_nu = reinterpret_cast<int *>(base), *s_nc = s_nu + (T + 1);
  double *s_out = reinterpret_cast<double *>(base + dims_bytes);
  // the array keeps the last call's contents: rewrite the dims, zero the result staging
  if (c.nu_t) {
    NA_CK(cudaMemcpyAsync(s_nu, c.nu_t, (size_t)(T + 1) * sizeof(int), cudaMemcpyHostToDevice, st));
    NA_CK(cudaMemcpyAsync(s_nc, c.nc_t, (size_t)(T + 1) * sizeof(int), cudaMemcpyHostToDevice, st));
  } else {
    noalloc_fill_dims_kernel<<<(T + 256) / 256, 256, 0, st>>>(s_nu, s_nc, T, c.nu, c.nc, c.nu_term, c.nc_term);
    NA_CK(cudaGetLastError());
  }
  NA_CK(cudaMemsetAsync(s_out, 0, out_bytes, st));
  DevParams p{};
  p.nx = c.nx; p.nu = c.nu; p.nc = c.nc; p.nth = c.nth;
  p.T = T; p.nc0 = c.nc0; p.batch = c.batch;
  p.nu_t = s_nu; p.nc_t = s_nc;
  p.knots = const_cast<double *>(d_knots);
  p.G0 = const_cast<double *>(d_G0); p.g0 = const_cast<double *>(d_g0);
  p.xs = const_cast<double *>(d_xs); p.us = const_cast<double *>(d_us);
  p.vs = const_cast<double *>(d_vs); p.lbdas = const_cast<double *>(d_lbdas);
  p.KL = garb200_make_knot_layout(c.nx, c.nu, c.nc, c.nth);
  const int rc = launch_kkt_error(p, c.nth > 0 ? d_theta : nullptr, mudyn, mueq, s_out, st);
  if (rc < 0) return set_error(rc == -1 ? GARB200_ERR_INVALID : GARB200_ERR_CUDA, "kkt_error launch failed");
  NA_CK(cudaMemcpyAsync(out, s_out, out_bytes, cudaMemcpyDeviceToHost, st));
  NA_CK(cudaStreamSynchronize(st));
  return GARB200_OK;
}
