// multi.cu — one LOGICAL batched solver over several GPUs of one box, behind the C ABI (SURVEY 8(b)
// "multi-GPU by opts.devices[]", 8(e)): the batch is cut into contiguous slices, GPU g owns problems
// [g B / G, (g + 1) B / G); there is NO communication inside backward or forward.  The only exchange is
// the gather of the trajectories after the rollout: to host buffers (each device copies its slice to its
// offset) or to the ROOT device with cudaMemcpyPeerAsync — copy engines over NVLink, no kernel, so nothing
// competes with the next sweep for SMs.  One host thread drives all devices (every call is asynchronous).
#include <cuda_runtime.h>
#include <string>
#include <vector>
#include "kernels.cuh"

namespace garb200 { int set_error(int code, const std::string &msg); }
using garb200::set_error;

struct garb200_multi {
  garb200_config cfg{};
  std::vector<garb200_solver *> shard;
  std::vector<int> device, first, count;
  std::vector<cudaStream_t> stream;
  std::vector<cudaEvent_t> done;      // per shard: rollout finished
  size_t kstride = 0;
};

#define MCK(call)                                                                              \
  do {                                                                                         \
    cudaError_t e_ = (call);                                                                   \
    if (e_ != cudaSuccess)                                                                     \
      return set_error(GARB200_ERR_CUDA, std::string(#call) + ": " + cudaGetErrorString(e_));  \
  } while (0)
#define MTRY(x) do { int rc_ = (x); if (rc_ != GARB200_OK) return rc_; } while (0)

extern "C" {

void garb200_multi_destroy(garb200_multi *m) {
  if (!m) return;
  for (size_t i = 0; i < m->shard.size(); i++) {
    if (m->shard[i]) garb200_destroy(m->shard[i]);
    cudaSetDevice(m->device[i]);
    if (i < m->done.size() && m->done[i]) cudaEventDestroy(m->done[i]);
    if (i < m->stream.size() && m->stream[i]) cudaStreamDestroy(m->stream[i]);
  }
  delete m;
}

int garb200_multi_create(const garb200_config *cfg, const int *devices, int ndev, garb200_multi **out) {
  if (!cfg || !devices || !out || ndev < 1) return set_error(GARB200_ERR_INVALID, "null argument");
  *out = nullptr;
  if (cfg->batch < ndev) return set_error(GARB200_ERR_INVALID, "fewer problems than devices");
  if (cfg->num_legs >= 2)   // the parallel-in-time solver does not shard: replicas only
    return set_error(GARB200_ERR_INVALID, "the parallel-in-time solver runs on one GPU: create one solver per device");
  garb200_multi *m = new garb200_multi();
  m->cfg = *cfg;
  const garb200_knot_layout KL = garb200_make_knot_layout(cfg->nx, cfg->nu, cfg->nc, cfg->nth);
  m->kstride = (size_t)KL.stride;
  for (int g = 0; g < ndev; g++) {
    const int b0 = (int)((long long)g * cfg->batch / ndev), b1 = (int)((long long)(g + 1) * cfg->batch / ndev);
    garb200_config c = *cfg;
    c.batch = b1 - b0;
    c.device = devices[g];
    garb200_solver *s = nullptr;
    const int rc = garb200_create(&c, &s);
    m->shard.push_back(s); m->device.push_back(devices[g]); m->first.push_back(b0); m->count.push_back(b1 - b0);
    m->stream.push_back(nullptr); m->done.push_back(nullptr);
    if (rc != GARB200_OK) { garb200_multi_destroy(m); return rc; }
    cudaStream_t st = nullptr;
    cudaEvent_t ev = nullptr;
    if (cudaSetDevice(devices[g]) != cudaSuccess || cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking) != cudaSuccess ||
        cudaEventCreateWithFlags(&ev, cudaEventDisableTiming) != cudaSuccess) {
      garb200_multi_destroy(m);
      return set_error(GARB200_ERR_CUDA, "stream/event creation failed");
    }
    m->stream[g] = st; m->done[g] = ev;
    garb200_set_stream(s, st);
  }
  // peer access root <-> shards for the device-side gather (ignored when unavailable: the copies then stage)
  for (int g = 1; g < ndev; g++)
    if (devices[g] != devices[0]) {
      int can = 0;
      if (cudaDeviceCanAccessPeer(&can, devices[g], devices[0]) == cudaSuccess && can) {
        cudaSetDevice(devices[g]);
        if (cudaDeviceEnablePeerAccess(devices[0], 0) != cudaSuccess) cudaGetLastError();
      }
    }
  *out = m;
  return GARB200_OK;
}

int garb200_multi_num_shards(const garb200_multi *m) { return m ? (int)m->shard.size() : 0; }

int garb200_multi_shard(garb200_multi *m, int i, garb200_solver **solver, int *device, int *first, int *count) {
  if (!m || i < 0 || i >= (int)m->shard.size()) return set_error(GARB200_ERR_INVALID, "shard index out of range");
  if (solver) *solver = m->shard[i];
  if (device) *device = m->device[i];
  if (first) *first = m->first[i];
  if (count) *count = m->count[i];
  return GARB200_OK;
}

int garb200_multi_upload_packed(garb200_multi *m, const double *knots, const double *G0, const double *g0) {
  if (!m || !knots) return set_error(GARB200_ERR_INVALID, "null argument");
  const size_t per = (size_t)(m->cfg.horizon + 1) * m->kstride, n = m->cfg.nx, c0 = m->cfg.nc0;
  for (size_t g = 0; g < m->shard.size(); g++) {
    const size_t b0 = (size_t)m->first[g];
    MTRY(garb200_upload_packed(m->shard[g], knots + b0 * per, G0 ? G0 + b0 * c0 * n : nullptr, g0 ? g0 + b0 * c0 : nullptr));
  }
  return GARB200_OK;
}

int garb200_multi_generate(garb200_multi *m, uint64_t seed, int dist) {
  if (!m) return set_error(GARB200_ERR_INVALID, "null argument");
  for (size_t g = 0; g < m->shard.size(); g++) MTRY(garb200_generate(m->shard[g], seed + g, dist));
  return GARB200_OK;
}

int garb200_multi_backward(garb200_multi *m, double mudyn, double mueq) {
  if (!m) return set_error(GARB200_ERR_INVALID, "null argument");
  for (auto *s : m->shard) MTRY(garb200_backward(s, mudyn, mueq));      // asynchronous on every device
  return GARB200_OK;
}

// forward on every shard, then the gather:
//   root_device < 0: outputs are HOST pointers for the GLOBAL batch (each device copies its slice to its offset)
//   root_device >= 0: outputs are DEVICE pointers on that device for the global batch; slices arrive by
//                     cudaMemcpyPeerAsync (copy engines, no kernel)
int garb200_multi_forward(garb200_multi *m, const double *theta, int root_device, double *xs, double *us,
                          double *vs, double *lbdas) {
  if (!m) return set_error(GARB200_ERR_INVALID, "null argument");
  const garb200_config &c = m->cfg;
  const size_t T = c.horizon, n = c.nx, nu = c.nu, nc = c.nc, c0 = c.nc0, nth = c.nth;
  for (size_t g = 0; g < m->shard.size(); g++) {
    garb200_solver *s = m->shard[g];
    const size_t b0 = (size_t)m->first[g], cnt = (size_t)m->count[g];
    MCK(cudaSetDevice(m->device[g]));
    double *dth = nullptr;
    if (theta && nth > 0) {                      // host theta [global batch][nth] -> this shard's device copy
      MTRY(garb200_device_theta(s, &dth));
      MCK(cudaMemcpyAsync(dth, theta + b0 * nth, cnt * nth * sizeof(double), cudaMemcpyHostToDevice, m->stream[g]));
    }
    MTRY(garb200_forward_device(s, dth, nullptr, nullptr, nullptr, nullptr));
    double *dx, *du, *dv, *dl;
    MTRY(garb200_device_trajectory(s, &dx, &du, &dv, &dl));
    struct Seg { double *dst; const double *src; size_t off, len; };
    const Seg seg[4] = {{xs, dx, b0 * (T + 1) * n, cnt * (T + 1) * n}, {us, du, b0 * (T + 1) * nu, cnt * (T + 1) * nu},
                        {vs, dv, b0 * (T + 1) * nc, cnt * (T + 1) * nc}, {lbdas, dl, b0 * (c0 + T * n), cnt * (c0 + T * n)}};
    for (const Seg &q : seg) {
      if (!q.dst || q.len == 0) continue;
      if (root_device < 0)
        MCK(cudaMemcpyAsync(q.dst + q.off, q.src, q.len * sizeof(double), cudaMemcpyDeviceToHost, m->stream[g]));
      else
        MCK(cudaMemcpyPeerAsync(q.dst + q.off, root_device, q.src, m->device[g], q.len * sizeof(double), m->stream[g]));
    }
    MCK(cudaEventRecord(m->done[g], m->stream[g]));
  }
  return GARB200_OK;
}

// Block until every shard has finished; GARB200_ERR_NUMERIC if any problem of any shard raised a flag.
int garb200_multi_synchronize(garb200_multi *m) {
  if (!m) return set_error(GARB200_ERR_INVALID, "null argument");
  int rc = GARB200_OK;
  for (auto *s : m->shard) {
    const int r = garb200_synchronize(s);
    if (r != GARB200_OK) rc = r;
  }
  return rc;
}

// KKT residuals of the last forward on every shard: out [global batch][3] (host)
int garb200_multi_kkt_error(garb200_multi *m, double mudyn, double mueq, const double *theta, double *out) {
  if (!m || !out) return set_error(GARB200_ERR_INVALID, "null argument");
  for (size_t g = 0; g < m->shard.size(); g++) {
    const size_t b0 = (size_t)m->first[g];
    MTRY(garb200_kkt_error(m->shard[g], mudyn, mueq, theta ? theta + b0 * m->cfg.nth : nullptr, out + 3 * b0));
  }
  return GARB200_OK;
}

} // extern "C"
