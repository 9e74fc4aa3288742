// api.cu -- the extern "C" boundary of libb200approx.so (see include/b200approx.h).
// Host buffers in, host buffers out; all compute runs in the CUDA kernels of
// this library.  There is deliberately no CPU path behind these entry points.
#include <stdarg.h>
#include <string.h>

#include <vector>

#include "common.cuh"
#include "../../include/b200approx_debug.h"

namespace b200 {

static thread_local char g_err[512] = "";

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

namespace {

__global__ void widen_idx_kernel(const int32_t* __restrict__ a, int64_t* __restrict__ b, int64_t n) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) b[i] = a[i];
}

struct Timer {
  b200_model* m;
  cudaStream_t st;
  double* acc;
  Timer(b200_model* m_, cudaStream_t s, double* a) : m(m_), st(s), acc(a) { cudaEventRecord(m->ev0, st); }
  ~Timer() {
    cudaEventRecord(m->ev1, st);
    cudaEventSynchronize(m->ev1);
    float ms = 0;
    cudaEventElapsedTime(&ms, m->ev0, m->ev1);
    *acc += ms;
  }
};

int use(b200_model* m) {
  if (!m) {
    set_error("null model handle");
    return 1;
  }
  B200_CUDA(cudaSetDevice(m->device));
  m->defer_fallback = false;  // only b200_nn_predict's chunk loop switches it on, and off again
  return 0;
}

// d2h of an int32 device array widened to the caller's int64 host array
int fetch_idx64(b200_model* m, const int32_t* src, int64_t count, int64_t* dst) {
  B200_TRY(m->ids64.reserve((size_t)count * 8));
  widen_idx_kernel<<<(unsigned)((count + 255) / 256), 256>>>(src, m->ids64.as<int64_t>(), count);
  B200_CUDA(cudaGetLastError());
  B200_CUDA(cudaMemcpy(dst, m->ids64.p, (size_t)count * 8, cudaMemcpyDeviceToHost));
  return 0;
}

int upload_queries(b200_model* m, const double* z, int64_t q) {
  if (!z && q > 0) {
    set_error("null query pointer");
    return 1;
  }
  B200_TRY(m->zq.reserve((size_t)q * m->d * 8 + 8));
  B200_CUDA(cudaMemcpy(m->zq.p, z, (size_t)q * m->d * 8, cudaMemcpyHostToDevice));
  return 0;
}

int upload_y(b200_model* m, const double* y, int r) {
  if (!y) return 0;
  if (r < 1) {
    set_error("y given with r=%d", r);
    return 1;
  }
  B200_TRY(m->y.reserve((size_t)m->n * r * 8));
  B200_CUDA(cudaMemcpy(m->y.p, y, (size_t)m->n * r * 8, cudaMemcpyHostToDevice));
  m->y_r = r;
  return 0;
}

// Response values of a predict call: the caller's y (uploaded now), or -- y == NULL with an output
// buffer -- the values left on the device by b200_set_values.  *have: weighted sums are wanted.
int resolve_y(b200_model* m, const double* y, int32_t* r, bool want_out, bool* have) {
  *have = false;
  if (y) {
    B200_TRY(upload_y(m, y, *r));
    *have = true;
  } else if (want_out) {
    if (m->y_r < 1) {
      set_error("no response values: pass y or call b200_set_values first");
      return 1;
    }
    *r = m->y_r;
    *have = true;
  }
  return 0;
}

// Sets *flag when any query coordinate is NaN or infinite (grid-stride over the q*d doubles).
__global__ void nonfinite_kernel(const double* __restrict__ z, int64_t count, int32_t* __restrict__ flag) {
  bool bad = false;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < count; i += (int64_t)gridDim.x * blockDim.x)
    bad |= !isfinite(z[i]);
  if (__syncthreads_or(bad) && threadIdx.x == 0) *flag = 1;
}

}  // namespace

// sklearn's BallTree.query (the reference's search, nearest_neighbor/__init__.py:75) raises on NaN or
// infinite input; so does every neighbour search here, instead of returning index -1 rows.
int reject_nonfinite(b200_model* m, const double* z_dev, int64_t q, cudaStream_t st) {
  B200_TRY(m->flags.reserve(256));
  int32_t* flag = m->flags.as<int32_t>() + 6;
  B200_CUDA(cudaMemsetAsync(flag, 0, 4, st));
  const int64_t count = q * m->d;
  const unsigned blocks = (unsigned)std::min<int64_t>((count + 1023) / 1024, (int64_t)m->sm_count * 8);
  nonfinite_kernel<<<blocks, 256, 0, st>>>(z_dev, count, flag);
  B200_CUDA(cudaGetLastError());
  int h = 0;
  B200_CUDA(cudaMemcpyAsync(&h, flag, 4, cudaMemcpyDeviceToHost, st));
  B200_CUDA(cudaStreamSynchronize(st));
  if (h) {
    set_error("query points contain NaN or infinity");
    return 1;
  }
  return 0;
}

int knn_dispatch(b200_model* m, const double* z_dev, int64_t q, int k, int64_t self_offset,
                 int32_t* idx_dev, double* d2_dev, cudaStream_t st) {
  if (q > 0) B200_TRY(reject_nonfinite(m, z_dev, q, st));
  bool tensor = false;
  if (m->knn_path == B200_PATH_TENSOR) {
    if (!tc_supported(m, k)) {
      set_error("tensor kNN path unsupported for d=%d k=%d", m->d, k);
      return 1;
    }
    tensor = true;
  } else if (m->knn_path == B200_PATH_AUTO) {
    // the tensor filter pays off once the distance matrix is large; in very low dimension its worst-case
    // error bound is loose against the tiny neighbour gaps and most queries would end in the exact
    // kernel anyway (measured cross-over between d = 4 and 5, profiles/microbench/dispatch_lowd.py)
    tensor = tc_supported(m, k) && (double)m->n * (double)q >= 2.5e8 && m->n >= 4096 && m->d >= 5;
  }
  if (tensor) return knn_tc(m, z_dev, q, k, self_offset, idx_dev, d2_dev, st);
  Timer t(m, st, &m->st.ms_exact);
  m->st.exact_queries += q;
  return knn_exact(m, z_dev, q, k, self_offset, idx_dev, d2_dev, nullptr, 0, st);
}

}  // namespace b200

using namespace b200;

extern "C" {

int b200_version(void) { return 100; }
const char* b200_last_error(void) { return g_err; }

int b200_device_count(void) {
  int c = 0;
  if (cudaGetDeviceCount(&c) != cudaSuccess) return 0;
  return c;
}

int b200_create(const double* x, int64_t n, int32_t d, int32_t device, b200_model** out) {
  if (!x || !out || n < 1 || d < 1) {
    set_error("b200_create: bad arguments (n=%lld d=%d)", (long long)n, d);
    return 1;
  }
  if (n > 0x7fffffffLL - 1024) {
    set_error("b200_create: n=%lld exceeds int32 row indices", (long long)n);
    return 1;
  }
  int cnt = 0;
  cudaError_t e = cudaGetDeviceCount(&cnt);
  if (e != cudaSuccess || cnt == 0) {
    set_error("no CUDA device available (%s); libb200approx has no CPU path", cudaGetErrorString(e));
    return 2;
  }
  if (device < 0 || device >= cnt) {
    set_error("b200_create: device %d out of range (0..%d)", device, cnt - 1);
    return 1;
  }
  B200_CUDA(cudaSetDevice(device));
  cudaDeviceProp prop;
  B200_CUDA(cudaGetDeviceProperties(&prop, device));
  if (prop.major != 10) {
    set_error("device %d is sm_%d%d; this library is built for sm_100a only", device, prop.major, prop.minor);
    return 2;
  }
  b200_model* m = new b200_model();
  m->device = device;
  m->n = n;
  m->d = d;
  m->sm_count = prop.multiProcessorCount;
  // any failure from here on releases what was created (b200_destroy copes with a partly built model)
  cudaError_t ce = cudaMalloc(&m->x, (size_t)n * d * 8);
  if (ce == cudaSuccess) ce = cudaMemcpy(m->x, x, (size_t)n * d * 8, cudaMemcpyHostToDevice);
  if (ce == cudaSuccess) ce = cudaEventCreate(&m->ev0);
  if (ce == cudaSuccess) ce = cudaEventCreate(&m->ev1);
  if (ce == cudaSuccess) ce = cudaEventCreate(&m->ev2);
  if (ce == cudaSuccess) ce = cudaEventCreate(&m->ev3);
  if (ce == cudaSuccess) ce = cudaStreamCreateWithFlags(&m->copy_stream, cudaStreamNonBlocking);
  if (ce == cudaSuccess) ce = cudaEventCreateWithFlags(&m->ev_copy, cudaEventDisableTiming);
  if (ce != cudaSuccess) {
    set_error("b200_create: %s", cudaGetErrorString(ce));
    b200_destroy(m);
    return 2;
  }
  *out = m;
  return 0;
}

int b200_destroy(b200_model* m) {
  if (!m) return 0;
  cudaSetDevice(m->device);
  cudaDeviceSynchronize();
  if (m->x) cudaFree(m->x);
  DevBuf* bufs[] = {&m->tc.mu, &m->tc.img, &m->tc.nlo, &m->tc.perm, &m->tc.xs, &m->tc_sm.mu, &m->tc_sm.img, &m->tc_sm.nlo, &m->tc_sm.perm, &m->vor_g, &m->x32, &m->rw, &m->lf, &m->la, &m->y, &m->zq, &m->idx32, &m->d2,
                    &m->wts, &m->cnt, &m->err, &m->out, &m->part_d2, &m->part_idx, &m->cand,
                    &m->cand_meta, &m->zimg, &m->flags, &m->misc, &m->ids64};
  for (DevBuf* b : bufs) b->release();
  if (m->ev0) cudaEventDestroy(m->ev0);
  if (m->ev1) cudaEventDestroy(m->ev1);
  if (m->ev2) cudaEventDestroy(m->ev2);
  if (m->ev3) cudaEventDestroy(m->ev3);
  if (m->ev_copy) cudaEventDestroy(m->ev_copy);
  if (m->copy_stream) cudaStreamDestroy(m->copy_stream);
  for (int r = 0; r < 4; ++r) {
    if (m->aux_event[r]) cudaEventDestroy(m->aux_event[r]);
    if (m->aux_stream[r]) cudaStreamDestroy(m->aux_stream[r]);
  }
  delete m;
  return 0;
}

int b200_set_option(b200_model* m, const char* key, double value) {
  B200_TRY(use(m));
  if (!strcmp(key, "knn_path")) m->knn_path = (int)value;
  else if (!strcmp(key, "margin")) m->margin = (int)value;
  else if (!strcmp(key, "tc_mode")) m->tc_mode = (int)value;
  else if (!strcmp(key, "tc_halves")) m->tc_halves = (int)value;
  else if (!strcmp(key, "tc_slots")) m->tc_slots = (int)value;
  else if (!strcmp(key, "tc_stats")) m->tc_stats = (int)value;
  else if (!strcmp(key, "tc_parts")) m->tc_parts = ((int)value == 2 || (int)value == 4) ? (int)value : 1;
  else if (!strcmp(key, "tail_split")) m->no_tail_split = value == 0.0;
  else if (!strcmp(key, "fallback")) m->force_split_fallback = (int)value == 1;
  else if (!strcmp(key, "rows_per_call")) m->rows_per_call = (int64_t)value;
  else if (!strcmp(key, "reset_stats")) {
    m->st = Stats();
    for (double& v : m->tc_stat_vals) v = 0;
  }
  else {
    set_error("unknown option '%s'", key);
    return 1;
  }
  return 0;
}

int b200_get_stat(b200_model* m, const char* key, double* value) {
  B200_TRY(use(m));
  const Stats& s = m->st;
  if (!strcmp(key, "uncertified")) *value = s.uncertified;
  else if (!strcmp(key, "ms_stage1")) *value = s.ms_stage1;
  else if (!strcmp(key, "ms_rerank")) *value = s.ms_rerank;
  else if (!strcmp(key, "ms_prep")) *value = s.ms_prep;
  else if (!strcmp(key, "ms_weights")) *value = s.ms_weights;
  else if (!strcmp(key, "ms_exact")) *value = s.ms_exact;
  else if (!strcmp(key, "launches")) *value = s.launches;
  else if (!strcmp(key, "tc_queries")) *value = s.tc_queries;
  else if (!strcmp(key, "exact_queries")) *value = s.exact_queries;
  else if (!strcmp(key, "tc_K")) *value = m->tc.K;
  else if (!strcmp(key, "tc_scale")) *value = m->tc.scale;
  else if (!strcmp(key, "tc_ymax")) *value = m->tc.ymax;
  else if (!strcmp(key, "tc_cap")) *value = s.last_cap;
  else if (!strcmp(key, "tc_keep")) *value = s.last_keep;
  else if (!strcmp(key, "tc_slots")) *value = s.last_slots;
  else if (!strcmp(key, "tc_halves")) *value = s.last_halves;
  else if (!strcmp(key, "sm_count")) *value = m->sm_count;
  else if (!strncmp(key, "tc_stat", 7) && atoi(key + 7) >= 0 && atoi(key + 7) < 16) *value = m->tc_stat_vals[atoi(key + 7)];
  else {
    set_error("unknown stat '%s'", key);
    return 1;
  }
  return 0;
}

int b200_set_values(b200_model* m, const double* y, int32_t r) {
  B200_TRY(use(m));
  if (!y) {
    m->y_r = 0;
    return 0;
  }
  return upload_y(m, y, r);
}

int b200_knn(b200_model* m, const double* z, int64_t q, int32_t k, int64_t self_offset,
             int64_t* idx, double* d2) {
  B200_TRY(use(m));
  if (q < 0 || !idx || k < 1) {
    set_error("b200_knn: bad arguments (q=%lld k=%d)", (long long)q, k);
    return 1;
  }
  if (q == 0) return 0;
  B200_TRY(upload_queries(m, z, q));
  B200_TRY(m->idx32.reserve((size_t)q * k * 4));
  B200_TRY(m->d2.reserve((size_t)q * k * 8));
  B200_TRY(knn_dispatch(m, m->zq.as<double>(), q, k, self_offset, m->idx32.as<int32_t>(),
                        m->d2.as<double>(), 0));
  B200_TRY(fetch_idx64(m, m->idx32.as<int32_t>(), q * k, idx));
  if (d2) B200_CUDA(cudaMemcpy(d2, m->d2.p, (size_t)q * k * 8, cudaMemcpyDeviceToHost));
  B200_CUDA(cudaDeviceSynchronize());
  return 0;
}

int b200_nn_predict(b200_model* m, const double* z, int64_t q, int32_t k, int32_t method,
                    const double* y, int32_t r, int64_t* ids, double* wts, int32_t* cnt,
                    int32_t* err, double* out) {
  B200_TRY(use(m));
  if (q < 0 || (y && !out) || k < 1) {
    set_error("b200_nn_predict: bad arguments (q=%lld k=%d)", (long long)q, k);
    return 1;
  }
  if (q == 0) return 0;
  bool have_y = false;
  B200_TRY(resolve_y(m, y, &r, out != nullptr, &have_y));
  B200_TRY(m->zq.reserve((size_t)q * m->d * 8 + 8));
  B200_TRY(m->idx32.reserve((size_t)q * k * 4));
  B200_TRY(m->d2.reserve((size_t)q * k * 8));
  B200_TRY(m->wts.reserve((size_t)q * k * 8));
  B200_TRY(m->cnt.reserve((size_t)q * 4));
  B200_TRY(m->err.reserve((size_t)q * 4));
  if (have_y) B200_TRY(m->out.reserve((size_t)q * r * 8));
  // Large batches from page-locked host memory go up in chunks on a copy stream, each chunk's
  // transfer under the previous chunk's kernels.  Chunks are whole waves of the tensor kernel's
  // persistent grid (sm_count blocks of 256 queries), a short first one so the kernels start early.
  const int64_t wave = (int64_t)m->sm_count * 256;
  bool chunked = false;
  if (q >= 2 * wave && method != B200_VORONOI) {
    cudaPointerAttributes at;
    chunked = cudaPointerGetAttributes(&at, z) == cudaSuccess && at.type == cudaMemoryTypeHost;
    cudaGetLastError();
  }
  const size_t row_bytes = (size_t)m->d * 8;
  int64_t done = 0, next = chunked ? (q >= 6 * wave ? 2 * wave : wave) : q;  // [done, next): chunk whose copy is in flight
  B200_CUDA(cudaMemcpyAsync(m->zq.p, z, (size_t)next * row_bytes, cudaMemcpyHostToDevice, m->copy_stream));
  m->deferred.clear();
  m->defer_fallback = chunked;
  while (done < q) {
    B200_CUDA(cudaEventRecord(m->ev_copy, m->copy_stream));
    B200_CUDA(cudaEventSynchronize(m->ev_copy));
    const int64_t lo = done, hi = next;
    if (hi < q) {  // start the next transfer before this chunk's kernels
      next = std::min<int64_t>(q, hi + 6 * wave);
      if (q - next < 2 * wave) next = q;
      B200_CUDA(cudaMemcpyAsync(m->zq.as<char>() + (size_t)hi * row_bytes,
                                reinterpret_cast<const char*>(z) + (size_t)hi * row_bytes,
                                (size_t)(next - hi) * row_bytes, cudaMemcpyHostToDevice, m->copy_stream));
    }
    const double* zc = m->zq.as<double>() + (size_t)lo * m->d;
    int32_t* ic = m->idx32.as<int32_t>() + (size_t)lo * k;
    double* dc = m->d2.as<double>() + (size_t)lo * k;
    m->defer_base = lo;
    {
      const int rc = knn_dispatch(m, zc, hi - lo, k, -1, ic, dc, 0);
      if (rc) {
        m->defer_fallback = false;
        return rc;
      }
    }
    {
      Timer t(m, 0, &m->st.ms_weights);
      B200_TRY(nn_weights(m, zc, hi - lo, k, method, ic, dc, have_y ? m->y.as<double>() : nullptr, r,
                          m->wts.as<double>() + (size_t)lo * k, m->cnt.as<int32_t>() + lo, m->err.as<int32_t>() + lo,
                          have_y ? m->out.as<double>() + (size_t)lo * r : nullptr, 0));
    }
    done = hi;
  }
  m->defer_fallback = false;
  m->defer_base = 0;
  if (!m->deferred.empty()) {
    // the uncertified queries of all chunks: one exact rescan, then their weights again
    const int nf = (int)m->deferred.size();
    B200_TRY(m->ids64.reserve((size_t)nf * 4 + 64));
    int32_t* fl = m->ids64.as<int32_t>();
    B200_CUDA(cudaMemcpy(fl, m->deferred.data(), (size_t)nf * 4, cudaMemcpyHostToDevice));
    {
      Timer t(m, 0, &m->st.ms_exact);
      B200_TRY(knn_rescan(m, m->zq.as<double>(), q, k, -1, m->idx32.as<int32_t>(), m->d2.as<double>(), fl, nf, 0));
    }
    for (int32_t i : m->deferred) {
      B200_TRY(nn_weights(m, m->zq.as<double>() + (size_t)i * m->d, 1, k, method, m->idx32.as<int32_t>() + (size_t)i * k,
                          m->d2.as<double>() + (size_t)i * k, have_y ? m->y.as<double>() : nullptr, r,
                          m->wts.as<double>() + (size_t)i * k, m->cnt.as<int32_t>() + i, m->err.as<int32_t>() + i,
                          have_y ? m->out.as<double>() + (size_t)i * r : nullptr, 0));
    }
    m->deferred.clear();
  }
  if (ids) {
    const int32_t* src = method == B200_VORONOI ? m->misc.as<int32_t>() : m->idx32.as<int32_t>();
    B200_TRY(fetch_idx64(m, src, q * k, ids));
  }
  if (wts) B200_CUDA(cudaMemcpy(wts, m->wts.p, (size_t)q * k * 8, cudaMemcpyDeviceToHost));
  if (cnt) B200_CUDA(cudaMemcpy(cnt, m->cnt.p, (size_t)q * 4, cudaMemcpyDeviceToHost));
  if (err) B200_CUDA(cudaMemcpy(err, m->err.p, (size_t)q * 4, cudaMemcpyDeviceToHost));
  if (have_y) B200_CUDA(cudaMemcpy(out, m->out.p, (size_t)q * r * 8, cudaMemcpyDeviceToHost));
  B200_CUDA(cudaDeviceSynchronize());
  return 0;
}

int b200_knn_dev(b200_model* m, const double* z_dev, int64_t q, int32_t k, int64_t self_offset,
                 int32_t* idx_dev, double* d2_dev, void* stream) {
  B200_TRY(use(m));
  if (q < 0 || k < 1 || !idx_dev) {
    set_error("b200_knn_dev: bad arguments (q=%lld k=%d)", (long long)q, k);
    return 1;
  }
  if (q == 0) return 0;
  double* d2 = d2_dev;
  if (!d2) {
    B200_TRY(m->d2.reserve((size_t)q * k * 8));
    d2 = m->d2.as<double>();
  }
  return knn_dispatch(m, z_dev, q, k, self_offset, idx_dev, d2, (cudaStream_t)stream);
}

int b200_nn_weights_dev(b200_model* m, const double* z_dev, int64_t q, int32_t k, int32_t method,
                        const int32_t* idx_dev, const double* d2_dev, const double* y_dev, int32_t r,
                        double* wts_dev, int32_t* cnt_dev, int32_t* err_dev, double* out_dev,
                        void* stream) {
  B200_TRY(use(m));
  return nn_weights(m, z_dev, q, k, method, idx_dev, d2_dev, y_dev, r, wts_dev, cnt_dev, err_dev, out_dev,
                    (cudaStream_t)stream);
}

int b200_tc_debug_scores(b200_model* m, const double* z, int64_t q, int32_t k, float* scores,
                         int64_t cols) {
  B200_TRY(use(m));
  const int64_t ncols = ((m->n + 127) / 128) * 128;
  if (cols != ncols || q < 1 || q > 4096) {
    set_error("b200_tc_debug_scores: need cols == %lld and 1 <= q <= 4096", (long long)ncols);
    return 1;
  }
  const int64_t qpad = ((q + 255) / 256) * 256;
  DevBuf dump;
  B200_TRY(dump.reserve((size_t)qpad * ncols * 4));
  B200_TRY(upload_queries(m, z, q));
  B200_TRY(m->idx32.reserve((size_t)q * k * 4));
  B200_TRY(m->d2.reserve((size_t)q * k * 8));
  m->dbg_scores = dump.as<float>();
  int rc = knn_tc(m, m->zq.as<double>(), q, k, -1, m->idx32.as<int32_t>(), m->d2.as<double>(), 0);
  m->dbg_scores = nullptr;
  // the dump is in image-position order: hand it back in fitted-row order (padding columns stay last)
  if (rc == 0) {
    std::vector<float> raw((size_t)q * ncols);
    std::vector<int32_t> perm((size_t)m->n);
    if (cudaMemcpy(raw.data(), dump.p, raw.size() * 4, cudaMemcpyDeviceToHost) != cudaSuccess ||
        cudaMemcpy(perm.data(), m->tc.perm.p, perm.size() * 4, cudaMemcpyDeviceToHost) != cudaSuccess) {
      set_error("b200_tc_debug_scores: copy back failed");
      rc = 2;
    } else {
      for (int64_t r = 0; r < q; ++r) {
        const float* src = raw.data() + (size_t)r * ncols;
        float* dst = scores + (size_t)r * ncols;
        for (int64_t c = 0; c < m->n; ++c) dst[perm[c]] = src[c];
        for (int64_t c = m->n; c < ncols; ++c) dst[c] = src[c];
      }
    }
  }
  dump.release();
  return rc;
}

int b200_merge_topk_dev(int32_t device, const double* d2_parts, const int32_t* idx_parts,
                        int32_t parts, int64_t q, int32_t k, double* d2_out, int32_t* idx_out,
                        void* stream) {
  B200_CUDA(cudaSetDevice(device));
  return merge_topk(d2_parts, idx_parts, parts, q, k, d2_out, idx_out, (cudaStream_t)stream);
}

int b200_shepard_predict(b200_model* m, const double* z, int64_t q, const double* y, int32_t r, double* wts,
                         double* out) {
  B200_TRY(use(m));
  if (q < 0 || (y && !out) || (!out && !wts)) {
    set_error("b200_shepard_predict: bad arguments");
    return 1;
  }
  if (q == 0) return 0;
  B200_TRY(upload_queries(m, z, q));
  bool have_y = false;
  B200_TRY(resolve_y(m, y, &r, out != nullptr, &have_y));
  if (have_y) B200_TRY(m->out.reserve((size_t)q * r * 8));
  if (wts) B200_TRY(m->wts.reserve((size_t)q * m->n * 8));
  {
    Timer t(m, 0, &m->st.ms_weights);
    B200_TRY(shepard_all(m, m->zq.as<double>(), q, have_y ? m->y.as<double>() : nullptr, r,
                         wts ? m->wts.as<double>() : nullptr, have_y ? m->out.as<double>() : nullptr, 0));
  }
  if (wts) B200_CUDA(cudaMemcpy(wts, m->wts.p, (size_t)q * m->n * 8, cudaMemcpyDeviceToHost));
  if (have_y) B200_CUDA(cudaMemcpy(out, m->out.p, (size_t)q * r * 8, cudaMemcpyDeviceToHost));
  B200_CUDA(cudaDeviceSynchronize());
  return 0;
}

int b200_shepmod_fit(b200_model* m, double* rw, int32_t* ier) {
  B200_TRY(use(m));
  int e = 0;
  B200_TRY(shepmod_fit(m, 0, &e));
  if (ier) *ier = e;
  if (e == 0 && rw) B200_CUDA(cudaMemcpy(rw, m->rw.p, (size_t)m->n * 8, cudaMemcpyDeviceToHost));
  B200_CUDA(cudaDeviceSynchronize());
  return 0;
}

int b200_eval_mesh(int32_t device, const double* int_prods, int32_t n, const double* app_int_prods, int64_t q,
                   double* mesh) {
  B200_CUDA(cudaSetDevice(device));
  if (n < 1 || q < 0 || !int_prods || (q > 0 && (!app_int_prods || !mesh))) {
    set_error("b200_eval_mesh: bad arguments");
    return 1;
  }
  if (q == 0) return 0;
  DevBuf ip, ap, out;
  int rc = ip.reserve((size_t)n * n * 8);
  if (!rc) rc = ap.reserve((size_t)q * n * 8);
  if (!rc) rc = out.reserve((size_t)q * n * 8);
  if (!rc && (cudaMemcpy(ip.p, int_prods, (size_t)n * n * 8, cudaMemcpyHostToDevice) != cudaSuccess ||
              cudaMemcpy(ap.p, app_int_prods, (size_t)q * n * 8, cudaMemcpyHostToDevice) != cudaSuccess)) {
    set_error("b200_eval_mesh: upload failed");
    rc = 2;
  }
  if (!rc) rc = eval_mesh(ip.as<double>(), n, ap.as<double>(), q, out.as<double>(), 0);
  if (!rc && cudaMemcpy(mesh, out.p, (size_t)q * n * 8, cudaMemcpyDeviceToHost) != cudaSuccess) {
    set_error("b200_eval_mesh: copy back failed");
    rc = 2;
  }
  ip.release();
  ap.release();
  out.release();
  return rc;
}

int b200_voronoi_predict(b200_model* m, const double* z, int64_t q, int32_t cap, int64_t* ids,
                         double* wts, int32_t* cnt, int32_t* err, const double* y, int32_t r,
                         double* out) {
  B200_TRY(use(m));
  if (q < 0 || cap < 1 || (y && !out)) {
    set_error("b200_voronoi_predict: bad arguments");
    return 1;
  }
  if (q == 0) return 0;
  B200_TRY(upload_queries(m, z, q));
  bool have_y = false;
  B200_TRY(resolve_y(m, y, &r, out != nullptr, &have_y));
  B200_TRY(m->idx32.reserve((size_t)q * cap * 4));
  B200_TRY(m->wts.reserve((size_t)q * cap * 8));
  B200_TRY(m->cnt.reserve((size_t)q * 4));
  B200_TRY(m->err.reserve((size_t)q * 4));
  if (have_y) B200_TRY(m->out.reserve((size_t)q * r * 8));
  {
    Timer t(m, 0, &m->st.ms_weights);
    B200_TRY(voronoi_all(m, m->zq.as<double>(), q, cap, have_y ? m->y.as<double>() : nullptr, r, m->idx32.as<int32_t>(),
                         m->wts.as<double>(), m->cnt.as<int32_t>(), m->err.as<int32_t>(),
                         have_y ? m->out.as<double>() : nullptr, 0));
  }
  if (ids) B200_TRY(fetch_idx64(m, m->idx32.as<int32_t>(), q * cap, ids));
  if (wts) B200_CUDA(cudaMemcpy(wts, m->wts.p, (size_t)q * cap * 8, cudaMemcpyDeviceToHost));
  if (cnt) B200_CUDA(cudaMemcpy(cnt, m->cnt.p, (size_t)q * 4, cudaMemcpyDeviceToHost));
  if (err) B200_CUDA(cudaMemcpy(err, m->err.p, (size_t)q * 4, cudaMemcpyDeviceToHost));
  if (have_y) B200_CUDA(cudaMemcpy(out, m->out.p, (size_t)q * r * 8, cudaMemcpyDeviceToHost));
  B200_CUDA(cudaDeviceSynchronize());
  return 0;
}

int b200_shepmod_predict(b200_model* m, const double* z, int64_t q, int32_t cap, int64_t* ids,
                         double* wts, int32_t* cnt, int32_t* ier, const double* y, int32_t r,
                         double* out) {
  B200_TRY(use(m));
  if (q < 0 || (y && !out) || cap < 1) {
    set_error("b200_shepmod_predict: bad arguments (q=%lld cap=%d)", (long long)q, cap);
    return 1;
  }
  if (q == 0) return 0;
  B200_TRY(upload_queries(m, z, q));
  bool have_y = false;
  B200_TRY(resolve_y(m, y, &r, out != nullptr, &have_y));
  B200_TRY(m->idx32.reserve((size_t)q * cap * 4));
  B200_TRY(m->wts.reserve((size_t)q * cap * 8));
  B200_TRY(m->cnt.reserve((size_t)q * 4));
  B200_TRY(m->err.reserve((size_t)q * 4));
  if (have_y) B200_TRY(m->out.reserve((size_t)q * r * 8));
  {
    Timer t(m, 0, &m->st.ms_weights);
    B200_TRY(shepmod_predict(m, m->zq.as<double>(), q, cap, m->idx32.as<int32_t>(), m->wts.as<double>(),
                             m->cnt.as<int32_t>(), m->err.as<int32_t>(), have_y ? m->y.as<double>() : nullptr,
                             r, have_y ? m->out.as<double>() : nullptr, 0));
  }
  if (ids) B200_TRY(fetch_idx64(m, m->idx32.as<int32_t>(), q * cap, ids));
  if (wts) B200_CUDA(cudaMemcpy(wts, m->wts.p, (size_t)q * cap * 8, cudaMemcpyDeviceToHost));
  if (cnt) B200_CUDA(cudaMemcpy(cnt, m->cnt.p, (size_t)q * 4, cudaMemcpyDeviceToHost));
  if (ier) B200_CUDA(cudaMemcpy(ier, m->err.p, (size_t)q * 4, cudaMemcpyDeviceToHost));
  if (have_y) B200_CUDA(cudaMemcpy(out, m->out.p, (size_t)q * r * 8, cudaMemcpyDeviceToHost));
  B200_CUDA(cudaDeviceSynchronize());
  return 0;
}

static int lshep_fit_host(b200_model* m, const double* f, int32_t r, double* a, double* rw, int32_t* ier, int mode) {
  B200_TRY(use(m));
  if (!f || r < 1) {
    set_error("b200_lshep_fit: bad arguments");
    return 1;
  }
  const int64_t n = m->n;
  std::vector<double> ft((size_t)r * n);  // (n,r) -> (r,n): one contiguous response per column
  for (int64_t i = 0; i < n; ++i)
    for (int c = 0; c < r; ++c) ft[(size_t)c * n + i] = f[i * r + c];
  B200_TRY(m->out.reserve((size_t)r * n * 8));
  B200_CUDA(cudaMemcpy(m->out.p, ft.data(), (size_t)r * n * 8, cudaMemcpyHostToDevice));
  int e = 0;
  B200_TRY(lshep_fit(m, m->out.as<double>(), r, 0, &e, mode));
  if (ier) *ier = e;
  if (e != 1) {
    if (a) B200_CUDA(cudaMemcpy(a, m->la.p, (size_t)r * n * m->d * 8, cudaMemcpyDeviceToHost));
    if (rw) B200_CUDA(cudaMemcpy(rw, m->rw.p, (size_t)n * 8, cudaMemcpyDeviceToHost));
  }
  B200_CUDA(cudaDeviceSynchronize());
  return 0;
}

int b200_lshep_fit(b200_model* m, const double* f, int32_t r, double* a, double* rw, int32_t* ier) {
  return lshep_fit_host(m, f, r, a, rw, ier, 0);
}

int b200_lshep_fit_robust(b200_model* m, const double* f, double* a, double* rw, int32_t* ier) {
  return lshep_fit_host(m, f, 1, a, rw, ier, 1);
}

int b200_ripple_fit(b200_model* m, const double* f, double* a, double* rw, int32_t* ier) {
  return lshep_fit_host(m, f, 1, a, rw, ier, 2);
}

int b200_lshep_predict(b200_model* m, double radius, const double* z, int64_t q, double* out,
                       int32_t* ier) {
  B200_TRY(use(m));
  if (q < 0 || !out) {
    set_error("b200_lshep_predict: bad arguments");
    return 1;
  }
  if (q == 0) return 0;
  const int r = m->lshep_r;
  B200_TRY(upload_queries(m, z, q));
  B200_TRY(m->out.reserve((size_t)q * r * 8));
  B200_TRY(m->err.reserve((size_t)q * r * 4));
  {
    Timer t(m, 0, &m->st.ms_weights);
    B200_TRY(lshep_predict(m, radius, m->zq.as<double>(), q, m->out.as<double>(), m->err.as<int32_t>(), 0));
  }
  B200_CUDA(cudaMemcpy(out, m->out.p, (size_t)q * r * 8, cudaMemcpyDeviceToHost));
  if (ier) B200_CUDA(cudaMemcpy(ier, m->err.p, (size_t)q * r * 4, cudaMemcpyDeviceToHost));
  B200_CUDA(cudaDeviceSynchronize());
  return 0;
}

}  // extern "C"
