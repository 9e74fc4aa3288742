// multi.cu -- one fitted model over several GPUs of one box, driven from ONE process.
//
// The reference is single-threaded and has no notion of devices; this is the B200 design for
// BASELINE.json configs[4] ("1M data (sharded) x 1M queries ... at 1/2/4/8 B200 with NCCL top-k merge")
// behind the C-ABI, so that NearestNeighbor(devices=[...]) needs neither torch nor a launcher:
//   * query sharding (the data set fits one GPU): every device holds all fitted points and takes a
//     contiguous slice of the queries -- no exchange at all;
//   * data sharding: device g holds rows [g n/G, (g+1) n/G) and all queries, finds its exact local top-k,
//     the lists travel in ONE all-to-all (grouped ncclSend / ncclRecv over NVLink) so that device g owns
//     the G lists of its slice of the queries, merges them on the device by (d2, global index) --
//     bit-identical to the single-GPU answer -- and weights its slice.
// One host thread per device; NCCL is opened with dlopen on first use (no link-time dependency, and a
// process that never shards data never loads it).
#include <dlfcn.h>
#include <nccl.h>

#include <mutex>
#include <string>
#include <thread>
#include <vector>

#include "common.cuh"

namespace b200 {
namespace {

struct NcclApi {
  void* h = nullptr;
  ncclResult_t (*CommInitAll)(ncclComm_t*, int, const int*) = nullptr;
  ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
  ncclResult_t (*GroupStart)() = nullptr;
  ncclResult_t (*GroupEnd)() = nullptr;
  ncclResult_t (*Send)(const void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*Recv)(void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
  const char* (*GetErrorString)(ncclResult_t) = nullptr;
};
NcclApi g_nccl;
std::mutex g_nccl_mu;

int load_nccl() {
  std::lock_guard<std::mutex> lk(g_nccl_mu);
  if (g_nccl.h) return 0;
  const char* names[] = {"libnccl.so.2", "libnccl.so"};
  void* h = nullptr;
  for (const char* nm : names) {
    h = dlopen(nm, RTLD_NOW | RTLD_LOCAL);
    if (h) break;
  }
  if (!h) {
    set_error("data sharding needs NCCL: dlopen(libnccl.so.2) failed: %s", dlerror());
    return 2;
  }
#define B200_SYM(field, name)                                                 \
  *reinterpret_cast<void**>(&g_nccl.field) = dlsym(h, name);                  \
  if (!g_nccl.field) {                                                        \
    set_error("NCCL symbol %s not found", name);                              \
    return 2;                                                                 \
  }
  B200_SYM(CommInitAll, "ncclCommInitAll")
  B200_SYM(CommDestroy, "ncclCommDestroy")
  B200_SYM(GroupStart, "ncclGroupStart")
  B200_SYM(GroupEnd, "ncclGroupEnd")
  B200_SYM(Send, "ncclSend")
  B200_SYM(Recv, "ncclRecv")
  B200_SYM(GetErrorString, "ncclGetErrorString")
#undef B200_SYM
  g_nccl.h = h;
  return 0;
}

__global__ void add_offset_kernel(int32_t* __restrict__ v, int64_t count, int32_t off) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < count && v[i] >= 0) v[i] += off;
}

}  // namespace
}  // namespace b200

struct b200_multi {
  int G = 0;
  bool shard_data = false;
  int64_t n = 0;
  int d = 0;
  std::vector<int> devs;
  std::vector<b200_model*> models;
  std::vector<int64_t> lo, hi;  // rows of each device (data sharding)
  std::vector<ncclComm_t> comms;
  std::vector<cudaStream_t> streams;
  std::vector<cudaEvent_t> ev;  // 4 per device: collective start/end, merge end, step end
  // per-device scratch of the data-sharded path
  std::vector<b200::DevBuf> z, idx, d2, ridx, rd2, midx, md2, out, wts, y;
  int y_r = 0;
  double ms_collective = 0, ms_merge = 0, ms_local = 0, calls = 0;
};

using namespace b200;

namespace {

// run f(g) on one host thread per device; the first error message wins
template <class F>
int per_device(b200_multi* mm, F f) {
  std::vector<int> rc(mm->G, 0);
  std::vector<std::string> msg(mm->G);
  std::vector<std::thread> th;
  for (int g = 0; g < mm->G; ++g)
    th.emplace_back([&, g]() {
      rc[g] = f(g);
      if (rc[g]) msg[g] = b200_last_error();  // thread-local text of the worker
    });
  for (auto& t : th) t.join();
  for (int g = 0; g < mm->G; ++g)
    if (rc[g]) {
      set_error("device %d: %s", mm->devs[g], msg[g].c_str());
      return rc[g];
    }
  return 0;
}

}  // namespace

extern "C" {

int b200_multi_destroy(b200_multi* mm) {
  if (!mm) return 0;
  for (int g = 0; g < (int)mm->models.size(); ++g) {
    cudaSetDevice(mm->devs[g]);
    cudaDeviceSynchronize();
    if (g < (int)mm->comms.size() && mm->comms[g] && g_nccl.CommDestroy) g_nccl.CommDestroy(mm->comms[g]);
    std::vector<std::vector<DevBuf>*> all = {&mm->z, &mm->idx, &mm->d2, &mm->ridx, &mm->rd2, &mm->midx, &mm->md2, &mm->out, &mm->wts, &mm->y};
    for (auto* v : all)
      if (g < (int)v->size()) (*v)[g].release();
    if (g < (int)mm->streams.size() && mm->streams[g]) cudaStreamDestroy(mm->streams[g]);
    for (int e = 0; e < 4; ++e)
      if (4 * g + e < (int)mm->ev.size() && mm->ev[4 * g + e]) cudaEventDestroy(mm->ev[4 * g + e]);
    b200_destroy(mm->models[g]);
  }
  delete mm;
  return 0;
}

int b200_multi_create(const double* x, int64_t n, int32_t d, int32_t n_dev, const int32_t* devs, int32_t shard_data,
                      b200_multi** out) {
  if (!x || !out || !devs || n_dev < 1 || n_dev > 64 || n < n_dev || d < 1) {
    set_error("b200_multi_create: bad arguments (n=%lld d=%d devices=%d)", (long long)n, d, n_dev);
    return 1;
  }
  b200_multi* mm = new b200_multi();
  mm->G = n_dev;
  mm->shard_data = shard_data != 0 && n_dev > 1;
  mm->n = n;
  mm->d = d;
  mm->devs.assign(devs, devs + n_dev);
  mm->models.assign(n_dev, nullptr);
  mm->lo.resize(n_dev);
  mm->hi.resize(n_dev);
  const int64_t per = (n + n_dev - 1) / n_dev;
  for (int g = 0; g < n_dev; ++g) {
    mm->lo[g] = mm->shard_data ? std::min<int64_t>(n, g * per) : 0;
    mm->hi[g] = mm->shard_data ? std::min<int64_t>(n, mm->lo[g] + per) : n;
  }
  int rc = per_device(mm, [&](int g) {
    return b200_create(x + (size_t)mm->lo[g] * d, mm->hi[g] - mm->lo[g], d, mm->devs[g], &mm->models[g]);
  });
  if (!rc && mm->shard_data) {
    rc = load_nccl();
    if (!rc) {
      mm->comms.assign(n_dev, nullptr);
      const ncclResult_t r = g_nccl.CommInitAll(mm->comms.data(), n_dev, mm->devs.data());
      if (r != ncclSuccess) {
        set_error("ncclCommInitAll: %s", g_nccl.GetErrorString(r));
        rc = 2;
      }
    }
    if (!rc) {
      mm->streams.assign(n_dev, nullptr);
      mm->ev.assign(4 * n_dev, nullptr);
      for (auto* v : {&mm->z, &mm->idx, &mm->d2, &mm->ridx, &mm->rd2, &mm->midx, &mm->md2, &mm->out, &mm->wts, &mm->y})
        v->resize(n_dev);
      for (int g = 0; g < n_dev && !rc; ++g) {
        if (cudaSetDevice(mm->devs[g]) != cudaSuccess || cudaStreamCreateWithFlags(&mm->streams[g], cudaStreamNonBlocking) != cudaSuccess)
          rc = 2;
        for (int e = 0; e < 4 && !rc; ++e)
          if (cudaEventCreate(&mm->ev[4 * g + e]) != cudaSuccess) rc = 2;
      }
      if (rc) set_error("b200_multi_create: stream / event creation failed");
    }
  }
  if (rc) {
    const std::string keep = b200_last_error();
    b200_multi_destroy(mm);
    set_error("%s", keep.c_str());
    return rc;
  }
  *out = mm;
  return 0;
}

int b200_multi_set_option(b200_multi* mm, const char* key, double value) {
  if (!mm) return 1;
  for (b200_model* m : mm->models) B200_TRY(b200_set_option(m, key, value));
  if (!strcmp(key, "reset_stats")) mm->ms_collective = mm->ms_merge = mm->ms_local = mm->calls = 0;
  return 0;
}

int b200_multi_get_stat(b200_multi* mm, const char* key, double* value) {
  if (!mm || !value) return 1;
  if (!strcmp(key, "ms_collective")) *value = mm->ms_collective;
  else if (!strcmp(key, "ms_merge")) *value = mm->ms_merge;
  else if (!strcmp(key, "ms_local")) *value = mm->ms_local;
  else if (!strcmp(key, "devices")) *value = mm->G;
  else if (!strcmp(key, "shard_data")) *value = mm->shard_data ? 1 : 0;
  else {  // sum over the devices' own counters
    double s = 0, v = 0;
    for (b200_model* m : mm->models) {
      B200_TRY(b200_get_stat(m, key, &v));
      s += v;
    }
    *value = s;
  }
  return 0;
}

int b200_multi_set_values(b200_multi* mm, const double* y, int32_t r) {
  if (!mm) return 1;
  if (!mm->shard_data) return per_device(mm, [&](int g) { return b200_set_values(mm->models[g], y, r); });
  // data sharding: the weights of a query slice can name any fitted row -> all n responses on every device
  if (!y) {
    mm->y_r = 0;
    return 0;
  }
  if (r < 1) {
    set_error("y given with r=%d", r);
    return 1;
  }
  const int rc = per_device(mm, [&](int g) {
    B200_CUDA(cudaSetDevice(mm->devs[g]));
    B200_TRY(mm->y[g].reserve((size_t)mm->n * r * 8));
    B200_CUDA(cudaMemcpy(mm->y[g].p, y, (size_t)mm->n * r * 8, cudaMemcpyHostToDevice));
    return 0;
  });
  if (!rc) mm->y_r = r;
  return rc;
}

// NearestNeighbor._predict + WeightedApproximator.predict over the devices of the handle
// (Average / Shepard weighting; see b200_nn_predict).  ids/wts (q,k) and out (q,r) may be NULL.
int b200_multi_nn_predict(b200_multi* mm, const double* z, int64_t q, int32_t k, int32_t method, int64_t* ids,
                          double* wts, double* out) {
  if (!mm || q < 0 || k < 1 || (q > 0 && !z)) {
    set_error("b200_multi_nn_predict: bad arguments");
    return 1;
  }
  if (q == 0) return 0;
  const int G = mm->G, d = mm->d;
  if (!mm->shard_data) {
    // ---- query sharding: contiguous slices, no exchange
    const int64_t per = (q + G - 1) / G;
    return per_device(mm, [&](int g) {
      const int64_t lo = std::min<int64_t>(q, g * per), hi = std::min<int64_t>(q, lo + per);
      if (hi == lo) return 0;
      int r = 0;
      return b200_nn_predict(mm->models[g], z + (size_t)lo * d, hi - lo, k, method, nullptr, r,
                             ids ? ids + (size_t)lo * k : nullptr, wts ? wts + (size_t)lo * k : nullptr, nullptr, nullptr,
                             out ? out + (size_t)lo * mm->models[g]->y_r : nullptr);
    });
  }
  // ---- data sharding
  if (method != B200_AVERAGE && method != B200_SHEPARD) {
    set_error("b200_multi_nn_predict: data sharding supports Average and Shepard weighting");
    return 1;
  }
  if (out && mm->y_r < 1) {
    set_error("no response values: call b200_multi_set_values first");
    return 1;
  }
  const int64_t qm = (q + G - 1) / G;      // queries merged and weighted per device
  const int64_t qp = qm * G;               // padded query count (padding rows repeat the last query)
  const int r = mm->y_r;
  std::vector<float> t_coll(G, 0.f), t_merge(G, 0.f), t_local(G, 0.f);
  const int rc = per_device(mm, [&](int g) {
    b200_model* m = mm->models[g];
    cudaStream_t st = mm->streams[g];
    B200_CUDA(cudaSetDevice(mm->devs[g]));
    B200_TRY(mm->z[g].reserve((size_t)qp * d * 8));
    B200_TRY(mm->idx[g].reserve((size_t)qp * k * 4));
    B200_TRY(mm->d2[g].reserve((size_t)qp * k * 8));
    B200_TRY(mm->ridx[g].reserve((size_t)G * qm * k * 4));
    B200_TRY(mm->rd2[g].reserve((size_t)G * qm * k * 8));
    B200_TRY(mm->midx[g].reserve((size_t)qm * k * 4));
    B200_TRY(mm->md2[g].reserve((size_t)qm * k * 8));
    B200_TRY(mm->wts[g].reserve((size_t)qm * k * 8));
    if (out) B200_TRY(mm->out[g].reserve((size_t)qm * r * 8));
    B200_CUDA(cudaMemcpyAsync(mm->z[g].p, z, (size_t)q * d * 8, cudaMemcpyHostToDevice, st));
    for (int64_t p = q; p < qp; ++p)  // padding
      B200_CUDA(cudaMemcpyAsync(mm->z[g].as<double>() + (size_t)p * d, z + (size_t)(q - 1) * d, (size_t)d * 8,
                                cudaMemcpyHostToDevice, st));
    cudaEventRecord(mm->ev[4 * g + 3], st);
    // exact local top-k of ALL queries against this device's rows (indices local to the shard)
    B200_TRY(b200_knn_dev(m, mm->z[g].as<double>(), qp, k, -1, mm->idx[g].as<int32_t>(), mm->d2[g].as<double>(), st));
    add_offset_kernel<<<(unsigned)((qp * k + 255) / 256), 256, 0, st>>>(mm->idx[g].as<int32_t>(), qp * k, (int32_t)mm->lo[g]);
    B200_CUDA(cudaGetLastError());
    cudaEventRecord(mm->ev[4 * g + 0], st);
    // all-to-all: my lists for peer p's query slice -> p; p's lists for my slice -> slot p
    ncclResult_t nr = g_nccl.GroupStart();
    for (int p = 0; p < G && nr == ncclSuccess; ++p) {
      nr = g_nccl.Send(mm->d2[g].as<double>() + (size_t)p * qm * k, (size_t)qm * k, ncclDouble, p, mm->comms[g], st);
      if (nr == ncclSuccess) nr = g_nccl.Recv(mm->rd2[g].as<double>() + (size_t)p * qm * k, (size_t)qm * k, ncclDouble, p, mm->comms[g], st);
      if (nr == ncclSuccess) nr = g_nccl.Send(mm->idx[g].as<int32_t>() + (size_t)p * qm * k, (size_t)qm * k, ncclInt32, p, mm->comms[g], st);
      if (nr == ncclSuccess) nr = g_nccl.Recv(mm->ridx[g].as<int32_t>() + (size_t)p * qm * k, (size_t)qm * k, ncclInt32, p, mm->comms[g], st);
    }
    if (nr == ncclSuccess) nr = g_nccl.GroupEnd();
    if (nr != ncclSuccess) {
      set_error("NCCL all-to-all: %s", g_nccl.GetErrorString(nr));
      return 2;
    }
    cudaEventRecord(mm->ev[4 * g + 1], st);
    B200_TRY(b200_merge_topk_dev(mm->devs[g], mm->rd2[g].as<double>(), mm->ridx[g].as<int32_t>(), G, qm, k,
                                 mm->md2[g].as<double>(), mm->midx[g].as<int32_t>(), st));
    cudaEventRecord(mm->ev[4 * g + 2], st);
    B200_TRY(b200_nn_weights_dev(m, nullptr, qm, k, method, mm->midx[g].as<int32_t>(), mm->md2[g].as<double>(),
                                 out ? mm->y[g].as<double>() : nullptr, r, mm->wts[g].as<double>(), nullptr, nullptr,
                                 out ? mm->out[g].as<double>() : nullptr, st));
    const int64_t lo = (int64_t)g * qm, cntq = std::max<int64_t>(0, std::min<int64_t>(q, lo + qm) - lo);
    if (cntq > 0) {
      if (wts) B200_CUDA(cudaMemcpyAsync(wts + (size_t)lo * k, mm->wts[g].p, (size_t)cntq * k * 8, cudaMemcpyDeviceToHost, st));
      if (out) B200_CUDA(cudaMemcpyAsync(out + (size_t)lo * r, mm->out[g].p, (size_t)cntq * r * 8, cudaMemcpyDeviceToHost, st));
    }
    B200_CUDA(cudaStreamSynchronize(st));
    if (ids && cntq > 0) {
      std::vector<int32_t> tmp((size_t)cntq * k);
      B200_CUDA(cudaMemcpy(tmp.data(), mm->midx[g].p, tmp.size() * 4, cudaMemcpyDeviceToHost));
      for (size_t i = 0; i < tmp.size(); ++i) ids[(size_t)lo * k + i] = tmp[i];
    }
    cudaEventElapsedTime(&t_local[g], mm->ev[4 * g + 3], mm->ev[4 * g + 0]);
    cudaEventElapsedTime(&t_coll[g], mm->ev[4 * g + 0], mm->ev[4 * g + 1]);
    cudaEventElapsedTime(&t_merge[g], mm->ev[4 * g + 1], mm->ev[4 * g + 2]);
    return 0;
  });
  if (rc) return rc;
  float a = 0, b = 0, c = 0;
  for (int g = 0; g < G; ++g) {
    a = std::max(a, t_coll[g]);
    b = std::max(b, t_merge[g]);
    c = std::max(c, t_local[g]);
  }
  mm->ms_collective += a;
  mm->ms_merge += b;
  mm->ms_local += c;
  mm->calls += 1;
  return 0;
}

}  // extern "C"
