// kzg_gpu_group: several devices of one box behind the C ABI (include/kzg_gpu.h, "Multi-GPU").
//
// Written purely on top of the single-device entry points of the same header, so it is also the proof that
// the ABI is sufficient for a multi-GPU host: one kzg_gpu_ctx with the replicated SRS per device, one more with
// that device's point-range slice, and one persistent host worker thread per device (every context call is
// synchronous, so devices only overlap when different threads drive them).  This stands where the reference has
// its std::thread fan-out over index ranges (src/trusted_setup.cpp:39-73).
#include <algorithm>
#include <condition_variable>
#include <cstring>
#include <functional>
#include <mutex>
#include <new>
#include <string>
#include <thread>
#include <vector>

#include "../../include/kzg_gpu.h"

namespace {

struct Worker {
  std::thread th;
  std::mutex m;
  std::condition_variable cv;
  std::function<int()> job;
  bool has_job = false, stop = false;
  int result = 0;

  Worker() : th([this] { loop(); }) {}
  ~Worker() {
    {
      std::lock_guard<std::mutex> l(m);
      stop = true;
    }
    cv.notify_all();
    th.join();
  }
  void loop() {
    std::unique_lock<std::mutex> l(m);
    for (;;) {
      cv.wait(l, [this] { return has_job || stop; });
      if (stop) return;
      std::function<int()> fn = std::move(job);
      l.unlock();
      const int rc = fn();
      l.lock();
      result = rc;
      has_job = false;
      cv.notify_all();
    }
  }
  void submit(std::function<int()> fn) {
    std::lock_guard<std::mutex> l(m);
    job = std::move(fn);
    has_job = true;
    cv.notify_all();
  }
  int wait() {
    std::unique_lock<std::mutex> l(m);
    cv.wait(l, [this] { return !has_job; });
    return result;
  }
};

}  // namespace

struct kzg_gpu_group {
  int curve = -1;
  int M = 0, SB = 0;
  std::vector<int> devices;
  std::vector<kzg_gpu_ctx*> full;    // replica of the whole SRS (blob sharding, single-device calls)
  std::vector<kzg_gpu_ctx*> slice;   // the member's point range of the SRS, G1 only (point-range split); may be null
  std::vector<uint64_t> slice_lo, slice_n;
  std::vector<Worker*> workers;
  uint64_t srs_n = 0;
  uint64_t split_min = 65536;
  std::vector<std::pair<std::string, long long>> options;  // replayed on slice contexts created later
  std::string last_error;
  std::mutex mu;

  int size() const { return (int)devices.size(); }

  // run fn(i) on every member's worker; first failure wins
  int for_all(const std::function<int(int)>& fn, int members = -1) {
    const int g = members < 0 ? size() : members;
    for (int i = 1; i < g; i++) workers[i]->submit([&fn, i] { return fn(i); });
    int rc = fn(0);  // the caller's thread drives member 0
    int which = 0;
    for (int i = 1; i < g; i++) {
      const int r = workers[i]->wait();
      if (r && !rc) { rc = r; which = i; }
    }
    if (rc) note_error(which, rc);
    return rc;
  }
  void note_error(int member, int rc) {
    last_error = std::string("device ") + std::to_string(devices[member]) + ": " + kzg_gpu_strerror(rc);
    for (kzg_gpu_ctx* c : {full[member], member < (int)slice.size() ? slice[member] : nullptr})
      if (c && *kzg_gpu_last_error(c)) last_error += std::string(" [") + kzg_gpu_last_error(c) + "]";
  }
  void drop_slices() {
    for (kzg_gpu_ctx*& c : slice) {
      if (c) kzg_gpu_ctx_destroy(c);
      c = nullptr;
    }
  }
  bool want_slices(uint64_t n) const { return size() > 1 && n >= split_min && n >= (uint64_t)size(); }
  int make_slice_ctx(int i, uint64_t n) {
    const uint64_t g = (uint64_t)size();
    slice_lo[i] = n * (uint64_t)i / g;
    slice_n[i] = n * (uint64_t)(i + 1) / g - slice_lo[i];
    if (!slice[i]) {
      int rc = kzg_gpu_ctx_create(&slice[i], curve, devices[i]);
      if (rc) return rc;
      for (auto& kv : options) kzg_gpu_set_option(slice[i], kv.first.c_str(), kv.second);
      kzg_gpu_set_option(slice[i], "finish_affine", 0);  // a member leaves its projective partial: one inversion per MSM, after the fold
      kzg_gpu_set_option(slice[i], "setup_g2", 0);
    }
    return KZG_GPU_OK;
  }
};

extern "C" {

int kzg_gpu_group_create(kzg_gpu_group** out, int curve, const int* devices, int n_devices) {
  if (!out) return KZG_GPU_ERR_ARG;
  *out = nullptr;
  const int visible = kzg_gpu_device_count();
  if (visible <= 0) return KZG_GPU_ERR_CUDA;  // no CPU fallback
  if (kzg_gpu_modbytes(curve) < 0) return KZG_GPU_ERR_ARG;
  std::vector<int> devs;
  if (devices) {
    if (n_devices <= 0) return KZG_GPU_ERR_ARG;
    devs.assign(devices, devices + n_devices);
  } else {
    const int n = n_devices > 0 ? n_devices : visible;
    for (int i = 0; i < n; i++) devs.push_back(i);
  }
  // (an ordinal may appear more than once: members that share a device, which is how the one-GPU tests reach every
  // multi-member code path)
  for (size_t i = 0; i < devs.size(); i++)
    if (devs[i] < 0 || devs[i] >= visible) return KZG_GPU_ERR_ARG;
  if (devs.empty() || devs.size() > 64) return KZG_GPU_ERR_ARG;
  kzg_gpu_group* g = new (std::nothrow) kzg_gpu_group();
  if (!g) return KZG_GPU_ERR_MEMORY;
  g->curve = curve;
  g->M = kzg_gpu_modbytes(curve);
  g->SB = kzg_gpu_scalar_bytes(curve);
  g->devices = devs;
  g->full.assign(devs.size(), nullptr);
  g->slice.assign(devs.size(), nullptr);
  g->slice_lo.assign(devs.size(), 0);
  g->slice_n.assign(devs.size(), 0);
  // one context per device, created side by side on the worker threads: a CUDA context plus this library's module
  // costs 0.3-1 s per device, which a process that runs ONE operation (the reference's demo CLI) would pay serially
  for (size_t i = 0; i < devs.size(); i++) g->workers.push_back(i == 0 ? nullptr : new Worker());
  const int rc_create = g->for_all([&](int i) -> int { return kzg_gpu_ctx_create(&g->full[i], curve, devs[i]); });
  if (rc_create) {
    kzg_gpu_group_destroy(g);
    return rc_create;
  }
  *out = g;
  return KZG_GPU_OK;
}

void kzg_gpu_group_destroy(kzg_gpu_group* g) {
  if (!g) return;
  for (Worker* w : g->workers) delete w;
  for (kzg_gpu_ctx* c : g->full)
    if (c) kzg_gpu_ctx_destroy(c);
  g->drop_slices();
  delete g;
}

int kzg_gpu_group_size(const kzg_gpu_group* g) { return g ? g->size() : 0; }
kzg_gpu_ctx* kzg_gpu_group_ctx(kzg_gpu_group* g, int i) { return (g && i >= 0 && i < g->size()) ? g->full[i] : nullptr; }
const char* kzg_gpu_group_last_error(const kzg_gpu_group* g) { return g ? g->last_error.c_str() : ""; }

int kzg_gpu_group_set_option(kzg_gpu_group* g, const char* name, long long value) {
  if (!g || !name) return KZG_GPU_ERR_ARG;
  std::lock_guard<std::mutex> lock(g->mu);
  if (std::string(name) == "split_min") {
    if (value < 1) return KZG_GPU_ERR_ARG;
    g->split_min = (uint64_t)value;
    return KZG_GPU_OK;
  }
  for (kzg_gpu_ctx* c : g->full) {
    int rc = kzg_gpu_set_option(c, name, value);
    if (rc) return rc;
  }
  if (std::string(name) != "finish_affine" && std::string(name) != "setup_g2") {
    g->options.emplace_back(name, value);
    for (kzg_gpu_ctx* c : g->slice)
      if (c) kzg_gpu_set_option(c, name, value);
  }
  return KZG_GPU_OK;
}

int kzg_gpu_group_srs_load(kzg_gpu_group* g, uint64_t n, const uint8_t* g1, const uint8_t* g2) {
  if (!g || !g1 || n == 0) return KZG_GPU_ERR_ARG;
  std::lock_guard<std::mutex> lock(g->mu);
  g->srs_n = 0;
  const bool slices = g->want_slices(n);
  if (!slices) g->drop_slices();
  const int M = g->M;
  int rc = g->for_all([&](int i) -> int {
    int r = kzg_gpu_srs_load(g->full[i], n, g1, g2);
    if (r || !slices) return r;
    r = g->make_slice_ctx(i, n);
    if (r) return r;
    return kzg_gpu_srs_load(g->slice[i], g->slice_n[i], g1 + g->slice_lo[i] * (uint64_t)(2 * M), nullptr);
  });
  if (!rc) g->srs_n = n;
  return rc;
}

int kzg_gpu_group_setup_powers(kzg_gpu_group* g, const uint8_t* s_le, uint64_t n, uint8_t* g1_out, uint8_t* g2_out) {
  if (!g || !s_le || n == 0) return KZG_GPU_ERR_ARG;
  std::lock_guard<std::mutex> lock(g->mu);
  g->srs_n = 0;
  const bool slices = g->want_slices(n);
  if (!slices) g->drop_slices();
  int rc = g->for_all([&](int i) -> int {
    int r = kzg_gpu_setup_powers(g->full[i], s_le, n, i == 0 ? g1_out : nullptr, i == 0 ? g2_out : nullptr);
    if (r || !slices) return r;
    r = g->make_slice_ctx(i, n);
    if (r) return r;
    return kzg_gpu_setup_powers_range(g->slice[i], s_le, g->slice_lo[i], g->slice_n[i], nullptr, nullptr);
  });
  if (!rc) g->srs_n = n;
  return rc;
}

int kzg_gpu_group_msm_g1(kzg_gpu_group* g, const uint8_t* scalars, uint64_t count, uint64_t base_offset, uint8_t* out_xy,
                         int* is_inf) {
  if (!g) return KZG_GPU_ERR_ARG;
  std::lock_guard<std::mutex> lock(g->mu);
  if (g->srs_n == 0) return KZG_GPU_ERR_STATE;
  if (base_offset + count > g->srs_n || base_offset + count < count) return KZG_GPU_ERR_ARG;
  const int G = g->size();
  if (G == 1 || count < g->split_min || !g->slice[0]) {
    int rc = kzg_gpu_msm_g1(g->full[0], scalars, count, base_offset, out_xy, is_inf);
    if (rc) g->note_error(0, rc);
    return rc;
  }
  // point-range split: member i takes the part of [base_offset, base_offset + count) that lies in its slice
  size_t n_words = 0;
  kzg_gpu_msm_fetch_partial(g->slice[0], nullptr, &n_words);
  std::vector<uint32_t> partials((size_t)G * n_words, 0u);
  std::vector<int> used(G, 0);
  const int SB = g->SB;
  int rc = g->for_all([&](int i) -> int {
    const uint64_t lo = std::max(base_offset, g->slice_lo[i]);
    const uint64_t hi = std::min(base_offset + count, g->slice_lo[i] + g->slice_n[i]);
    if (hi <= lo) return KZG_GPU_OK;
    int inf = 0;
    int r = kzg_gpu_msm_g1(g->slice[i], scalars + (lo - base_offset) * (uint64_t)SB, hi - lo, lo - g->slice_lo[i], nullptr, &inf);
    if (r) return r;
    size_t nw = 0;
    r = kzg_gpu_msm_fetch_partial(g->slice[i], partials.data() + (size_t)i * n_words, &nw);
    used[i] = 1;
    return r;
  });
  if (rc) return rc;
  std::vector<uint32_t> packed;
  uint64_t np = 0;
  for (int i = 0; i < G; i++)
    if (used[i]) {
      packed.insert(packed.end(), partials.begin() + (size_t)i * n_words, partials.begin() + (size_t)(i + 1) * n_words);
      np++;
    }
  rc = kzg_gpu_g1_sum_partials(g->full[0], packed.data(), np, out_xy, is_inf);
  if (rc) g->note_error(0, rc);
  return rc;
}

static int shard_batch(kzg_gpu_group* g, uint64_t batch, const std::function<int(int, uint64_t, uint64_t)>& run) {
  const uint64_t G = (uint64_t)g->size();
  const int members = (int)std::min<uint64_t>(G, batch);
  return g->for_all([&](int i) -> int {
    const uint64_t b0 = batch * (uint64_t)i / (uint64_t)members, b1 = batch * (uint64_t)(i + 1) / (uint64_t)members;
    return b1 > b0 ? run(i, b0, b1) : KZG_GPU_OK;
  }, members);
}

int kzg_gpu_group_msm_g1_batch(kzg_gpu_group* g, const uint8_t* scalars, uint64_t count, uint64_t batch, uint8_t* out_xy,
                               int* is_inf) {
  if (!g || batch == 0) return KZG_GPU_ERR_ARG;
  std::lock_guard<std::mutex> lock(g->mu);
  if (g->srs_n == 0) return KZG_GPU_ERR_STATE;
  const uint64_t SB = (uint64_t)g->SB, PB = (uint64_t)(2 * g->M);
  return shard_batch(g, batch, [&](int i, uint64_t b0, uint64_t b1) -> int {
    return kzg_gpu_msm_g1_batch(g->full[i], scalars ? scalars + b0 * count * SB : nullptr, count, b1 - b0,
                                out_xy ? out_xy + b0 * PB : nullptr, is_inf ? is_inf + b0 : nullptr);
  });
}

int kzg_gpu_group_prove_batch(kzg_gpu_group* g, const uint8_t* coeffs, uint64_t n, const int64_t* first_roots, uint64_t k,
                              uint64_t batch, uint8_t* out_xy, int* is_inf) {
  if (!g || batch == 0 || !coeffs || !first_roots) return KZG_GPU_ERR_ARG;
  std::lock_guard<std::mutex> lock(g->mu);
  if (g->srs_n == 0) return KZG_GPU_ERR_STATE;
  const uint64_t SB = (uint64_t)g->SB, PB = (uint64_t)(2 * g->M);
  return shard_batch(g, batch, [&](int i, uint64_t b0, uint64_t b1) -> int {
    return kzg_gpu_prove_batch(g->full[i], coeffs + b0 * n * SB, n, first_roots + b0, k, b1 - b0,
                               out_xy ? out_xy + b0 * PB : nullptr, is_inf ? is_inf + b0 : nullptr);
  });
}

}  // extern "C"
