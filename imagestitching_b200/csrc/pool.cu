// StitchPool: one process drives every GPU of a box -- one host thread + one StitchCtx per GPU, a shared job queue,
// no collective (independent stitch jobs: BASELINE configs[4]).  The reference draws one job at a time on
// one phone (isStitching, P:772,781); this is the job-level data parallelism the north star asks for, behind the same
// StitchJob / StitchResult the single-GPU entry points take.
#include <cuda_runtime.h>

#include <atomic>
#include <condition_variable>
#include <cstring>
#include <deque>
#include <memory>
#include <mutex>
#include <string>
#include <thread>
#include <unordered_map>
#include <vector>

#include "stitch.h"

namespace {

struct PoolJob {
  uint64_t ticket = 0;
  StitchJob job;
  std::vector<StitchImage> images;
  StitchResult* result = nullptr;
  int32_t device = -1;        // index into the pool's device list, -1 = any
  int kind = 0;               // 0 = stitch job, 1 = digest command
  const void* dig_ptr = nullptr;
  size_t dig_bytes = 0;
  uint64_t* dig_out = nullptr;
};

struct Done {
  int status = STITCH_OK;
  std::string error;
  int32_t device = -1;
};

}  // namespace

struct StitchPool {
  std::vector<int32_t> devices;
  std::vector<StitchCtx*> ctxs;
  std::vector<std::thread> workers;
  int32_t lanes = 2;
  std::mutex mu;
  std::condition_variable cv_work, cv_done;
  std::deque<std::unique_ptr<PoolJob>> any_q;                 // jobs any GPU may take (host buffers)
  std::vector<std::deque<std::unique_ptr<PoolJob>>> dev_q;    // jobs bound to one GPU (device buffers)
  std::unordered_map<uint64_t, Done> done;                    // finished, not yet waited for
  uint64_t next_ticket = 1, submitted = 0, finished = 0;
  bool stop = false;
  std::vector<int64_t> jobs_run;                              // per device, for reports / tests
  std::string last_error;
};

namespace {

void worker(StitchPool* P, int idx) {
  StitchCtx* ctx = P->ctxs[(size_t)idx];
  cudaSetDevice(P->devices[(size_t)idx]);
  struct InFlight { std::unique_ptr<PoolJob> pj; uint64_t ctx_ticket; };
  std::deque<InFlight> inflight;
  auto retire = [&](InFlight& f, int st) {
    Done d;
    d.status = st;
    d.device = idx;
    if (st != STITCH_OK) d.error = stitch_last_error(ctx);
    std::lock_guard<std::mutex> lk(P->mu);
    P->done[f.pj->ticket] = std::move(d);
    P->finished++;
    P->jobs_run[(size_t)idx]++;
    P->cv_done.notify_all();
  };
  for (;;) {
    std::unique_ptr<PoolJob> pj;
    {
      std::unique_lock<std::mutex> lk(P->mu);
      auto have = [&] { return !P->dev_q[(size_t)idx].empty() || !P->any_q.empty(); };
      if (inflight.empty()) P->cv_work.wait(lk, [&] { return P->stop || have(); });
      if (P->stop && !have() && inflight.empty()) return;
      if ((int)inflight.size() < P->lanes && have()) {
        auto& q = !P->dev_q[(size_t)idx].empty() ? P->dev_q[(size_t)idx] : P->any_q;
        pj = std::move(q.front());
        q.pop_front();
      }
    }
    if (pj) {
      if (pj->kind == 1) {   // digest: ordered after everything this worker has in flight
        while (!inflight.empty()) {
          retire(inflight.front(), stitch_wait(ctx, inflight.front().ctx_ticket));
          inflight.pop_front();
        }
        InFlight f{std::move(pj), 0};
        retire(f, stitch_device_digest(ctx, f.pj->dig_ptr, f.pj->dig_bytes, f.pj->dig_out));
        continue;
      }
      pj->job.images = pj->images.data();
      uint64_t t = 0;
      const int st = stitch_submit(ctx, &pj->job, pj->result, &t);
      InFlight f{std::move(pj), t};
      if (st != STITCH_OK) retire(f, st);
      else inflight.push_back(std::move(f));
      continue;
    }
    if (!inflight.empty()) {   // lanes full, or nothing queued: complete the oldest job
      retire(inflight.front(), stitch_wait(ctx, inflight.front().ctx_ticket));
      inflight.pop_front();
    }
  }
}

int enqueue(StitchPool* P, std::unique_ptr<PoolJob> pj, uint64_t* ticket) {
  std::lock_guard<std::mutex> lk(P->mu);
  if (P->stop) return STITCH_E_BUSY;
  pj->ticket = P->next_ticket++;
  if (ticket) *ticket = pj->ticket;
  P->submitted++;
  if (pj->device >= 0) {
    P->dev_q[(size_t)pj->device].push_back(std::move(pj));
    P->cv_work.notify_all();
  } else {
    P->any_q.push_back(std::move(pj));
    P->cv_work.notify_one();
  }
  return STITCH_OK;
}

}  // namespace

extern "C" int stitch_pool_create(const int32_t* devices, int32_t n_devices, int32_t lanes_per_device, StitchPool** out) {
  if (!out) return STITCH_E_BAD_ARG;
  *out = nullptr;
  setenv("CUDA_DEVICE_MAX_CONNECTIONS", "32", 0);   // see stitch_ctx_create
  int count = 0;
  if (cudaGetDeviceCount(&count) != cudaSuccess || count <= 0) return STITCH_E_CUDA;
  std::unique_ptr<StitchPool> P(new StitchPool());
  if (!devices || n_devices <= 0) {   // every GPU of the box
    for (int d = 0; d < count; ++d) P->devices.push_back(d);
  } else {
    for (int i = 0; i < n_devices; ++i) {
      if (devices[i] < 0 || devices[i] >= count) return STITCH_E_BAD_ARG;
      P->devices.push_back(devices[i]);
    }
  }
  P->lanes = lanes_per_device < 1 ? 2 : (lanes_per_device > STITCH_MAX_LANES ? STITCH_MAX_LANES : lanes_per_device);
  const size_t n = P->devices.size();
  P->dev_q.resize(n);
  P->jobs_run.assign(n, 0);
  for (size_t i = 0; i < n; ++i) {
    StitchCtx* c = nullptr;
    int st = stitch_ctx_create(P->devices[i], nullptr, &c);
    if (st == STITCH_OK) st = stitch_ctx_set_lanes(c, P->lanes);
    if (st != STITCH_OK) {
      if (c) stitch_ctx_destroy(c);
      for (StitchCtx* x : P->ctxs) stitch_ctx_destroy(x);
      return st;
    }
    P->ctxs.push_back(c);
  }
  StitchPool* raw = P.release();
  for (size_t i = 0; i < n; ++i) raw->workers.emplace_back(worker, raw, (int)i);
  *out = raw;
  return STITCH_OK;
}

extern "C" int stitch_pool_destroy(StitchPool* P) {
  if (!P) return STITCH_E_BAD_ARG;
  {
    std::lock_guard<std::mutex> lk(P->mu);
    P->stop = true;
    P->cv_work.notify_all();
  }
  for (std::thread& t : P->workers) t.join();   // workers drain their queues before they leave
  for (size_t i = 0; i < P->ctxs.size(); ++i) {
    cudaSetDevice(P->devices[i]);
    stitch_ctx_destroy(P->ctxs[i]);
  }
  delete P;
  return STITCH_OK;
}

extern "C" int32_t stitch_pool_device_count(StitchPool* P) { return P ? (int32_t)P->devices.size() : 0; }

extern "C" StitchCtx* stitch_pool_ctx(StitchPool* P, int32_t index) {
  return (P && index >= 0 && (size_t)index < P->ctxs.size()) ? P->ctxs[(size_t)index] : nullptr;
}

extern "C" int stitch_pool_submit(StitchPool* P, const StitchJob* job, StitchResult* result, int32_t device, uint64_t* ticket) {
  if (!P || !job || !result || !ticket) return STITCH_E_BAD_ARG;
  if (device >= (int32_t)P->devices.size()) return STITCH_E_BAD_ARG;
  if (job->n > 0 && !job->images) return STITCH_E_BAD_ARG;
  if (device < 0) {   // a job that may run anywhere must not name memory of one GPU
    if (job->out_mem_kind == STITCH_MEM_DEVICE) return STITCH_E_BAD_ARG;
    for (int i = 0; i < job->n; ++i)
      if (job->images[i].mem_kind == STITCH_MEM_DEVICE) return STITCH_E_BAD_ARG;
  }
  std::unique_ptr<PoolJob> pj(new PoolJob());
  pj->job = *job;
  if (job->n > 0) pj->images.assign(job->images, job->images + job->n);
  pj->result = result;
  pj->device = device < 0 ? -1 : device;
  return enqueue(P, std::move(pj), ticket);
}

extern "C" int stitch_pool_digest(StitchPool* P, int32_t device, const void* dev_ptr, size_t bytes, uint64_t* out, uint64_t* ticket) {
  if (!P || device < 0 || device >= (int32_t)P->devices.size() || !dev_ptr || !out || !ticket) return STITCH_E_BAD_ARG;
  std::unique_ptr<PoolJob> pj(new PoolJob());
  pj->kind = 1;
  pj->device = device;
  pj->dig_ptr = dev_ptr;
  pj->dig_bytes = bytes;
  pj->dig_out = out;
  return enqueue(P, std::move(pj), ticket);
}

extern "C" int stitch_pool_wait(StitchPool* P, uint64_t ticket, int32_t* device_out) {
  if (!P) return STITCH_E_BAD_ARG;
  std::unique_lock<std::mutex> lk(P->mu);
  if (ticket == 0 || ticket >= P->next_ticket) return STITCH_E_BAD_ARG;
  P->cv_done.wait(lk, [&] { return P->done.count(ticket) != 0; });
  Done d = std::move(P->done[ticket]);
  P->done.erase(ticket);
  if (device_out) *device_out = d.device;
  if (d.status != STITCH_OK) P->last_error = d.error;
  return d.status;
}

extern "C" int stitch_pool_wait_all(StitchPool* P) {
  if (!P) return STITCH_E_BAD_ARG;
  std::unique_lock<std::mutex> lk(P->mu);
  P->cv_done.wait(lk, [&] { return P->finished == P->submitted; });
  int st = STITCH_OK;
  for (auto& kv : P->done)
    if (kv.second.status != STITCH_OK && st == STITCH_OK) {
      st = kv.second.status;
      P->last_error = kv.second.error;
    }
  P->done.clear();
  return st;
}

extern "C" int64_t stitch_pool_jobs_run(StitchPool* P, int32_t index) {
  if (!P || index < 0 || (size_t)index >= P->jobs_run.size()) return -1;
  std::lock_guard<std::mutex> lk(P->mu);
  return P->jobs_run[(size_t)index];
}

extern "C" const char* stitch_pool_last_error(StitchPool* P) { return P ? P->last_error.c_str() : ""; }
