// One handle over several GPUs of a box (SURVEY.md 8b "device_mask", 8e): streams are split into contiguous shards,
// one LPCNetBatch per GPU, one feeder thread per GPU that owns the device for the life of the handle.  Streams share
// only read-only weights (src/lpcnet.c:146-167 touches one LPCNetState), so there is no data-path collective: a call
// hands every feeder its rows of the caller's host buffers and returns when all of them are done.
#include <condition_variable>
#include <functional>
#include <memory>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

#include "../../include/lpcnet_b200.h"
#include "engine.h"

using namespace lpcb;

namespace {

struct Feeder {
  int device = 0, first = 0, count = 0;     // this GPU's shard: streams [first, first + count)
  LPCNetBatch *batch = nullptr;
  std::thread th;
  std::mutex mu;
  std::condition_variable cv;
  std::function<int()> job;                 // runs on the feeder thread, with its device current
  bool has_job = false, done = false, quit = false;
  int rc = 0;
  std::string err;
};

void feeder_main(Feeder *f) {
  cudaSetDevice(f->device);
  for (;;) {
    std::function<int()> job;
    {
      std::unique_lock<std::mutex> lk(f->mu);
      f->cv.wait(lk, [&] { return f->has_job || f->quit; });
      if (f->quit) return;
      job = f->job;
      f->has_job = false;
    }
    const int rc = job();
    std::string err = rc ? std::string(lpcnet_b200_last_error()) : std::string();   // this thread's error text
    {
      std::lock_guard<std::mutex> lk(f->mu);
      f->rc = rc; f->err = err; f->done = true;
    }
    f->cv.notify_all();
  }
}

}  // namespace

struct LPCNetMultiBatch {
  int n_streams = 0;
  std::vector<std::unique_ptr<Feeder>> feeders;
};

namespace {

// every feeder runs make_job(feeder); -1 with the first error text if any failed
int run_all(LPCNetMultiBatch *mb, const std::function<std::function<int()>(Feeder &)> &make_job) {
  for (auto &fp : mb->feeders) {
    Feeder &f = *fp;
    {
      std::lock_guard<std::mutex> lk(f.mu);
      f.job = make_job(f); f.has_job = true; f.done = false;
    }
    f.cv.notify_all();
  }
  int rc = 0;
  std::string err;
  for (auto &fp : mb->feeders) {
    Feeder &f = *fp;
    std::unique_lock<std::mutex> lk(f.mu);
    f.cv.wait(lk, [&] { return f.done; });
    if (f.rc && !rc) { rc = f.rc; err = "GPU " + std::to_string(f.device) + ": " + f.err; }
  }
  return rc ? set_last_error(err) : 0;
}

Feeder *owner_of(LPCNetMultiBatch *mb, int stream) {
  for (auto &fp : mb->feeders)
    if (stream >= fp->first && stream < fp->first + fp->count) return fp.get();
  return nullptr;
}

}  // namespace

extern "C" {

LPCNetMultiBatch *lpcnet_batch_create_multi(const unsigned char *blob, int len, int n_streams, unsigned device_mask,
                                            int features_delay, float lpc_gamma) {
  if (n_streams <= 0) { set_last_error("n_streams must be positive"); return nullptr; }
  int ndev = lpcnet_b200_device_count();
  if (ndev == 0) { set_last_error("no CUDA device available (liblpcnet_b200 has no CPU fallback)"); return nullptr; }
  if (ndev > 32) ndev = 32;
  std::vector<int> devs;
  for (int d = 0; d < ndev; d++)
    if (device_mask == 0 || (device_mask >> d) & 1u) devs.push_back(d);
  if (devs.empty()) { set_last_error("device_mask selects no visible CUDA device"); return nullptr; }
  if ((int)devs.size() > n_streams) devs.resize(n_streams);           // at least one stream per GPU
  LPCNetMultiBatch *mb = new LPCNetMultiBatch();
  mb->n_streams = n_streams;
  const int world = (int)devs.size(), base = n_streams / world, rem = n_streams % world;
  for (int k = 0; k < world; k++) {                                    // contiguous shards, sizes differ by at most 1
    std::unique_ptr<Feeder> f(new Feeder());
    f->device = devs[k];
    f->first = k * base + (k < rem ? k : rem);
    f->count = base + (k < rem ? 1 : 0);
    f->th = std::thread(feeder_main, f.get());
    mb->feeders.push_back(std::move(f));
  }
  const int rc = run_all(mb, [&](Feeder &f) {
    return std::function<int()>([&f, blob, len, features_delay, lpc_gamma]() {
      f.batch = lpcnet_batch_create(blob, len, f.count, f.device, features_delay, lpc_gamma);
      return f.batch ? 0 : -1;
    });
  });
  if (rc) {
    const std::string err = lpcnet_b200_last_error();
    lpcnet_multi_destroy(mb);
    set_last_error(err);
    return nullptr;
  }
  return mb;
}

void lpcnet_multi_destroy(LPCNetMultiBatch *mb) {
  if (!mb) return;
  run_all(mb, [](Feeder &f) {
    return std::function<int()>([&f]() {
      if (f.batch) lpcnet_batch_destroy(f.batch);
      f.batch = nullptr;
      return 0;
    });
  });
  for (auto &fp : mb->feeders) {
    { std::lock_guard<std::mutex> lk(fp->mu); fp->quit = true; }
    fp->cv.notify_all();
    if (fp->th.joinable()) fp->th.join();
  }
  delete mb;
}

int lpcnet_multi_n_streams(const LPCNetMultiBatch *mb) { return mb ? mb->n_streams : 0; }
int lpcnet_multi_n_devices(const LPCNetMultiBatch *mb) { return mb ? (int)mb->feeders.size() : 0; }

int lpcnet_multi_shard(const LPCNetMultiBatch *mb, int k, int *device, int *first_stream, int *n_streams) {
  if (!mb || k < 0 || k >= (int)mb->feeders.size()) return set_last_error("lpcnet_multi_shard: no such shard");
  if (device) *device = mb->feeders[k]->device;
  if (first_stream) *first_stream = mb->feeders[k]->first;
  if (n_streams) *n_streams = mb->feeders[k]->count;
  return 0;
}

int lpcnet_multi_set_seed(LPCNetMultiBatch *mb, int stream, const unsigned char *data, int n) {
  Feeder *f = mb ? owner_of(mb, stream) : nullptr;
  if (!f) return set_last_error("lpcnet_multi_set_seed: stream index out of range");
  return lpcnet_batch_set_seed(f->batch, stream - f->first, data, n);     // (lpcnet_batch_* select their device themselves)
}

int lpcnet_multi_reset(LPCNetMultiBatch *mb, int stream) {
  if (!mb) return set_last_error("null batch");
  if (stream < 0) return run_all(mb, [](Feeder &f) { return std::function<int()>([&f]() { return lpcnet_batch_reset(f.batch, -1); }); });
  Feeder *f = owner_of(mb, stream);
  if (!f) return set_last_error("lpcnet_multi_reset: stream index out of range");
  return lpcnet_batch_reset(f->batch, stream - f->first);
}

int lpcnet_multi_set_codebooks(LPCNetMultiBatch *mb, const float *codebooks, int n_floats) {
  if (!mb) return set_last_error("null batch");
  return run_all(mb, [=](Feeder &f) { return std::function<int()>([&f, codebooks, n_floats]() { return lpcnet_batch_set_codebooks(f.batch, codebooks, n_floats); }); });
}

int lpcnet_multi_synthesize(LPCNetMultiBatch *mb, const float *features, int feature_stride, int n_frames, short *pcm) {
  if (!mb) return set_last_error("null batch");
  if (n_frames <= 0) return 0;
  if (!features || !pcm) return set_last_error("null buffer");
  return run_all(mb, [=](Feeder &f) {
    return std::function<int()>([&f, features, feature_stride, n_frames, pcm]() {
      return lpcnet_batch_synthesize(f.batch, features + (size_t)f.first * n_frames * feature_stride, feature_stride, n_frames,
                                     pcm + (size_t)f.first * n_frames * 160, nullptr, nullptr, nullptr);
    });
  });
}

int lpcnet_multi_decode(LPCNetMultiBatch *mb, const unsigned char *packets, int n_packets, short *pcm) {
  if (!mb) return set_last_error("null batch");
  if (n_packets <= 0) return 0;
  if (!packets || !pcm) return set_last_error("null buffer");
  return run_all(mb, [=](Feeder &f) {
    return std::function<int()>([&f, packets, n_packets, pcm]() {
      return lpcnet_batch_decode(f.batch, packets + (size_t)f.first * n_packets * 8, n_packets, pcm + (size_t)f.first * n_packets * 640, nullptr);
    });
  });
}

}  // extern "C"
