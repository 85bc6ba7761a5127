// host_pipeline.cu — b2pt_render_host: Technique::render through HOST memory, at the reference's contract.
//
// The reference's caller (Application::render, Application.cpp:41-79; Framework::runBatch, framework.cpp:426-437) asks for ONE pass of
// one sample per pixel per call and expects it added into its dvec4 view when the call returns.  One pass of 1024 x 1024 is a batch of
// 2^20 paths: too small to fill a B200 (the wavefront stages run at half their rate on batches of that size), and a
// 32 MB film crossing PCIe per call on top.  Passes are pure functions of (seed, sample index), so the library renders AHEAD of its
// caller:
//
//   * a call for pass j of a sequence (same parameters, consecutive sample indices) is served from a batch of per-pass films
//     [j, j + L) rendered in one go; L starts at half of 2^24 / pixels passes (or what the call asks for, if that is more) and
//     doubles per batch up to 2^24 / pixels — the batch size the integrators are efficient at — so a short run wastes little;
//   * while the caller consumes batch A — device->host copy in chunks on a copy stream, each chunk added into the view by a pool of
//     persistent host threads as it lands — a worker thread already renders batch B = [j + L, j + 2L) on its own streams;
//   * a call whose parameters do not continue the sequence drops what was rendered ahead and starts a new sequence.
//
// What a call returns is exactly what the synchronous path returned: the passes with the sample indices it named.  Rendering ahead
// costs at most 2 L passes of GPU time at the end of a run.  B2PT_HOST_LOOKAHEAD=0 turns it off (every call renders its own passes).
#include <algorithm>
#include <atomic>
#include <condition_variable>
#include <functional>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

#include "render_common.cuh"

namespace b2pt {

int render_passes(b2pt_scene_s* scene, const B2ptRenderParams* params, double* accum_dev, size_t pass_stride, float* path_radiance_dev,
                  cudaStream_t stream, B2ptStats* stats);

namespace {

// out[i] = (accumulate ? out[i] : 0) + sum over k < m of films[k * stride + i]
__global__ void __launch_bounds__(256) film_sum_kernel(const double* __restrict__ films, uint32_t m, size_t stride, size_t n, double* out, int accumulate) {
  for (size_t i = size_t(blockIdx.x) * blockDim.x + threadIdx.x; i < n; i += size_t(gridDim.x) * blockDim.x) {
    double s = accumulate ? out[i] : 0.0;
    for (uint32_t k = 0; k < m; ++k) s += films[size_t(k) * stride + i];
    out[i] = s;
  }
}

// Persistent host threads that run the tasks 0 .. n-1 of one parallel loop at a time (the caller's thread takes part).
class TaskPool {
 public:
  explicit TaskPool(int workers) {
    for (int i = 0; i < workers; ++i) threads_.emplace_back([this] { loop(); });
  }
  ~TaskPool() {
    { std::lock_guard<std::mutex> g(m_); quit_ = true; }
    cv_.notify_all();
    for (auto& t : threads_) t.join();
  }
  template <class F>
  void run(int n, F&& f) {
    fn_ = [&](int i) { f(i); };
    {
      std::lock_guard<std::mutex> g(m_);
      n_ = n; next_.store(0); done_.store(0); ++generation_;
    }
    cv_.notify_all();
    work();
    std::unique_lock<std::mutex> g(m_);
    done_cv_.wait(g, [&] { return done_.load() >= n_; });
    n_ = 0;
  }

 private:
  void work() {
    for (;;) {
      const int i = next_.fetch_add(1);
      if (i >= n_) return;
      fn_(i);
      if (done_.fetch_add(1) + 1 >= n_) { std::lock_guard<std::mutex> g(m_); done_cv_.notify_all(); }
    }
  }
  void loop() {
    uint64_t seen = 0;
    for (;;) {
      {
        std::unique_lock<std::mutex> g(m_);
        cv_.wait(g, [&] { return quit_ || generation_ != seen; });
        if (quit_) return;
        seen = generation_;
      }
      work();
    }
  }
  std::vector<std::thread> threads_;
  std::mutex m_;
  std::condition_variable cv_, done_cv_;
  std::function<void(int)> fn_;
  std::atomic<int> next_{0}, done_{0};
  int n_ = 0;
  uint64_t generation_ = 0;
  bool quit_ = false;
};

struct PassBatch {
  double* films = nullptr;         // capacity * film_doubles
  uint32_t start = 0, passes = 0;  // sequence positions [start, start + passes)
  enum State { kEmpty, kRendering, kReady } state = kEmpty;
  B2ptStats stats{};
  int rc = B2PT_OK;
  std::string error;
};

constexpr int kChunks = 8;

struct HostPipeline {
  b2pt_scene_s* scene = nullptr;
  int device = 0;
  B2ptRenderParams key{};  // the sequence: first_sample = sample index of position 0; num_samples unused
  bool has_key = false;
  uint32_t cursor = 0;     // next position the caller will ask for
  size_t film_doubles = 0;
  uint32_t capacity = 0;   // films per batch
  uint32_t lookahead = 0;  // passes per batch rendered ahead (0 = off)
  uint32_t last_batch = 0; // size of the batch launched last in this sequence: batches double from what the caller asks for up to `lookahead`,
                           // so a short run wastes at most about as many passes as it consumed
  PassBatch batch[2];
  double* stage_dev = nullptr;
  double* pinned = nullptr;
  cudaStream_t render_stream = nullptr, copy_stream = nullptr;
  cudaEvent_t chunk_event[kChunks] = {};
  std::thread worker;
  std::mutex m;
  std::condition_variable cv;
  int job = -1;  // batch the worker should render
  bool quit = false;
  TaskPool* adders = nullptr;

  void worker_loop() {
    cudaSetDevice(device);
    for (;;) {
      int b;
      {
        std::unique_lock<std::mutex> g(m);
        cv.wait(g, [&] { return quit || job >= 0; });
        if (quit) return;
        b = job;
      }
      PassBatch& pb = batch[b];
      B2ptRenderParams p = key;
      const uint32_t stride = key.sample_stride ? key.sample_stride : 1u;
      p.first_sample = key.first_sample + pb.start * stride;
      p.num_samples = pb.passes;
      int rc = B2PT_OK;
      cudaError_t e = cudaMemsetAsync(pb.films, 0, size_t(pb.passes) * film_doubles * sizeof(double), render_stream);
      if (e != cudaSuccess) rc = cuda_fail(e, "cudaMemsetAsync(films)", __FILE__, __LINE__);
      B2ptStats st{};
      if (rc == B2PT_OK) rc = render_passes(scene, &p, pb.films, film_doubles, nullptr, render_stream, &st);
      if (rc == B2PT_OK) { e = cudaStreamSynchronize(render_stream); if (e != cudaSuccess) rc = cuda_fail(e, "cudaStreamSynchronize", __FILE__, __LINE__); }
      {
        std::lock_guard<std::mutex> g(m);
        pb.stats = st; pb.rc = rc;
        pb.error = rc == B2PT_OK ? std::string() : std::string(b2pt_last_error());
        pb.state = PassBatch::kReady;
        job = -1;
      }
      cv.notify_all();
    }
  }

  // no batch is being rendered when this returns
  void quiesce() {
    std::unique_lock<std::mutex> g(m);
    cv.wait(g, [&] { return job < 0; });
  }

  void release_buffers() {
    for (auto& b : batch) { if (b.films) cudaFree(b.films); b.films = nullptr; b.state = PassBatch::kEmpty; }
    if (stage_dev) cudaFree(stage_dev);
    if (pinned) cudaFreeHost(pinned);
    stage_dev = nullptr; pinned = nullptr; capacity = 0; film_doubles = 0;
  }

  ~HostPipeline() {
    { std::lock_guard<std::mutex> g(m); }
    quiesce();
    { std::lock_guard<std::mutex> g(m); quit = true; }
    cv.notify_all();
    if (worker.joinable()) worker.join();
    delete adders;
    release_buffers();
    for (auto& e : chunk_event) if (e) cudaEventDestroy(e);
    if (render_stream) cudaStreamDestroy(render_stream);
    if (copy_stream) cudaStreamDestroy(copy_stream);
  }
};

void pipeline_free(void* p) { delete (HostPipeline*)p; }
void pipeline_quiesce(b2pt_scene_s* s) { if (s->pipeline) ((HostPipeline*)s->pipeline)->quiesce(); }

bool same_sequence(const B2ptRenderParams& a, const B2ptRenderParams& b) {
  B2ptRenderParams x = a, y = b;
  x.first_sample = y.first_sample = 0;
  x.num_samples = y.num_samples = 0;
  return memcmp(&x, &y, sizeof x) == 0;
}

int pipeline_create(b2pt_scene_s* scene, HostPipeline** out) {
  HostPipeline* h = new HostPipeline();
  scene->pipeline = h; scene->pipeline_free = pipeline_free; scene->pipeline_quiesce = pipeline_quiesce;
  h->scene = scene;
  B2_CUDA(cudaGetDevice(&h->device));
  B2_CUDA(cudaStreamCreateWithFlags(&h->render_stream, cudaStreamNonBlocking));
  B2_CUDA(cudaStreamCreateWithFlags(&h->copy_stream, cudaStreamNonBlocking));
  for (auto& e : h->chunk_event) B2_CUDA(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
  int threads = int(std::min<unsigned>(8u, std::max(2u, std::thread::hardware_concurrency() / 2u)));
  { const char* e = getenv("B2PT_HOST_THREADS"); if (e) threads = std::max(1, std::min(64, atoi(e))); }
  h->adders = new TaskPool(threads - 1);
  h->worker = std::thread([h] { h->worker_loop(); });
  *out = h;
  return B2PT_OK;
}

int pipeline_buffers(HostPipeline* h, size_t film_doubles, uint32_t capacity) {
  if (h->film_doubles == film_doubles && h->capacity >= capacity) return B2PT_OK;
  h->quiesce();
  h->release_buffers();
  h->film_doubles = film_doubles; h->capacity = capacity;
  for (auto& b : h->batch) B2_CUDA(cudaMalloc(&b.films, size_t(capacity) * film_doubles * sizeof(double)));
  B2_CUDA(cudaMalloc(&h->stage_dev, film_doubles * sizeof(double)));
  B2_CUDA(cudaMallocHost((void**)&h->pinned, film_doubles * sizeof(double)));
  return B2PT_OK;
}

// hands batch `b` = [start, start + passes) to the worker; the caller holds no lock
void pipeline_launch(HostPipeline* h, int b, uint32_t start, uint32_t passes) {
  {
    std::lock_guard<std::mutex> g(h->m);
    PassBatch& pb = h->batch[b];
    pb.start = start; pb.passes = passes; pb.state = PassBatch::kRendering; pb.rc = B2PT_OK; pb.error.clear();
    h->job = b;
  }
  h->cv.notify_all();
}

}  // namespace
}  // namespace b2pt

extern "C" int b2pt_render_host(b2pt_scene_t scene, const B2ptRenderParams* params, double* view_host, B2ptStats* stats) {
  using namespace b2pt;
  B2_CHECK(scene && params && view_host, B2PT_ERR_INVALID, "b2pt_render_host: null argument");
  B2_CHECK(params->width > 0 && params->height > 0, B2PT_ERR_INVALID, "b2pt_render_host: empty window");
  { int n = 0; if (cudaGetDeviceCount(&n) != cudaSuccess || n == 0) { cudaGetLastError(); B2_CHECK(false, B2PT_ERR_NO_DEVICE, "b2pt_render_host: no CUDA device (there is no CPU fallback)"); } }
  B2ptStats total{};
  if (params->num_samples == 0) {  // flush: wait for what is being rendered ahead and drop it
    if (HostPipeline* p = (HostPipeline*)scene->pipeline) {
      p->quiesce();
      for (auto& b : p->batch) b.state = PassBatch::kEmpty;
      p->has_key = false;
    }
    if (stats) *stats = total;
    return B2PT_OK;
  }
  HostPipeline* h = (HostPipeline*)scene->pipeline;
  if (!h) { if (int rc = pipeline_create(scene, &h)) return rc; }
  const size_t fd = size_t(params->width) * params->height * 4;
  const uint32_t n = params->num_samples;
  // passes per batch: what fills the GPU (2^24 paths), at least what this call asks for
  uint32_t look = uint32_t(std::min<size_t>(64, std::max<size_t>(1, (size_t(1) << 24) / (fd / 4))));
  { const char* e = getenv("B2PT_HOST_LOOKAHEAD"); if (e) look = uint32_t(std::max(0, std::min(256, atoi(e)))); }
  const uint32_t per_batch = std::max(look, n);
  if (int rc = pipeline_buffers(h, fd, per_batch)) return rc;
  h->lookahead = look;

  // ---- does the call continue the sequence that is being rendered ahead?
  const uint32_t stride = params->sample_stride ? params->sample_stride : 1u;
  const bool continues = h->has_key && same_sequence(h->key, *params) &&
                         params->first_sample == h->key.first_sample + h->cursor * stride;
  if (!continues) {
    h->quiesce();
    for (auto& b : h->batch) b.state = PassBatch::kEmpty;
    h->key = *params; h->has_key = true; h->cursor = 0; h->last_batch = 0;
  }

  // ---- gather the n passes [cursor, cursor + n) into `src` (a single film is read in place, several are summed on the device)
  const double* src = nullptr;
  uint32_t left = n, summed = 0;
  while (left) {
    int b = -1;
    {
      std::unique_lock<std::mutex> g(h->m);
      for (int k = 0; k < 2; ++k) {
        PassBatch& pb = h->batch[k];
        if (pb.state != PassBatch::kEmpty && h->cursor >= pb.start && h->cursor < pb.start + pb.passes) b = k;
      }
      if (b >= 0) h->cv.wait(g, [&] { return h->batch[b].state == PassBatch::kReady; });
    }
    if (b < 0) {  // nothing rendered (or being rendered) covers the cursor: render a batch that starts at it
      h->quiesce();
      int slot = h->batch[0].state == PassBatch::kEmpty ? 0 : 1;
      if (h->batch[slot].state != PassBatch::kEmpty) { h->batch[0].state = PassBatch::kEmpty; h->batch[1].state = PassBatch::kEmpty; slot = 0; }
      const uint32_t size = std::min(per_batch, std::max(std::max(left, (look + 1u) / 2u), 2u * h->last_batch));
      h->last_batch = size;
      pipeline_launch(h, slot, h->cursor, size);
      continue;
    }
    PassBatch& pb = h->batch[b];
    if (pb.rc != B2PT_OK) {
      const int rc = pb.rc;
      set_error(pb.error);
      pb.state = PassBatch::kEmpty; h->has_key = false;
      return rc;
    }
    // keep the GPU busy: the other slot renders what follows this batch while the caller drains it
    PassBatch& other = h->batch[b ^ 1];
    if (look > 0) {
      bool go;
      {
        std::lock_guard<std::mutex> g(h->m);
        const bool other_done = other.state == PassBatch::kEmpty || (other.state == PassBatch::kReady && other.start + other.passes <= h->cursor);
        go = h->job < 0 && other_done;
      }
      if (go) {
        const uint32_t size = std::min(per_batch, std::max(n, 2u * h->last_batch));
        h->last_batch = size;
        pipeline_launch(h, b ^ 1, pb.start + pb.passes, size);
      }
    }
    const uint32_t take = std::min(left, pb.start + pb.passes - h->cursor);
    const double* first = pb.films + size_t(h->cursor - pb.start) * fd;
    if (take == n && take == 1) {
      src = first;  // one pass: its film is read in place
    } else {        // several passes (possibly from two batches): summed on the device first
      film_sum_kernel<<<kNumSMs * 4, 256, 0, h->copy_stream>>>(first, take, fd, fd, h->stage_dev, summed ? 1 : 0);
      B2_CUDA(cudaGetLastError());
      src = h->stage_dev;
      summed += take;
    }
    // statistics of the call: the batch's, in proportion to the passes taken from it
    const double share = double(take) / double(pb.passes);
    total.num_basic_rays += uint64_t(double(pb.stats.num_basic_rays) * share);
    total.num_shadow_rays += uint64_t(double(pb.stats.num_shadow_rays) * share);
    total.numeric_errors += uint64_t(double(pb.stats.numeric_errors) * share + 0.5);
    total.kernel_launches += uint64_t(double(pb.stats.kernel_launches) * share);
    total.trace_launches += uint64_t(double(pb.stats.trace_launches) * share);
    total.shadow_launches += uint64_t(double(pb.stats.shadow_launches) * share);
    total.device_ms += float(pb.stats.device_ms * share);
    total.trace_ms += float(pb.stats.trace_ms * share); total.shadow_ms += float(pb.stats.shadow_ms * share); total.shade_ms += float(pb.stats.shade_ms * share);
    h->cursor += take;
    left -= take;
  }
  total.num_samples = n;
  total.num_paths = uint64_t(fd / 4) * n;

  // ---- device -> host in chunks on the copy stream; every chunk is added into the caller's view as it lands
  const size_t chunk = (fd + kChunks - 1) / kChunks;
  for (int c = 0; c < kChunks; ++c) {
    const size_t lo = std::min(fd, chunk * c), hi = std::min(fd, chunk * (c + 1));
    if (hi > lo) B2_CUDA(cudaMemcpyAsync(h->pinned + lo, src + lo, (hi - lo) * sizeof(double), cudaMemcpyDeviceToHost, h->copy_stream));
    B2_CUDA(cudaEventRecord(h->chunk_event[c], h->copy_stream));
  }
  std::atomic<int> failed{0};
  const double* pin = h->pinned;
  const int parts = 2;  // each chunk is added by two tasks: more tasks than threads keeps the pool busy while later chunks are in flight
  h->adders->run(kChunks * parts, [&](int t) {
    const int c = t / parts, k = t % parts;
    if (cudaEventSynchronize(h->chunk_event[c]) != cudaSuccess) { failed.store(1); return; }
    const size_t lo = std::min(fd, chunk * c), hi = std::min(fd, chunk * (c + 1));
    const size_t a = lo + (hi - lo) * k / parts, b = lo + (hi - lo) * (k + 1) / parts;
    for (size_t i = a; i < b; ++i) view_host[i] += pin[i];  // Technique::_commit_images adds into the caller's dvec4 buffer
  });
  B2_CHECK(!failed.load(), B2PT_ERR_CUDA, "b2pt_render_host: device->host copy of the film failed");
  if (stats) *stats = total;
  return B2PT_OK;
}
