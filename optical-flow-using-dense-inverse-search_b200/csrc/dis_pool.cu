// dis_pool.cu — several GPUs of one box behind one handle (include/dis_b200.h, dis_pool_*).
//
// The reference's unit of independent work is one iteration of the per-pair loop, src/main.cpp:102: the body shares
// no state between pairs.  A pool is therefore N engines (one per GPU, each with its own streams and device memory)
// driven by N persistent host threads; a batch or a sequence is cut into N CONTIGUOUS blocks of pairs (sizes differ by
// at most one; neighbouring blocks of a sequence share one frame, so every GPU builds each of its pyramids once) and
// every block runs through the single-GPU entry points.  No collective, no peer traffic: results land in the
// caller's buffer directly from each GPU.  Each worker thread binds itself to the CPUs of its GPU's NUMA node before
// it creates its engine, so pinned staging memory it allocates (dis_pool_host_alloc) is local to that GPU's root port.
#include <cuda_runtime.h>
#include <pthread.h>
#include <sched.h>
#include <stdio.h>
#include <string.h>

#include <chrono>
#include <condition_variable>
#include <deque>
#include <functional>
#include <map>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

#include "../../include/dis_b200.h"

namespace {

thread_local std::string g_pool_err;
int pfail(int code, const std::string& msg) { g_pool_err = msg; return code; }

// CPUs of the NUMA node a CUDA device hangs off (empty: unknown / single node)
std::vector<int> device_cpus(int device) {
  std::vector<int> cpus;
  char bus[32] = {0};
  if (cudaDeviceGetPCIBusId(bus, sizeof bus, device) != cudaSuccess) { cudaGetLastError(); return cpus; }
  for (char* c = bus; *c; ++c) if (*c >= 'A' && *c <= 'Z') *c = (char)(*c - 'A' + 'a');
  char path[128];
  snprintf(path, sizeof path, "/sys/bus/pci/devices/%s/numa_node", bus);
  int node = -1;
  if (FILE* f = fopen(path, "r")) { if (fscanf(f, "%d", &node) != 1) node = -1; fclose(f); }
  if (node < 0) return cpus;
  snprintf(path, sizeof path, "/sys/devices/system/node/node%d/cpulist", node);
  FILE* f = fopen(path, "r");
  if (!f) return cpus;
  char list[4096] = {0};
  if (!fgets(list, sizeof list, f)) list[0] = 0;
  fclose(f);
  for (char* tok = strtok(list, ",\n"); tok; tok = strtok(nullptr, ",\n")) {
    int a = 0, b = 0;
    const int n = sscanf(tok, "%d-%d", &a, &b);
    if (n == 1) b = a;
    if (n >= 1) for (int c = a; c <= b; ++c) cpus.push_back(c);
  }
  return cpus;
}

struct Worker {
  int device = 0, numa_cpus = 0;
  dis_engine* eng = nullptr;
  std::thread th;
  std::mutex m;
  std::condition_variable cv;
  struct Task { std::function<int(dis_engine*)> fn; int ticket; };   // ticket < 0: a blocking call waits on `done`
  std::deque<Task> q;
  int backlog = 0;      // queued + running tasks (under m)
  dis_pool* pool = nullptr;
  bool done = false, quit = false;
  int rc = DIS_OK;
  std::string err;
  double ms = 0;       // wall time of the last job on this worker
  int lo = 0, hi = 0;  // pairs of the last job
};

}  // namespace

struct dis_pool {
  std::vector<Worker*> w;
  dis_params prm;
  double wall_ms = 0;
  bool peers_ready = false;
  // asynchronous submissions (dis_pool_submit_*): one record per ticket until it is waited for
  struct Ticket { bool done = false; int rc = DIS_OK; std::string err; int worker = -1; dis_stats stats; double ms = 0; };
  std::mutex tm;
  std::condition_variable tcv;
  std::map<int, Ticket> tickets;
  int next_ticket = 1, last_worker = -1;
};

namespace {

void worker_main(Worker* w, const dis_params* prm, bool bind) {
  if (bind) {
    const std::vector<int> cpus = device_cpus(w->device);
    if (!cpus.empty()) {
      cpu_set_t set;
      CPU_ZERO(&set);
      for (int c : cpus) if (c < CPU_SETSIZE) CPU_SET(c, &set);
      if (pthread_setaffinity_np(pthread_self(), sizeof set, &set) == 0) w->numa_cpus = (int)cpus.size();
    }
  }
  int rc = dis_create(prm, w->device, &w->eng);
  {
    std::lock_guard<std::mutex> g(w->m);
    w->rc = rc;
    if (rc) w->err = dis_last_error();
    w->done = true;
  }
  w->cv.notify_all();
  if (rc) return;
  for (;;) {
    Worker::Task task;
    {
      std::unique_lock<std::mutex> g(w->m);
      w->cv.wait(g, [&] { return !w->q.empty() || w->quit; });
      if (w->q.empty()) break;          // quit: only once the queue has drained
      task = std::move(w->q.front());
      w->q.pop_front();
    }
    const auto t0 = std::chrono::steady_clock::now();
    rc = task.fn(w->eng);
    const double ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
    if (task.ticket >= 0) {             // asynchronous submission: the result goes to the ticket
      dis_stats st;
      memset(&st, 0, sizeof st);
      const std::string err = rc ? dis_last_error() : "";
      if (!rc) dis_get_stats(w->eng, &st);
      { std::lock_guard<std::mutex> g(w->m); --w->backlog; }
      {
        std::lock_guard<std::mutex> g(w->pool->tm);
        dis_pool::Ticket& t = w->pool->tickets[task.ticket];
        t.done = true; t.rc = rc; t.err = err; t.stats = st; t.ms = ms;
      }
      w->pool->tcv.notify_all();
      continue;
    }
    {
      std::lock_guard<std::mutex> g(w->m);
      --w->backlog;
      w->rc = rc;
      w->err = rc ? dis_last_error() : "";
      w->ms = ms;
      w->done = true;
    }
    w->cv.notify_all();
  }
  dis_destroy(w->eng);
  w->eng = nullptr;
}

void submit(Worker* w, std::function<int(dis_engine*)> job, int ticket = -1) {
  {
    std::lock_guard<std::mutex> g(w->m);
    w->q.push_back(Worker::Task{std::move(job), ticket});
    ++w->backlog;
    if (ticket < 0) w->done = false;
  }
  w->cv.notify_all();
}

// asynchronous submission: the whole call goes to the worker with the shortest queue (ties: the one after the last
// used), so equal jobs alternate between the engines
int submit_ticket(dis_pool* p, std::function<int(dis_engine*)> job, int* ticket) {
  const int R = (int)p->w.size();
  int best = -1, best_n = 0;
  for (int k = 1; k <= R; ++k) {
    const int i = (p->last_worker + k) % R;
    int n;
    { std::lock_guard<std::mutex> g(p->w[i]->m); n = p->w[i]->backlog; }
    if (best < 0 || n < best_n) { best = i; best_n = n; }
  }
  p->last_worker = best;
  int t;
  {
    std::lock_guard<std::mutex> g(p->tm);
    t = p->next_ticket++;
    p->tickets[t].worker = best;
  }
  submit(p->w[best], std::move(job), t);
  *ticket = t;
  return DIS_OK;
}

int wait_done(Worker* w) {
  std::unique_lock<std::mutex> g(w->m);
  w->cv.wait(g, [&] { return w->done; });
  return w->rc;
}

// run one job per worker (null: none), wait for all; first failure wins
int run_all(dis_pool* p, const std::vector<std::function<int(dis_engine*)>>& jobs) {
  const auto t0 = std::chrono::steady_clock::now();
  for (size_t i = 0; i < p->w.size(); ++i) {
    p->w[i]->ms = 0;
    if (jobs[i]) submit(p->w[i], jobs[i]);
  }
  int rc = DIS_OK;
  std::string err;
  for (size_t i = 0; i < p->w.size(); ++i)
    if (jobs[i]) {
      const int r = wait_done(p->w[i]);
      if (r && !rc) { rc = r; err = "GPU " + std::to_string(p->w[i]->device) + ": " + p->w[i]->err; }
    }
  p->wall_ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
  return rc ? pfail(rc, err) : DIS_OK;
}

}  // namespace

extern "C" {

const char* dis_pool_last_error(void) { return g_pool_err.c_str(); }

void dis_pool_shard(int n_pairs, int parts, int part, int* lo, int* hi) {
  if (parts < 1 || part < 0 || part >= parts || n_pairs < 0) { if (lo) *lo = 0; if (hi) *hi = 0; return; }
  const int q = n_pairs / parts, r = n_pairs % parts;
  const int l = part * q + (part < r ? part : r);
  if (lo) *lo = l;
  if (hi) *hi = l + q + (part < r ? 1 : 0);
}

int dis_pool_create(const dis_params* p, const int* devices, int n_devices, dis_pool** out) {
  if (!p || !out || n_devices < 1 || n_devices > 64) return pfail(DIS_ERR_PARAM, "dis_pool_create: need 1..64 devices");
  *out = nullptr;
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess || n <= 0) {
    cudaGetLastError();
    return pfail(DIS_ERR_CUDA, "no CUDA device: this engine has no CPU path");
  }
  dis_pool* pool = new dis_pool();
  pool->prm = *p;
  const bool bind = getenv("DIS_POOL_NO_AFFINITY") == nullptr;
  for (int i = 0; i < n_devices; ++i) {
    Worker* w = new Worker();
    w->device = devices ? devices[i] : i;
    w->pool = pool;
    pool->w.push_back(w);
    w->th = std::thread(worker_main, w, &pool->prm, bind);
  }
  int rc = DIS_OK;
  std::string err;
  for (Worker* w : pool->w) {
    const int r = wait_done(w);
    if (r && !rc) { rc = r; err = "GPU " + std::to_string(w->device) + ": " + w->err; }
  }
  if (rc) { dis_pool_destroy(pool); return pfail(rc, err); }
  // Engines that share a GPU hide each other's copies, so the output slices of the host-buffer calls (each costs a launch
  // tail at the finest level) only cost GPU time: 31.1 against 32.0 ms per 49-pair step on one B200 (bench.py, 10 steps).
  // What stays exposed without them is the last download of a run (3 ms at 57 GB/s, 15 ms where eight GPUs share the
  // host links at 11 GB/s): DIS_POOL_SLICES=1 keeps the slices for short runs on slow links.
  if (!getenv("DIS_POOL_SLICES"))
    for (Worker* w : pool->w) {
      int same = 0;
      for (Worker* v : pool->w) same += v->device == w->device;
      if (same > 1) dis_set_host_slices(w->eng, 0);
    }
  *out = pool;
  return DIS_OK;
}

void dis_pool_destroy(dis_pool* p) {
  if (!p) return;
  for (Worker* w : p->w) {
    { std::lock_guard<std::mutex> g(w->m); w->quit = true; }
    w->cv.notify_all();
    if (w->th.joinable()) w->th.join();
    delete w;
  }
  delete p;
}

int dis_pool_size(const dis_pool* p) { return p ? (int)p->w.size() : 0; }

int dis_pool_device(const dis_pool* p, int i, int* device, int* numa_cpus) {
  if (!p || i < 0 || i >= (int)p->w.size()) return pfail(DIS_ERR_PARAM, "worker index out of range");
  if (device) *device = p->w[i]->device;
  if (numa_cpus) *numa_cpus = p->w[i]->numa_cpus;
  return DIS_OK;
}

int dis_pool_run_batch_u8(dis_pool* p, const uint8_t* const* a, const uint8_t* const* b, int n_pairs, int w, int h,
                          int stride, float* flow_out) {
  if (!p || !a || !b || !flow_out) return pfail(DIS_ERR_PARAM, "null argument");
  if (n_pairs <= 0 || w <= 0 || h <= 0 || stride < w) return pfail(DIS_ERR_PARAM, "bad batch geometry");
  const int R = (int)p->w.size();
  std::vector<std::function<int(dis_engine*)>> jobs(R);
  for (int r = 0; r < R; ++r) {
    int lo, hi;
    dis_pool_shard(n_pairs, R, r, &lo, &hi);
    p->w[r]->lo = lo; p->w[r]->hi = hi;
    if (hi > lo)
      jobs[r] = [=](dis_engine* e) {
        return dis_run_batch_u8(e, a + lo, b + lo, hi - lo, w, h, stride, flow_out + (size_t)lo * 2 * w * h);
      };
  }
  return run_all(p, jobs);
}

int dis_pool_run_sequence_u8(dis_pool* p, const uint8_t* frames, int n_frames, int w, int h, int stride,
                             size_t frame_stride, float* flow_out) {
  if (!p || !frames || !flow_out) return pfail(DIS_ERR_PARAM, "null argument");
  if (n_frames < 2 || w <= 0 || h <= 0 || stride < w || frame_stride < (size_t)stride * h)
    return pfail(DIS_ERR_PARAM, "bad sequence geometry");
  const int R = (int)p->w.size();
  std::vector<std::function<int(dis_engine*)>> jobs(R);
  for (int r = 0; r < R; ++r) {
    int lo, hi;
    dis_pool_shard(n_frames - 1, R, r, &lo, &hi);
    p->w[r]->lo = lo; p->w[r]->hi = hi;
    if (hi > lo)   // pairs lo .. hi-1 need frames lo .. hi: one frame shared with the next block
      jobs[r] = [=](dis_engine* e) {
        return dis_run_sequence_u8(e, frames + frame_stride * (size_t)lo, hi - lo + 1, w, h, stride, frame_stride,
                                   flow_out + (size_t)lo * 2 * w * h);
      };
  }
  return run_all(p, jobs);
}

// ---- asynchronous form: whole calls queued to the engines ---------------------------------------------------------
// A submission is ONE dis_run_sequence_u8 / dis_run_batch_u8 on ONE engine (not sharded): the pool hands it to the
// worker with the shortest queue and returns a ticket at once.  With a GPU listed twice in dis_pool_create, two
// consecutive submissions run on two engines of that GPU: the uploads and downloads of one overlap with the kernels of
// the other, and the latency-bound level tails of one are filled by the other's blocks.
int dis_pool_submit_sequence_u8(dis_pool* p, const uint8_t* frames, int n_frames, int w, int h, int stride,
                                size_t frame_stride, float* flow_out, int* ticket) {
  if (!p || !frames || !flow_out || !ticket) return pfail(DIS_ERR_PARAM, "null argument");
  if (n_frames < 2 || w <= 0 || h <= 0 || stride < w || frame_stride < (size_t)stride * h)
    return pfail(DIS_ERR_PARAM, "bad sequence geometry");
  return submit_ticket(p, [=](dis_engine* e) { return dis_run_sequence_u8(e, frames, n_frames, w, h, stride, frame_stride, flow_out); }, ticket);
}

int dis_pool_submit_batch_u8(dis_pool* p, const uint8_t* const* a, const uint8_t* const* b, int n_pairs, int w, int h,
                             int stride, float* flow_out, int* ticket) {
  if (!p || !a || !b || !flow_out || !ticket) return pfail(DIS_ERR_PARAM, "null argument");
  if (n_pairs <= 0 || w <= 0 || h <= 0 || stride < w) return pfail(DIS_ERR_PARAM, "bad batch geometry");
  std::vector<const uint8_t*> pa(a, a + n_pairs), pb(b, b + n_pairs);   // the pointer arrays may be temporaries of the caller
  return submit_ticket(p, [=](dis_engine* e) { return dis_run_batch_u8(e, pa.data(), pb.data(), n_pairs, w, h, stride, flow_out); }, ticket);
}

int dis_pool_wait(dis_pool* p, int ticket, dis_stats* stats, int* worker) {
  if (!p) return pfail(DIS_ERR_PARAM, "null pool");
  std::unique_lock<std::mutex> g(p->tm);
  auto it = p->tickets.find(ticket);
  if (it == p->tickets.end()) return pfail(DIS_ERR_STATE, "unknown ticket (never issued, or already waited for)");
  p->tcv.wait(g, [&] { return it->second.done; });
  const dis_pool::Ticket t = it->second;
  p->tickets.erase(it);
  g.unlock();
  if (stats) *stats = t.stats;
  if (worker) *worker = t.worker;
  return t.rc ? pfail(t.rc, "GPU " + std::to_string(p->w[t.worker]->device) + ": " + t.err) : DIS_OK;
}

int dis_pool_wait_all(dis_pool* p) {
  if (!p) return pfail(DIS_ERR_PARAM, "null pool");
  int rc = DIS_OK;
  std::string err;
  for (;;) {
    int t;
    { std::lock_guard<std::mutex> g(p->tm); if (p->tickets.empty()) break; t = p->tickets.begin()->first; }
    const int r = dis_pool_wait(p, t, nullptr, nullptr);
    if (r && !rc) { rc = r; err = g_pool_err; }
  }
  return rc ? pfail(rc, err) : DIS_OK;
}

// ---- one large pair over the pool's GPUs: bands of grid rows + halo exchange -----------------------
// Needs of a band follow from the reference's geometry: a pixel row y of a level is covered by the grid rows
// ceil((y - ps/2 + 1 - offh) / steps) .. floor((y + ps/2 - offh) / steps) (src/patch_grid.cpp:121-160), and a patch with
// reference row ry is initialised from row floor(ry / 2) of the coarser dense flow (src/patch_grid.cpp:112-117).
namespace {
int ceil_div(int a, int b) { return a >= 0 ? (a + b - 1) / b : -((-a) / b); }
int floor_div(int a, int b) { return a >= 0 ? a / b : -((-a + b - 1) / b); }
}  // namespace

int dis_pool_run_pair_tiled_u8(dis_pool* p, const uint8_t* a, const uint8_t* b, int w, int h, int stride, float* flow_out) {
  if (!p || !a || !b || !flow_out) return pfail(DIS_ERR_PARAM, "null argument");
  if (w <= 0 || h <= 0 || stride < w) return pfail(DIS_ERR_PARAM, "bad frame geometry");
  if (p->prm.finest_scale != 0) return pfail(DIS_ERR_PARAM, "the row-banded path needs finest_scale == 0");
  const int R = (int)p->w.size();
  const auto t0 = std::chrono::steady_clock::now();
  struct DeviceGuard {   // this function issues work to every GPU from the caller's thread: leave its current device as it was
    int dev = -1;
    DeviceGuard() { if (cudaGetDevice(&dev) != cudaSuccess) { cudaGetLastError(); dev = -1; } }
    ~DeviceGuard() { if (dev >= 0) cudaSetDevice(dev); }
  } guard;
  // peer access between every two distinct devices of the pool (once)
  if (!p->peers_ready) {
    for (int i = 0; i < R; ++i)
      for (int j = 0; j < R; ++j) {
        const int di = p->w[i]->device, dj = p->w[j]->device;
        if (di == dj) continue;
        int can = 0;
        if (cudaDeviceCanAccessPeer(&can, di, dj) != cudaSuccess || !can) {
          cudaGetLastError();
          return pfail(DIS_ERR_CUDA, "GPU " + std::to_string(di) + " cannot access GPU " + std::to_string(dj) + " as a peer");
        }
        cudaSetDevice(di);
        const cudaError_t er = cudaDeviceEnablePeerAccess(dj, 0);
        if (er != cudaSuccess && er != cudaErrorPeerAccessAlreadyEnabled) return pfail(DIS_ERR_CUDA, std::string("cudaDeviceEnablePeerAccess: ") + cudaGetErrorString(er));
        cudaGetLastError();
      }
    p->peers_ready = true;
  }
  // 1. every GPU gets both frames and builds its pyramids (in parallel, from the worker threads)
  std::vector<std::function<int(dis_engine*)>> jobs(R);
  for (int r = 0; r < R; ++r) jobs[r] = [=](dis_engine* e) { return dis_tiled_begin(e, a, b, w, h, stride); };
  if (int rc = run_all(p, jobs)) return rc;
  // 2. the level loop, issued from this thread for all GPUs; ordering between GPUs is by events only
  std::vector<dis_engine*> E(R);
  for (int r = 0; r < R; ++r) E[r] = p->w[r]->eng;
  const int nl = dis_num_levels(E[0]);
  const int ps = p->prm.patch_size, half = ps / 2;
  auto fail_here = [&](int rc) { for (int r = 0; r < R; ++r) dis_tiled_end(E[r]); return pfail(rc, dis_last_error()); };
  dis_level_info top_info;
  dis_get_level_info(E[0], 0, &top_info);
  const int top = (top_info.height - h) / 2;                      // divisibility padding above the frame (src/main.cpp:152)
  std::vector<int> band0(R), band1(R);
  for (int level = nl - 1; level >= 0; --level) {
    dis_level_info I, F;
    dis_get_level_info(E[0], level, &I);
    for (int r = 0; r < R; ++r) {
      dis_pool_shard(I.nph, R, r, &band0[r], &band1[r]);
      if (int rc = dis_tiled_search(E[r], level, band0[r], band1[r])) return fail_here(rc);
    }
    if (level > 0) dis_get_level_info(E[0], level - 1, &F);
    for (int r = 0; r < R; ++r) {
      // pixel rows [ya, yb) of this level that GPU r must densify
      int ya = 0, yb = 0;
      if (level == 0) {
        int o0, o1;
        dis_pool_shard(h, R, r, &o0, &o1);
        p->w[r]->lo = o0; p->w[r]->hi = o1;
        ya = top + o0; yb = top + o1;
      } else {
        int f0, f1;
        dis_pool_shard(F.nph, R, r, &f0, &f1);                   // its band at the next-finer level
        if (f1 > f0) { ya = (f0 * F.steps + F.offh) / 2; yb = ((f1 - 1) * F.steps + F.offh) / 2 + 1; }
      }
      if (yb <= ya) continue;
      yb = yb > I.height ? I.height : yb;
      int g0 = ceil_div(ya - half + 1 - I.offh, I.steps), g1 = floor_div(yb - 1 + half - I.offh, I.steps) + 1;   // covering grid rows [g0, g1)
      g0 = g0 < 0 ? 0 : g0; g1 = g1 > I.nph ? I.nph : g1;
      for (int q = 0; q < R; ++q) {                              // halo rows owned by other GPUs
        if (q == r) continue;
        const int lo = g0 > band0[q] ? g0 : band0[q], hi = g1 < band1[q] ? g1 : band1[q];
        if (hi > lo) if (int rc = dis_tiled_fetch_u_rows(E[r], level, E[q], lo, hi)) return fail_here(rc);
      }
      if (int rc = dis_tiled_densify_rows(E[r], level, ya, yb)) return fail_here(rc);
    }
  }
  // 3. every GPU downloads its own band of the result; then wait for all
  for (int r = 0; r < R; ++r)
    if (int rc = dis_tiled_finish_rows(E[r], flow_out, p->w[r]->lo, p->w[r]->hi)) return fail_here(rc);
  int rc = DIS_OK;
  std::string err;
  for (int r = 0; r < R; ++r) {
    const int c = dis_tiled_end(E[r]);
    if (c && !rc) { rc = c; err = "GPU " + std::to_string(p->w[r]->device) + ": " + dis_last_error(); }
    p->w[r]->lo = 0; p->w[r]->hi = 1;                             // took part (dis_pool_get_stats)
  }
  p->wall_ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
  return rc ? pfail(rc, err) : DIS_OK;
}

int dis_pool_last_tiled_device_ms(dis_pool* p, double* max_ms) {
  if (!p || !max_ms) return pfail(DIS_ERR_PARAM, "null argument");
  double m = 0;
  for (Worker* w : p->w) { float ms = 0.f; dis_tiled_last_ms(w->eng, &ms); if (ms > m) m = ms; }
  *max_ms = m;
  return DIS_OK;
}

int dis_pool_get_stats(dis_pool* p, dis_stats* out) {
  if (!p || !out) return pfail(DIS_ERR_PARAM, "null argument");
  memset(out, 0, sizeof *out);
  for (Worker* w : p->w) {
    if (w->hi <= w->lo) continue;
    dis_stats s;
    if (int rc = dis_get_stats(w->eng, &s)) return pfail(rc, dis_last_error());
    out->patches += s.patches; out->patch_iterations += s.patch_iterations; out->sampled_iterations += s.sampled_iterations;
    out->resets += s.resets; out->start_oob += s.start_oob; out->kernel_launches += s.kernel_launches;
    out->sampled_converge += s.sampled_converge; out->sampled_walk += s.sampled_walk;
  }
  return DIS_OK;
}

int dis_pool_last_ms(dis_pool* p, double* wall_ms, double* per_worker_ms) {
  if (!p) return pfail(DIS_ERR_PARAM, "null argument");
  if (wall_ms) *wall_ms = p->wall_ms;
  if (per_worker_ms) for (size_t i = 0; i < p->w.size(); ++i) per_worker_ms[i] = p->w[i]->ms;
  return DIS_OK;
}

int dis_pool_host_alloc(dis_pool* p, int worker, size_t bytes, void** out) {
  if (!p || !out || worker < 0 || worker >= (int)p->w.size()) return pfail(DIS_ERR_PARAM, "bad argument");
  *out = nullptr;
  std::vector<std::function<int(dis_engine*)>> jobs(p->w.size());
  void** dst = out;
  jobs[worker] = [=](dis_engine*) {   // on the worker thread: bound to the GPU's NUMA node, so the pages are local to it
    void* q = nullptr;
    if (cudaHostAlloc(&q, bytes, cudaHostAllocPortable) != cudaSuccess) { cudaGetLastError(); return (int)DIS_ERR_ALLOC; }
    memset(q, 0, bytes);
    *dst = q;
    return (int)DIS_OK;
  };
  const double keep = p->wall_ms;
  const int rc = run_all(p, jobs);
  p->wall_ms = keep;
  return rc;
}

int dis_pool_host_free(dis_pool* p, void* ptr) {
  (void)p;
  if (ptr && cudaFreeHost(ptr) != cudaSuccess) { cudaGetLastError(); return pfail(DIS_ERR_CUDA, "cudaFreeHost failed"); }
  return DIS_OK;
}

}  // extern "C"
