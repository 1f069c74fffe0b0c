// Device half of the C ABI: graph upload, workspace planning and the launch
// sequences of the loss / path-sum / Viterbi / counting entry points.
#include <cuda_runtime.h>

#include <algorithm>
#include <atomic>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <vector>

#include "device_utils.cuh"
#include "kernels.cuh"

using namespace ctcg;

namespace {

constexpr size_t kAlign = 256;
inline size_t align_up(size_t x) { return (x + kAlign - 1) / kAlign * kAlign; }

// What failed, for ctcg_last_error_detail(): the status codes stay coarse (CTCG_ECUDA), the detail says which
// runtime call, where, and the CUDA error string.  One slot per host thread.
thread_local char g_detail[320] = "";
void set_detail(const char* what, const char* file, int line, cudaError_t e) {
  const char* base = std::strrchr(file, '/');
  std::snprintf(g_detail, sizeof(g_detail), "%s:%d: %s -> %s (%s)", base ? base + 1 : file, line, what,
                cudaGetErrorName(e), cudaGetErrorString(e));
}

#define CTCG_CUDA(expr)                                  \
  do {                                                   \
    cudaError_t e__ = (expr);                            \
    if (e__ != cudaSuccess) {                            \
      set_detail(#expr, __FILE__, __LINE__, e__);        \
      (void)cudaGetLastError();                          \
      return CTCG_ECUDA;                                 \
    }                                                    \
  } while (0)

struct Plan {
  int nlp = 0;       // compact row pitch (floats), multiple of 4
  int arow_num = 0;  // alpha row pitch of the numerator
  int arow_den = 0;
  size_t lens = 0, loss_stage = 0, xg = 0, gg = 0, alpha_num = 0, beta_num = 0, cost_num = 0, rowlse = 0, alpha_den = 0, cost_den = 0;
  size_t pspill = 0, ckpt = 0;
  size_t spill2 = 0, ckpt2 = 0, fin = 0;
  BisweepPlan sw{};  // sw.ok: the denominator runs on the factorised bi-char kernels (kernels_bisweep.cu)
  size_t total = 0;
};

// The factorised kernels cover the bi-char decoding graph of ctcg_decoding_graph (shared by the batch).
BisweepPlan structured_plan(int T, int B, int V, const ctcg_graph* den) {
  BisweepPlan none{};
  if (!den || den->structured.kind != 2 || den->num_graphs != 1) return none;
  if (den->structured.S * den->structured.S != V) return none;
  return bisweep_plan(T, B, V, den->structured.S);
}

Plan make_plan(int T, int B, int V, const ctcg_graph* num, const ctcg_graph* den) {
  Plan p;
  size_t off = 0;
  auto take = [&](size_t bytes) {
    const size_t o = off;
    off += align_up(bytes);
    return o;
  };
  const size_t TB = (size_t)std::max(T, 1) * (size_t)std::max(B, 1);
  p.lens = take((size_t)B * 4);
  p.loss_stage = take((size_t)B * 4);  // captured steps write the loss here; it is copied to the caller's pointer
  if (num) {
    p.nlp = pad4(std::max(num->max_labels, 1));
    p.arow_num = pad4(2 * num->max_states + 2);  // split numerator: (integer, fraction) per state
    p.xg = take(TB * p.nlp * 8);  // float columns (generic kernel) or (exponent, mantissa) pairs (split sweeps)
    p.gg = take((TB + 16 * (size_t)std::max(B, 1)) * p.nlp * 4);  // segment-major layout: T rounded up to the segment
    const size_t T1B = (size_t)(std::max(T, 1) + 1) * (size_t)std::max(B, 1);  // the split sweeps spill T + 1 rows
    p.alpha_num = take(T1B * p.arow_num * 4);
    p.beta_num = take(T1B * p.arow_num * 4);
    p.cost_num = take((size_t)B * 8);
  }
  p.rowlse = take(TB * 4);
  if (den) {
    p.sw = structured_plan(T, B, V, den);
    p.cost_den = take((size_t)B * 8);
    if (p.sw.ok) {
      p.pspill = take(p.sw.spill_bytes);
      p.ckpt = take(p.sw.ckpt_bytes);
      p.spill2 = take(p.sw.spill_bytes);
      p.ckpt2 = take(p.sw.ckpt_bytes);
      p.fin = take((size_t)B * sizeof(BiFin));
    } else {
      p.arow_den = pad4(den->max_states + 2);
      p.alpha_den = take((TB + (size_t)std::max(B, 1)) * p.arow_den * 4);  // T + 1 rows per utterance
    }
  }
  p.total = off;
  return p;
}

template <typename T>
void append(std::vector<unsigned char>* blob, const std::vector<T>& v, size_t* off) {
  *off = blob->size();
  const size_t bytes = v.size() * sizeof(T);
  blob->resize(align_up(blob->size() + std::max<size_t>(bytes, 1)));
  if (bytes) std::memcpy(blob->data() + *off, v.data(), bytes);
}

int check_graph(const ctcg_graph* g, int B) {
  if (!g) return CTCG_EINVAL;
  if (g->num_graphs != 1 && g->num_graphs != B) return CTCG_EINVAL;
  int dev = -1;
  if (cudaGetDevice(&dev) != cudaSuccess) return CTCG_ECUDA;
  if (!g->dev_blob || g->device != dev) return CTCG_ENOTUPLOADED;
  return CTCG_OK;
}

// Pinned staging buffers for graph uploads (one small ring per host thread): the host copy
// into a slot is synchronous, the H2D copy out of it is asynchronous; a slot is reused only
// after its previous copy has completed.
cudaError_t stage_and_copy(void* dst, const void* src, size_t bytes, cudaStream_t st) {
  struct Slot {
    void* p = nullptr;
    size_t cap = 0;
    cudaEvent_t done = nullptr;
  };
  // two rings per host thread AND device (a slot's event belongs to the device it was created on): 32 small slots
  // for the per-call lengths, so that the host can run that many calls ahead of the GPU before it waits for a slot,
  // and 4 large ones for graph blobs (their pinned buffers are allocated on first use: not 32 of those)
  constexpr int kSmall = 32, kLarge = 4, kMaxDev = 64;
  constexpr size_t kSmallBytes = size_t{1} << 14;
  thread_local Slot small_ring[kMaxDev][kSmall], large_ring[kMaxDev][kLarge];
  thread_local int small_next[kMaxDev] = {}, large_next[kMaxDev] = {};
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return e;
  if (dev < 0 || dev >= kMaxDev) return cudaErrorInvalidDevice;
  const bool is_small = bytes <= kSmallBytes;
  if (is_small && !small_ring[dev][0].p) {  // the small slots are carved out of one pinned block, made once
    void* block = nullptr;
    if ((e = cudaHostAlloc(&block, kSmall * kSmallBytes, cudaHostAllocDefault)) != cudaSuccess) return e;
    for (int i = 0; i < kSmall; ++i) {
      small_ring[dev][i].p = static_cast<unsigned char*>(block) + (size_t)i * kSmallBytes;
      small_ring[dev][i].cap = kSmallBytes;
      if ((e = cudaEventCreateWithFlags(&small_ring[dev][i].done, cudaEventDisableTiming)) != cudaSuccess) return e;
    }
  }
  Slot& s = is_small ? small_ring[dev][small_next[dev]] : large_ring[dev][large_next[dev]];
  if (is_small) small_next[dev] = (small_next[dev] + 1) % kSmall;
  else large_next[dev] = (large_next[dev] + 1) % kLarge;
  if (!s.done && (e = cudaEventCreateWithFlags(&s.done, cudaEventDisableTiming)) != cudaSuccess) return e;
  if ((e = cudaEventSynchronize(s.done)) != cudaSuccess) return e;  // never-recorded events are complete
  if (s.cap < bytes) {
    if (s.p) cudaFreeHost(s.p);
    s.p = nullptr;
    s.cap = 0;
    const size_t cap = is_small ? kSmallBytes : std::max<size_t>(bytes + bytes / 2, size_t{1} << 20);
    if ((e = cudaHostAlloc(&s.p, cap, cudaHostAllocDefault)) != cudaSuccess) return e;
    s.cap = cap;
  }
  std::memcpy(s.p, src, bytes);
  if ((e = cudaMemcpyAsync(dst, s.p, bytes, cudaMemcpyHostToDevice, st)) != cudaSuccess) return e;
  return cudaEventRecord(s.done, st);
}

// The blob is freed in stream order on the stream it was uploaded on: every kernel that read it
// was launched on that stream (or joined back into it) before the handle is released.
void free_blob(ctcg_graph* g) {
  if (!g->dev_blob) return;
  if (!g->dev_async || cudaFreeAsync(g->dev_blob, static_cast<cudaStream_t>(g->dev_stream)) != cudaSuccess) {
    (void)cudaGetLastError();
    cudaFree(g->dev_blob);
  }
  (void)cudaGetLastError();
}

}  // namespace

extern "C" {

int ctcg_graph_upload(ctcg_graph* g, void* stream) {
  if (!g) return CTCG_EINVAL;
  int dev = -1;
  CTCG_CUDA(cudaGetDevice(&dev));
  if (g->dev_blob && g->device == dev) return CTCG_OK;
  if (g->dev_blob) {
    free_blob(g);
    g->dev_blob = nullptr;
  }
  std::vector<unsigned char> blob;
  size_t o[20];
  append(&blob, g->state_off, &o[0]);
  append(&blob, g->lab_off, &o[1]);
  append(&blob, g->lab_list, &o[2]);
  append(&blob, g->final_w, &o[3]);
  append(&blob, g->in_ptr, &o[4]);
  append(&blob, g->in_src, &o[5]);
  append(&blob, g->in_lab, &o[6]);
  append(&blob, g->in_clab, &o[7]);
  append(&blob, g->in_w, &o[8]);
  append(&blob, g->out_ptr, &o[9]);
  append(&blob, g->out_dst, &o[10]);
  append(&blob, g->out_lab, &o[11]);
  append(&blob, g->out_clab, &o[12]);
  append(&blob, g->out_w, &o[13]);
  append(&blob, g->lab_ptr, &o[14]);
  append(&blob, g->lab_src, &o[15]);
  append(&blob, g->lab_dst, &o[16]);
  append(&blob, g->lab_w, &o[17]);
  append(&blob, g->slab_ptr, &o[18]);
  append(&blob, g->slab_state, &o[19]);
  // transposed arc tables for wide shared graphs (see GraphDev): built here, from the CSR
  const bool ell = g->num_graphs == 1 && g->max_states > 0 && (long long)g->max_arcs >= 16LL * g->max_states &&
                   (long long)(g->max_in_degree + g->max_out_degree) * g->max_states <= (1LL << 22);  // <= 50 MB of tables
  const int ell_np = ell ? (g->max_states + 31) / 32 * 32 : 0;
  size_t oe[6] = {0, 0, 0, 0, 0, 0};
  if (ell) {
    const int n = g->max_states, kin = g->max_in_degree, kout = g->max_out_degree;
    std::vector<int32_t> es((size_t)kin * ell_np, 0), el((size_t)kin * ell_np, 0);
    std::vector<float> ew((size_t)kin * ell_np, -INFINITY);
    for (int s = 0; s < n; ++s)
      for (int e = g->in_ptr[s], k = 0; e < g->in_ptr[s + 1]; ++e, ++k) {
        es[(size_t)k * ell_np + s] = g->in_src[e];
        el[(size_t)k * ell_np + s] = g->in_lab[e];
        ew[(size_t)k * ell_np + s] = g->in_w[e];
      }
    append(&blob, es, &oe[0]);
    append(&blob, el, &oe[1]);
    append(&blob, ew, &oe[2]);
    std::vector<int32_t> od((size_t)kout * ell_np, 0), ol((size_t)kout * ell_np, 0);
    std::vector<float> ow((size_t)kout * ell_np, -INFINITY);
    for (int s = 0; s < n; ++s)
      for (int e = g->out_ptr[s], k = 0; e < g->out_ptr[s + 1]; ++e, ++k) {
        od[(size_t)k * ell_np + s] = g->out_dst[e];
        ol[(size_t)k * ell_np + s] = g->out_lab[e];
        ow[(size_t)k * ell_np + s] = g->out_w[e];
      }
    append(&blob, od, &oe[3]);
    append(&blob, ol, &oe[4]);
    append(&blob, ow, &oe[5]);
  }
  // Stream-ordered allocation and a pinned staging ring: nothing here waits for work already
  // queued on the stream (e.g. the H2D copy of the logits), so graph upload overlaps it.
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  void* d = nullptr;
  bool async_alloc = true;
  if (cudaMallocAsync(&d, blob.size(), st) != cudaSuccess) {
    (void)cudaGetLastError();
    async_alloc = false;
    CTCG_CUDA(cudaMalloc(&d, blob.size()));
  }
  cudaError_t e = stage_and_copy(d, blob.data(), blob.size(), st);
  if (e != cudaSuccess) {
    if (async_alloc) cudaFreeAsync(d, st);
    else cudaFree(d);
    (void)cudaGetLastError();
    return CTCG_ECUDA;
  }
  g->dev_stream = stream;
  g->dev_async = async_alloc;
  unsigned char* p = static_cast<unsigned char*>(d);
  GraphDev& v = g->dev;
  v.num_graphs = g->num_graphs;
  v.max_states = g->max_states;
  v.max_labels = g->max_labels;
  v.max_in_degree = g->max_in_degree;
  v.max_out_degree = g->max_out_degree;
  v.state_off = reinterpret_cast<const int32_t*>(p + o[0]);
  v.lab_off = reinterpret_cast<const int32_t*>(p + o[1]);
  v.lab_list = reinterpret_cast<const int32_t*>(p + o[2]);
  v.final_w = reinterpret_cast<const float*>(p + o[3]);
  v.in_ptr = reinterpret_cast<const int32_t*>(p + o[4]);
  v.in_src = reinterpret_cast<const int32_t*>(p + o[5]);
  v.in_lab = reinterpret_cast<const int32_t*>(p + o[6]);
  v.in_clab = reinterpret_cast<const int32_t*>(p + o[7]);
  v.in_w = reinterpret_cast<const float*>(p + o[8]);
  v.out_ptr = reinterpret_cast<const int32_t*>(p + o[9]);
  v.out_dst = reinterpret_cast<const int32_t*>(p + o[10]);
  v.out_lab = reinterpret_cast<const int32_t*>(p + o[11]);
  v.out_clab = reinterpret_cast<const int32_t*>(p + o[12]);
  v.out_w = reinterpret_cast<const float*>(p + o[13]);
  v.lab_ptr = reinterpret_cast<const int32_t*>(p + o[14]);
  v.lab_src = reinterpret_cast<const int32_t*>(p + o[15]);
  v.lab_dst = reinterpret_cast<const int32_t*>(p + o[16]);
  v.lab_w = reinterpret_cast<const float*>(p + o[17]);
  v.slab_ptr = reinterpret_cast<const int32_t*>(p + o[18]);
  v.slab_state = reinterpret_cast<const int32_t*>(p + o[19]);
  v.ell_np = ell_np;
  v.ell_kin = ell ? g->max_in_degree : 0;
  v.ell_kout = ell ? g->max_out_degree : 0;
  v.ell_in_src = ell ? reinterpret_cast<const int32_t*>(p + oe[0]) : nullptr;
  v.ell_in_lab = ell ? reinterpret_cast<const int32_t*>(p + oe[1]) : nullptr;
  v.ell_in_w = ell ? reinterpret_cast<const float*>(p + oe[2]) : nullptr;
  v.ell_out_dst = ell ? reinterpret_cast<const int32_t*>(p + oe[3]) : nullptr;
  v.ell_out_lab = ell ? reinterpret_cast<const int32_t*>(p + oe[4]) : nullptr;
  v.ell_out_w = ell ? reinterpret_cast<const float*>(p + oe[5]) : nullptr;
  v.max_arcs = g->max_arcs;
  v.unit_weights = g->unit_weights;
  v.in_label_uniform = g->in_label_uniform;
  g->dev_blob = d;
  {
    static std::atomic<unsigned long long> uid{0};
    g->upload_uid = ++uid;
  }
  g->dev_bytes = blob.size();
  g->device = dev;
  return CTCG_OK;
}

void ctcg_graph_free(ctcg_graph* g) {
  if (!g) return;
  if (g->dev_blob) free_blob(g);
  delete g;
}

size_t ctcg_workspace_bytes(int32_t T, int32_t B, int32_t V, const ctcg_graph* num, const ctcg_graph* den) {
  size_t need = make_plan(T, B, V, num, den).total;
  // ctcg_path_sum on a decoding graph runs the fused bi-char kernel, planned in the denominator slot
  if (num && !den) need = std::max(need, make_plan(T, B, V, nullptr, num).total);
  return need;
}

size_t ctcg_workspace_bytes_norm(int32_t T, int32_t B, int32_t V, const ctcg_graph* num, const ctcg_graph* den,
                                 int32_t normalize) {
  size_t need = align_up(ctcg_workspace_bytes(T, B, V, num, den));
  // per-context normalisation in front of a denominator: one normalised copy of the scores
  if (normalize == CTCG_NORM_CONTEXT && den) need += align_up((size_t)T * B * V * 4);
  return need;
}

uint64_t ctcg_launch_count(void) { return g_launches; }

void ctcg_profile_enable(int on) { profile_enable(on != 0); }

int ctcg_profile_collect(char* names, size_t names_cap, float* ms, int max_records) {
  return profile_collect(names, names_cap, ms, max_records);
}

}  // extern "C"

// ---------------------------------------------------------------------------
namespace {

BisweepArgs bichar_args(const Plan& p, unsigned char* ws, const float* x, int T, int B, int V, const int32_t* lens,
                        const ctcg_graph* den, double* cost, float* grad, const float* grad_scale) {
  BisweepArgs s{};
  s.x = x;
  s.total = (long long)T * B * V;
  s.T = T;
  s.B = B;
  s.V = V;
  s.S = den->structured.S;
  s.lens = lens;
  s.loops = den->structured.allow_nonblank_selfloops;
  s.ctx = den->structured.use_contextual_blanks;
  s.spill[0] = reinterpret_cast<float*>(ws + p.pspill);
  s.spill[1] = reinterpret_cast<float*>(ws + p.spill2);
  s.ck[0] = reinterpret_cast<float*>(ws + p.ckpt);
  s.ck[1] = reinterpret_cast<float*>(ws + p.ckpt2);
  s.prowf = p.sw.prowf;
  s.nck = p.sw.nck;
  s.K = p.sw.K;
  s.cost = cost;
  s.fin = reinterpret_cast<BiFin*>(ws + p.fin);
  s.grad = grad;
  s.g_scale = grad_scale;
  s.nmain = p.sw.nmain;
  s.rowf = p.sw.rowf;
  s.estride = p.sw.estride;
  s.epad = p.sw.epad;
  return s;
}

int run_bichar(const Plan& p, unsigned char* ws, const float* x, int T, int B, int V, const int32_t* lens,
               const ctcg_graph* den, double* cost, float* grad, const float* grad_scale, cudaStream_t st) {
  const BisweepArgs s = bichar_args(p, ws, x, T, B, V, lens, den, cost, grad, grad_scale);
  CTCG_CUDA(launch_bisweep(s, p.sw, st));
  CTCG_CUDA(launch_bicombine(s, p.sw, st));
  return CTCG_OK;
}

// The numerator (B small CTAs) and the denominator (B large CTAs) are independent until
// the final subtraction and neither fills 148 SMs at training batch sizes, so the
// denominator runs on a side stream, forked from and joined back into the caller's stream
// with events (legal inside a stream capture).  One side stream per host thread and
// device; CTCG_SERIAL=1 keeps everything on the caller's stream.
constexpr int kGroups = 16;  // most length groups of the denominator (streams of their own)
struct SideStream {
  cudaStream_t stream[kGroups] = {};
  cudaEvent_t fork = nullptr, numdone = nullptr;
  cudaEvent_t join[kGroups] = {};
  bool ok = false;
};

SideStream* side_stream() {
  static const bool serial = [] {
    const char* e = std::getenv("CTCG_SERIAL");
    return e && e[0] == '1';
  }();
  if (serial) return nullptr;
  constexpr int kMaxDev = 64;
  thread_local SideStream table[kMaxDev];
  int dev = -1;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= kMaxDev) return nullptr;
  SideStream& s = table[dev];
  if (!s.ok) {
    bool good = cudaEventCreateWithFlags(&s.fork, cudaEventDisableTiming) == cudaSuccess &&
                cudaEventCreateWithFlags(&s.numdone, cudaEventDisableTiming) == cudaSuccess;
    for (int g = 0; g < kGroups && good; ++g)
      good = cudaStreamCreateWithFlags(&s.stream[g], cudaStreamNonBlocking) == cudaSuccess &&
             cudaEventCreateWithFlags(&s.join[g], cudaEventDisableTiming) == cudaSuccess;
    if (!good) {
      (void)cudaGetLastError();
      return nullptr;
    }
    s.ok = true;
  }
  return &s;
}

// Length groups: the caller's batch is sorted by length (the reference asserts it, fst_utils.py:382,432), so
// contiguous index ranges are length classes.  Each group's sweeps and combine go to a stream of their own:
// the sweeps of the short utterances end at 0.6-0.7 of the longest one's, and their combine starts right
// then, on the SMs those sweeps free, instead of after the whole batch's sweeps (cfg3: 0.413 -> 0.378 ms with
// 4 groups).  Only while every sweep CTA of the batch is resident at once (2 B <= SMs): beyond that the sweeps
// run in waves, the SMs are never idle and smaller grids only fragment the waves (cfg5, B=256: 2.75 -> 3.05 ms).
// CTCG_GROUPS=n forces n groups (1 turns it off).
int sm_count() {
  static const int sms = [] {
    int dev = 0, n = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess) {
      (void)cudaGetLastError();
      n = 148;
    }
    return n;
  }();
  return sms;
}

int group_count(int B, const int32_t* lens_host) {
  static const int forced = [] {
    const char* e = std::getenv("CTCG_GROUPS");
    return e ? std::atoi(e) : 0;
  }();
  const int sms = sm_count();
  static const int per_group = [] {
    const char* e = std::getenv("CTCG_GROUP_MIN");
    return e ? std::max(1, std::atoi(e)) : 8;
  }();
  int g = forced > 0 ? std::min(forced, kGroups) : 8;
  if (forced <= 0 && 2 * B > sms) g = 1;
  while (g > 1 && B < per_group * g) g /= 2;     // at least 8 utterances per group
  for (int b = 1; b < B && g > 1; ++b)
    if (lens_host[b] > lens_host[b - 1]) g = 1;  // not sorted: index ranges are no length classes
  return g;
}

inline bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; }

// While a step is being captured into a CUDA graph (step_graph below) the lengths come from a pinned buffer that
// lives as long as the graph: the staging ring recycles its slots and waits on events, neither of which a captured
// copy node may depend on.
thread_local const int32_t* tls_capture_lens = nullptr;

int upload_lens(const Plan& p, void* ws, const int32_t* lens_host, int B, cudaStream_t st, int32_t** out) {
  int32_t* d = reinterpret_cast<int32_t*>(static_cast<unsigned char*>(ws) + p.lens);
  if (tls_capture_lens) {
    CTCG_CUDA(cudaMemcpyAsync(d, tls_capture_lens, (size_t)B * 4, cudaMemcpyHostToDevice, st));
    *out = d;
    return CTCG_OK;
  }
  CTCG_CUDA(stage_and_copy(d, lens_host, (size_t)B * 4, st));  // pinned staging: no implicit stream sync
  *out = d;
  return CTCG_OK;
}

}  // namespace

extern "C" {

static int loss_fwd_bwd_impl(const float* logits, int32_t T, int32_t B, int32_t V, const int32_t* lens_host,
                             const ctcg_graph* num, const ctcg_graph* den, int32_t normalize, int32_t num_symbols,
                             const float* grad_scale, float* loss, float* grad, void* workspace,
                             size_t workspace_bytes, void* stream) {
  if (!logits || !lens_host || !loss || !workspace || T < 1 || B < 1 || V < 1) return CTCG_EINVAL;
  if (normalize < CTCG_NORM_NONE || normalize > CTCG_NORM_CONTEXT) return CTCG_EINVAL;
  int st_ = check_graph(num, B);
  if (st_ != CTCG_OK) return st_;
  if (den && (st_ = check_graph(den, B)) != CTCG_OK) return st_;
  if (normalize == CTCG_NORM_CONTEXT && (num_symbols < 1 || V % num_symbols != 0)) return CTCG_EINVAL;
  for (int b = 0; b < B; ++b)
    if (lens_host[b] < 0 || lens_host[b] > T) return CTCG_EINVAL;
  if (normalize == CTCG_NORM_CONTEXT && den) {
    // With a denominator the dense posteriors pass through the block softmax too (no reference configuration
    // combines the two, the code allows it: ctc_losses.py:29-43 feeding advanced_decoder.py:488-500): the
    // normalised scores are materialised behind the ordinary workspace, the un-normalised loss runs on them and
    // its gradient is pulled back through the softmax in place.  Workspace: ctcg_workspace_bytes_norm.
    const size_t base = align_up(ctcg_workspace_bytes(T, B, V, num, den));
    const size_t extra = align_up((size_t)T * B * V * 4);
    if (workspace_bytes < base + extra) return CTCG_EWORKSPACE;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    unsigned char* ws = static_cast<unsigned char*>(workspace);
    float* lp = reinterpret_cast<float*>(ws + base);
    // the lengths are needed on the device before the inner call uploads them: same slot, same bytes
    const Plan p0 = make_plan(T, B, V, num, den);
    int32_t* lens0 = nullptr;
    if ((st_ = upload_lens(p0, workspace, lens_host, B, st, &lens0)) != CTCG_OK) return st_;
    CTCG_CUDA(launch_ctx_norm(false, logits, lp, nullptr, T, B, V, num_symbols, lens0, st));
    st_ = ctcg_loss_fwd_bwd(lp, T, B, V, lens_host, num, den, CTCG_NORM_NONE, num_symbols, grad_scale, loss, grad,
                            workspace, base, stream);
    if (st_ != CTCG_OK || !grad) return st_;
    CTCG_CUDA(launch_ctx_norm(true, logits, lp, grad, T, B, V, num_symbols, lens0, st));
    return CTCG_OK;
  }
  if (num->max_label >= V || (den && den->max_label >= V)) return CTCG_EINVAL;
  if (B > 65535) return CTCG_EUNSUPPORTED;  // the combine kernels index the utterance with blockIdx.y
  const Plan p = make_plan(T, B, V, num, den);
  if (workspace_bytes < p.total) return CTCG_EWORKSPACE;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  unsigned char* ws = static_cast<unsigned char*>(workspace);
  int32_t* lens = nullptr;
  if ((st_ = upload_lens(p, workspace, lens_host, B, st, &lens)) != CTCG_OK) return st_;
  float* xg = reinterpret_cast<float*>(ws + p.xg);
  float* gg = reinterpret_cast<float*>(ws + p.gg);
  double* cost_num = reinterpret_cast<double*>(ws + p.cost_num);
  double* cost_den = den ? reinterpret_cast<double*>(ws + p.cost_den) : nullptr;
  float* rowlse = reinterpret_cast<float*>(ws + p.rowlse);

  // A per-frame softmax normaliser cancels between numerator and denominator.
  const bool softmax = (normalize == CTCG_NORM_SOFTMAX) && !den;
  const bool fused_den = den && p.sw.ok && aligned16(logits) && aligned16(grad);
  if (den && p.sw.ok && !fused_den) {  // misaligned tensors: the workspace was planned for the fused kernel
    std::snprintf(g_detail, sizeof(g_detail), "logits / grad must be 16-byte aligned for the bi-char denominator (%p, %p)",
                  (const void*)logits, (const void*)grad);
    return CTCG_EUNSUPPORTED;
  }
  // 1. compact the label columns each utterance touches (the split sweeps take them as (exponent, mantissa)
  // pairs).  Before the fork: next to the denominator sweeps, which hold most of every SM's registers, these
  // wide streaming kernels get one CTA per SM and take three times as long.
  const bool split = [&] {
    PathArgs a{};
    a.g = num->dev;
    a.x_stride_t = pad4(std::max(num->max_labels, 1));
    a.alpha_row = pad4(2 * num->max_states + 2);
    a.beta_ws = reinterpret_cast<float*>(ws);
    return num_split_supported(a);
  }();
  CTCG_CUDA(launch_gather(logits, T, B, V, lens, num->dev, xg, p.nlp, normalize == CTCG_NORM_CONTEXT ? 2 : 0,
                          num_symbols, split, st));
  SideStream* side = fused_den ? side_stream() : nullptr;
  // With the factorised denominator the numerator's posteriors leave in the denominator's combine kernel
  // (one store of gamma_den - gamma_num): sweeps || numerator, then the combine.
  const bool fold = fused_den && grad != nullptr && split;
  BisweepArgs den_args{};
  const int ngroups = (fused_den && side && fold) ? group_count(B, lens_host) : 1;
  auto group_args = [&](int g) {  // utterances [g*B/G, (g+1)*B/G)
    BisweepArgs a = den_args;
    if (ngroups > 1) {
      a.b0 = (int)((long long)g * B / ngroups);
      a.nb = (int)((long long)(g + 1) * B / ngroups) - a.b0;
    }
    return a;
  };
  auto group_stream = [&](int g) { return side ? side->stream[g] : st; };
  if (fused_den) {
    // 1'. factorised bi-char denominator sweeps (kernels_bisweep.cu); concurrent with 1-2
    den_args = bichar_args(p, ws, logits, T, B, V, lens, den, cost_den, grad, grad_scale);
    if (fold) {
      den_args.num_gg = gg;
      den_args.num_nlp = p.nlp;
      den_args.num_graphs = num->dev.num_graphs;
      den_args.num_lab_off = num->dev.lab_off;
      den_args.num_lab_list = num->dev.lab_list;
    }
    if (side) CTCG_CUDA(cudaEventRecord(side->fork, st));
    for (int g = 0; g < ngroups && st_ == CTCG_OK; ++g) {
      if (side) CTCG_CUDA(cudaStreamWaitEvent(side->stream[g], side->fork, 0));
      const cudaError_t le = launch_bisweep(group_args(g), p.sw, group_stream(g));
      if (le != cudaSuccess) {
        set_detail("launch_bisweep", __FILE__, __LINE__, le);
        (void)cudaGetLastError();
        st_ = CTCG_ECUDA;
      }
    }
    if (side && st_ != CTCG_OK)
      for (int g = 0; g < ngroups; ++g) CTCG_CUDA(cudaEventRecord(side->join[g], side->stream[g]));
  }
  // joins the side stream back (also on failure, so that a capture is never left forked)
  auto join = [&]() -> int {
    if (!side) return CTCG_OK;
    for (int g = 0; g < ngroups; ++g) CTCG_CUDA(cudaStreamWaitEvent(st, side->join[g], 0));
    return CTCG_OK;
  };
  if (st_ != CTCG_OK) {
    (void)join();
    return st_;
  }
  st_ = [&]() -> int {
    // 2. numerator forward-backward on the compact columns
    PathArgs a{};
    a.g = num->dev;
    a.x = xg;
    a.x_stride_t = p.nlp;
    a.x_stride_b = (long long)T * p.nlp;
    a.x_scale = kLog2e;
    a.lens = lens;
    a.T = T;
    a.B = B;
    a.alpha_ws = reinterpret_cast<float*>(ws + p.alpha_num);
    a.alpha_row = p.arow_num;
    a.cost = cost_num;
    a.gout = gg;
    a.g_stride_t = p.nlp;
    a.g_stride_b = (long long)T * p.nlp;
    a.g_width = num->max_labels;
    a.g_sign = 1.f;
    a.g_scale = nullptr;
    a.write_grad = grad != nullptr;
    a.beta_ws = reinterpret_cast<float*>(ws + p.beta_num);
    if (split && fold) {  // the denominator's combine reads the posteriors one K-frame segment at a time
      a.g_seg = p.sw.K;
      a.g_stride_b = (long long)p.sw.nck * p.nlp * p.sw.K;
    }
    if (split) {
      CTCG_CUDA(launch_num_sweeps(a, st));
      CTCG_CUDA(launch_num_combine(a, st));
    } else {
      CTCG_CUDA(launch_path_fwd_bwd(a, true, st));  // graphs the split kernels do not take: generic CSR kernel
    }
    return CTCG_OK;
  }();
  if (fused_den) {
    // 3a'. combine: gamma_den - gamma_num, after the numerator's compact rows; shortest group first
    int cs = CTCG_OK;
    if (st_ == CTCG_OK && side && fold && cudaEventRecord(side->numdone, st) != cudaSuccess) {
      (void)cudaGetLastError();
      cs = CTCG_ECUDA;
    }
    for (int g = ngroups - 1; g >= 0; --g) {
      cudaStream_t gs = group_stream(g);
      if (st_ == CTCG_OK && cs == CTCG_OK && side && fold && cudaStreamWaitEvent(gs, side->numdone, 0) != cudaSuccess) {
        (void)cudaGetLastError();
        cs = CTCG_ECUDA;
      }
      if (st_ == CTCG_OK && cs == CTCG_OK && launch_bicombine(group_args(g), p.sw, gs) != cudaSuccess) {
        (void)cudaGetLastError();
        cs = CTCG_ECUDA;
      }
      if (side && cudaEventRecord(side->join[g], gs) != cudaSuccess) {
        (void)cudaGetLastError();
        cs = CTCG_ECUDA;
      }
    }
    const int js = join();
    if (st_ != CTCG_OK) return st_;
    if (cs != CTCG_OK) return cs;
    if (js != CTCG_OK) return js;
  } else if (st_ != CTCG_OK) {
    return st_;
  }

  if (fused_den) {
    // folded: the combine wrote g_scale * (gamma_den - gamma_num); a numerator of the generic kernel is subtracted here
    if (grad && !fold) CTCG_CUDA(launch_scatter_sub(T, B, V, lens, num->dev, gg, p.nlp, grad_scale, grad, st));
  } else if (den) {
    // 3a. denominator path sum over the shared decoding graph, writes +gamma_den rows
    PathArgs d{};
    d.g = den->dev;
    d.x = logits;
    d.x_stride_t = (long long)B * V;
    d.x_stride_b = V;
    d.x_scale = kLog2e;
    d.lens = lens;
    d.T = T;
    d.B = B;
    d.alpha_ws = reinterpret_cast<float*>(ws + p.alpha_den);
    d.alpha_row = p.arow_den;
    d.cost = cost_den;
    d.gout = grad;
    d.g_stride_t = (long long)B * V;
    d.g_stride_b = V;
    d.g_width = V;
    d.g_sign = 1.f;
    d.g_scale = grad_scale;
    d.write_grad = grad != nullptr;
    d.alpha_rows_plus1 = 1;
    CTCG_CUDA(launch_path_fwd_bwd(d, false, st));
    if (grad) CTCG_CUDA(launch_scatter_sub(T, B, V, lens, num->dev, gg, p.nlp, grad_scale, grad, st));
  } else if (grad && normalize == CTCG_NORM_CONTEXT) {
    // 3c. block softmax chain rule: softmax_block(x) * sum_block(gamma_num) - gamma_num
    CTCG_CUDA(launch_rows_ctx(logits, T, B, V, num_symbols, lens, num->dev, gg, p.nlp, grad_scale, grad, st));
  } else if (grad || softmax) {
    // 3b. one streaming pass: softmax(x) - gamma_num (or -gamma_num), row LSE for the loss
    CTCG_CUDA(launch_rows(logits, T, B, V, lens, num->dev, gg, p.nlp, softmax ? 1 : 0, grad_scale, grad,
                          rowlse, st));
  }
  // 4. per-utterance loss, fixed order, fp64 accumulation
  CTCG_CUDA(launch_final(T, B, lens, cost_num, cost_den, softmax ? rowlse : nullptr, loss, st));
  return CTCG_OK;
}

// ---- captured steps ---------------------------------------------------------------------------------------
// One step is ~25 kernel launches on up to 17 streams plus ~50 event records / waits: 0.1 ms of host time, a third
// of the step's GPU time.  A call whose every argument -- pointers, sizes, lengths, graph uploads, stream -- equals an
// earlier call's is replayed from a CUDA graph captured on the third sighting (one cudaGraphLaunch; the logits'
// CONTENTS are of course read anew).  That is the shape of a loop over a fixed set of buffers: evaluation passes,
// gradient accumulation over a cached batch, benchmarks; a trainer whose lengths and transcripts change every
// step never repeats a signature and pays one 16-entry look-up per call.  The caller's stream may be the legacy
// default stream, which cannot be captured: the graph runs on a stream of the library, ordered against the
// caller's by two events.  Off while profiling (ctcg_profile_enable) and with CTCG_STEP_GRAPHS=0.
namespace {

struct StepSig {
  const void *logits, *grad, *ws, *gscale, *stream;  // (not the loss pointer: captured steps stage the loss)
  unsigned long long num_uid, den_uid;
  size_t ws_bytes;
  int T, B, V, normalize, num_symbols;
};

struct StepGraph {
  StepSig sig{};
  std::vector<int32_t> lens;
  int seen = 0;
  bool failed = false;
  cudaGraphExec_t exec = nullptr;
  int32_t* pinned_lens = nullptr;
  unsigned long long launches = 0, stamp = 0;
  void reset() {
    if (exec) cudaGraphExecDestroy(exec);
    if (pinned_lens) cudaFreeHost(pinned_lens);
    exec = nullptr;
    pinned_lens = nullptr;
    seen = 0;
    failed = false;
    launches = 0;
    lens.clear();
  }
};

struct StepGraphCache {
  static constexpr int kEntries = 16;
  StepGraph e[kEntries];
  unsigned long long clock = 0;
  cudaStream_t stream = nullptr;
  cudaEvent_t in = nullptr, out = nullptr;
  bool ok = false, broken = false;
};

std::atomic<unsigned long long> g_step_replays{0}, g_step_captures{0}, g_step_ordinary{0};

StepGraphCache* step_graphs() {
  static const bool on = [] {
    const char* e = std::getenv("CTCG_STEP_GRAPHS");
    return !(e && e[0] == '0');
  }();
  if (!on || profile_is_on()) return nullptr;
  constexpr int kMaxDev = 64;
  thread_local StepGraphCache table[kMaxDev];
  int dev = -1;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= kMaxDev) return nullptr;
  StepGraphCache& c = table[dev];
  if (c.broken) return nullptr;
  if (!c.ok) {
    if (cudaStreamCreateWithFlags(&c.stream, cudaStreamNonBlocking) != cudaSuccess ||
        cudaEventCreateWithFlags(&c.in, cudaEventDisableTiming) != cudaSuccess ||
        cudaEventCreateWithFlags(&c.out, cudaEventDisableTiming) != cudaSuccess) {
      (void)cudaGetLastError();
      c.broken = true;
      return nullptr;
    }
    c.ok = true;
  }
  return &c;
}

}  // namespace

int ctcg_loss_fwd_bwd(const float* logits, int32_t T, int32_t B, int32_t V, const int32_t* lens_host,
                      const ctcg_graph* num, const ctcg_graph* den, int32_t normalize, int32_t num_symbols,
                      const float* grad_scale, float* loss, float* grad, void* workspace,
                      size_t workspace_bytes, void* stream) {
  StepGraphCache* c = nullptr;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  // Only steps short enough for the host to matter (every sweep CTA resident at once, the rule of the length
  // groups): at B = 256 a step is 3 ms, the 0.1 ms of launches is hidden anyway, and the graph's own scheduling of
  // the concurrent branches loses 10 % against the streams' (cfg4 2.95 -> 3.24 ms, cfg5 2.76 -> 3.04 ms).
  if (logits && lens_host && loss && workspace && num && B >= 1 && 2 * B <= sm_count() && !tls_capture_lens &&
      !(normalize == CTCG_NORM_CONTEXT && den)) {
    cudaStreamCaptureStatus cs = cudaStreamCaptureStatusNone;
    if (cudaStreamIsCapturing(st, &cs) == cudaSuccess && cs == cudaStreamCaptureStatusNone) c = step_graphs();
    else (void)cudaGetLastError();
  }
  if (!c)
    return loss_fwd_bwd_impl(logits, T, B, V, lens_host, num, den, normalize, num_symbols, grad_scale, loss, grad,
                             workspace, workspace_bytes, stream);
  StepSig sig{};
  sig.logits = logits, sig.grad = grad, sig.ws = workspace, sig.gscale = grad_scale, sig.stream = stream;
  sig.num_uid = num->upload_uid, sig.den_uid = den ? den->upload_uid : 0;
  sig.ws_bytes = workspace_bytes;
  sig.T = T, sig.B = B, sig.V = V, sig.normalize = normalize, sig.num_symbols = num_symbols;
  auto same = [&](const StepGraph& g) {
    const StepSig& a = g.sig;
    return a.logits == sig.logits && a.grad == sig.grad && a.ws == sig.ws && a.gscale == sig.gscale &&
           a.stream == sig.stream && a.num_uid == sig.num_uid && a.den_uid == sig.den_uid && a.ws_bytes == sig.ws_bytes &&
           a.T == T && a.B == B && a.V == V && a.normalize == normalize && a.num_symbols == num_symbols &&
           (int)g.lens.size() == B && std::memcmp(g.lens.data(), lens_host, (size_t)B * 4) == 0;
  };
  if (workspace_bytes < make_plan(T, B, V, num, den).total)
    return loss_fwd_bwd_impl(logits, T, B, V, lens_host, num, den, normalize, num_symbols, grad_scale, loss, grad,
                             workspace, workspace_bytes, stream);  // (answers CTCG_EWORKSPACE)
  float* loss_stage = reinterpret_cast<float*>(static_cast<unsigned char*>(workspace) + make_plan(T, B, V, num, den).loss_stage);
  StepGraph* hit = nullptr;
  for (StepGraph& g : c->e)
    if (g.seen > 0 && same(g)) {
      hit = &g;
      break;
    }
  ++c->clock;
  auto replay = [&](StepGraph& g) -> int {
    CTCG_CUDA(cudaEventRecord(c->in, st));
    CTCG_CUDA(cudaStreamWaitEvent(c->stream, c->in, 0));
    CTCG_CUDA(cudaGraphLaunch(g.exec, c->stream));
    CTCG_CUDA(cudaEventRecord(c->out, c->stream));
    CTCG_CUDA(cudaStreamWaitEvent(st, c->out, 0));
    // the loss tensor is a fresh small allocation of the caller's on every call: keeping its address out of the
    // signature is what lets a training-style loop (new loss tensor, same big buffers) hit the cache
    CTCG_CUDA(cudaMemcpyAsync(loss, loss_stage, (size_t)B * 4, cudaMemcpyDeviceToDevice, st));
    g_launches += g.launches;
    g.stamp = c->clock;
    ++g_step_replays;
    return CTCG_OK;
  };
  if (hit && hit->exec) return replay(*hit);
  if (hit && !hit->failed && hit->seen >= 2) {
    // third sighting: capture the launch sequence (nothing runs during the capture; it costs milliseconds), then run it
    StepGraph& g = *hit;
    bool ok = cudaHostAlloc(reinterpret_cast<void**>(&g.pinned_lens), (size_t)B * 4, cudaHostAllocDefault) == cudaSuccess;
    int rc = CTCG_ECUDA;
    const unsigned long long l0 = g_launches;
    if (ok) {
      std::memcpy(g.pinned_lens, lens_host, (size_t)B * 4);
      ok = cudaStreamBeginCapture(c->stream, cudaStreamCaptureModeRelaxed) == cudaSuccess;
    }
    if (ok) {
      tls_capture_lens = g.pinned_lens;
      rc = loss_fwd_bwd_impl(logits, T, B, V, lens_host, num, den, normalize, num_symbols, grad_scale, loss_stage, grad,
                             workspace, workspace_bytes, c->stream);
      tls_capture_lens = nullptr;
      cudaGraph_t graph = nullptr;
      const cudaError_t ce = cudaStreamEndCapture(c->stream, &graph);
      ok = rc == CTCG_OK && ce == cudaSuccess && graph != nullptr &&
           cudaGraphInstantiate(&g.exec, graph, 0) == cudaSuccess;
      if (graph) cudaGraphDestroy(graph);
    }
    g.launches = g_launches - l0;
    g_launches = l0;
    ++g_step_captures;
    if (ok) return replay(g);
    (void)cudaGetLastError();
    if (g.exec) cudaGraphExecDestroy(g.exec);
    g.exec = nullptr;
    g.failed = true;  // this signature stays on the ordinary path
    if (rc != CTCG_OK && rc != CTCG_ECUDA) return rc;
  } else if (!hit) {
    // victim: the least recently used entry WITHOUT a captured graph (a stream of never-repeating signatures, e.g. a
    // trainer's, must not push captured graphs out -- destroying one costs as much as a step); a captured one only
    // when every entry holds a graph
    StepGraph* slot = nullptr;
    for (StepGraph& g : c->e)
      if (!g.exec && (!slot || g.stamp < slot->stamp)) slot = &g;
    if (!slot) {
      slot = &c->e[0];
      for (StepGraph& g : c->e)
        if (g.stamp < slot->stamp) slot = &g;
    }
    slot->reset();
    slot->sig = sig;
    slot->lens.assign(lens_host, lens_host + B);
    slot->seen = 1;
    slot->stamp = c->clock;
  } else {
    hit->stamp = c->clock;
    if (!hit->failed) ++hit->seen;
  }
  ++g_step_ordinary;
  return loss_fwd_bwd_impl(logits, T, B, V, lens_host, num, den, normalize, num_symbols, grad_scale, loss, grad,
                           workspace, workspace_bytes, stream);
}

void ctcg_step_graph_info(uint64_t* replays, uint64_t* captures, uint64_t* ordinary) {
  if (replays) *replays = g_step_replays;
  if (captures) *captures = g_step_captures;
  if (ordinary) *ordinary = g_step_ordinary;
}

int ctcg_scale_grad(float* grad, int32_t T, int32_t B, int32_t V, const float* scale, int32_t scale_stride,
                    void* stream) {
  if (!grad || !scale || T < 1 || B < 1 || V < 1 || scale_stride < 0) return CTCG_EINVAL;
  CTCG_CUDA(launch_scale_grad(grad, T, B, V, scale, scale_stride, static_cast<cudaStream_t>(stream)));
  return CTCG_OK;
}

int ctcg_logits_post_width(int32_t V, int32_t flags) {
  if (V < 1) return 0;
  if (flags & CTCG_POST_IMITATE_BIPHONES) return (V <= 46340) ? V * V : 0;
  if (flags & (CTCG_POST_BLOCK_NORMALIZE | CTCG_POST_BLOCK_MARGINALIZE)) {
    int32_t S = 1;
    while ((long long)S * S < V) ++S;
    if (S * S != V) return 0;
    return (flags & CTCG_POST_BLOCK_NORMALIZE) ? V : S;
  }
  return V;
}

int ctcg_logits_post(const float* log_probs, int32_t T, int32_t B, int32_t V, int32_t flags, float* out,
                     void* stream) {
  if (!log_probs || !out || T < 1 || B < 1 || V < 1 || log_probs == out) return CTCG_EINVAL;
  const int32_t W = ctcg_logits_post_width(V, flags);
  if (W < 1) return CTCG_EINVAL;
  const bool transfer = (flags & CTCG_POST_TRANSFER_HASH) != 0;
  if (transfer && V < 4) return CTCG_EINVAL;
  int mode = 0, S = 0, tile_norm = 0;
  if (flags & CTCG_POST_IMITATE_BIPHONES) {
    mode = 1;
    tile_norm = (flags & CTCG_POST_BLOCK_NORMALIZE) ? 0 : 1;
  } else if (flags & CTCG_POST_BLOCK_NORMALIZE) {
    mode = 2;
    S = ctcg_logits_post_width(V, CTCG_POST_BLOCK_MARGINALIZE);
  } else if (flags & CTCG_POST_BLOCK_MARGINALIZE) {
    mode = 3;
    S = W;
  }
  CTCG_CUDA(launch_post(log_probs, (long long)T * B, V, S, mode, transfer ? 1 : 0, tile_norm, out,
                        static_cast<cudaStream_t>(stream)));
  return CTCG_OK;
}

int ctcg_path_sum(const float* log_probs, int32_t T, int32_t B, int32_t V, const int32_t* lens_host,
                  const ctcg_graph* graph, float* cost, float* grad, void* workspace,
                  size_t workspace_bytes, void* stream) {
  if (!log_probs || !lens_host || !cost || !workspace || T < 1 || B < 1 || V < 1) return CTCG_EINVAL;
  int st_ = check_graph(graph, B);
  if (st_ != CTCG_OK) return st_;
  for (int b = 0; b < B; ++b)
    if (lens_host[b] < 0 || lens_host[b] > T) return CTCG_EINVAL;
  if (graph->max_label >= V) return CTCG_EINVAL;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  unsigned char* ws = static_cast<unsigned char*>(workspace);
  int32_t* lens = nullptr;
  const Plan ps = make_plan(T, B, V, nullptr, graph);
  if (ps.sw.ok && aligned16(log_probs) && aligned16(grad)) {
    if (workspace_bytes < ps.total) return CTCG_EWORKSPACE;
    if ((st_ = upload_lens(ps, workspace, lens_host, B, st, &lens)) != CTCG_OK) return st_;
    double* c64 = reinterpret_cast<double*>(ws + ps.cost_den);
    if ((st_ = run_bichar(ps, ws, log_probs, T, B, V, lens, graph, c64, grad, nullptr, st)) != CTCG_OK) return st_;
    CTCG_CUDA(launch_cost_to_float(B, c64, cost, st));
    return CTCG_OK;
  }
  const Plan p = make_plan(T, B, V, graph, nullptr);
  if (workspace_bytes < p.total) return CTCG_EWORKSPACE;
  if ((st_ = upload_lens(p, workspace, lens_host, B, st, &lens)) != CTCG_OK) return st_;
  double* cost64 = reinterpret_cast<double*>(ws + p.cost_num);
  PathArgs d{};
  d.g = graph->dev;
  d.x = log_probs;
  d.x_stride_t = (long long)B * V;
  d.x_stride_b = V;
  d.x_scale = kLog2e;
  d.lens = lens;
  d.T = T;
  d.B = B;
  d.alpha_ws = reinterpret_cast<float*>(ws + p.alpha_num);
  d.alpha_row = p.arow_num;
  d.cost = cost64;
  d.gout = grad;
  d.g_stride_t = (long long)B * V;
  d.g_stride_b = V;
  d.g_width = V;
  d.g_sign = 1.f;
  d.g_scale = nullptr;
  d.write_grad = grad != nullptr;
  d.alpha_rows_plus1 = 1;  // alpha_num is planned with T + 1 rows
  CTCG_CUDA(launch_path_fwd_bwd(d, false, st));
  CTCG_CUDA(launch_cost_to_float(B, cost64, cost, st));
  return CTCG_OK;
}

int ctcg_viterbi(const float* log_probs, int32_t T, int32_t B, int32_t V, const int32_t* lens_host,
                 const ctcg_graph* graph, float* score, int32_t* frame_labels, void* workspace,
                 size_t workspace_bytes, void* stream) {
  if (!log_probs || !lens_host || !score || !frame_labels || !workspace || T < 1 || B < 1 || V < 1)
    return CTCG_EINVAL;
  int st_ = check_graph(graph, B);
  if (st_ != CTCG_OK) return st_;
  for (int b = 0; b < B; ++b)
    if (lens_host[b] < 0 || lens_host[b] > T) return CTCG_EINVAL;
  if (graph->max_label >= V) return CTCG_EINVAL;
  const Plan p = make_plan(T, B, V, graph, nullptr);
  if (workspace_bytes < p.total) return CTCG_EWORKSPACE;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  unsigned char* ws = static_cast<unsigned char*>(workspace);
  int32_t* lens = nullptr;
  if ((st_ = upload_lens(p, workspace, lens_host, B, st, &lens)) != CTCG_OK) return st_;
  // the bi-char decoding graph takes the factorised kernel (its back-pointers, 2 bytes per state and frame, fit
  // the generic kernel's 4-byte slots planned above); CTCG_VITERBI_GENERIC=1 keeps the generic one (cross-checks)
  static const bool force_generic = [] {
    const char* e = std::getenv("CTCG_VITERBI_GENERIC");
    return e && e[0] == '1';
  }();
  if (!force_generic && graph->structured.kind == 2 && graph->num_graphs == 1 &&
      graph->structured.S * graph->structured.S == V && graph->structured.S <= 64) {
    CTCG_CUDA(launch_biviterbi(log_probs, T, B, graph->structured.S, lens, graph->structured.allow_nonblank_selfloops,
                               graph->structured.use_contextual_blanks,
                               reinterpret_cast<unsigned short*>(ws + p.alpha_num), score, frame_labels, st));
    return CTCG_OK;
  }
  CTCG_CUDA(launch_viterbi(log_probs, T, B, V, lens, graph->dev, reinterpret_cast<int32_t*>(ws + p.alpha_num),
                           p.arow_num, score, frame_labels, st));
  return CTCG_OK;
}

int ctcg_path_count(int32_t B, const int32_t* lens_host, const ctcg_graph* graph, uint64_t* counts,
                    void* workspace, size_t workspace_bytes, void* stream) {
  if (!lens_host || !counts || !workspace || B < 1) return CTCG_EINVAL;
  int st_ = check_graph(graph, B);
  if (st_ != CTCG_OK) return st_;
  for (int b = 0; b < B; ++b)
    if (lens_host[b] < 0) return CTCG_EINVAL;
  const Plan p = make_plan(1, B, 1, nullptr, nullptr);
  if (workspace_bytes < p.total) return CTCG_EWORKSPACE;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  int32_t* lens = nullptr;
  if ((st_ = upload_lens(p, workspace, lens_host, B, st, &lens)) != CTCG_OK) return st_;
  CTCG_CUDA(launch_path_count(B, lens, graph->dev, counts, st));
  return CTCG_OK;
}

}  // extern "C"

// ---------------------------------------------------------------------------
// Host-buffer entry point with cached device scratch.
namespace {

struct HostCache {
  std::mutex mu;
  void* logits = nullptr;
  void* grad = nullptr;
  void* loss = nullptr;
  void* ws = nullptr;
  size_t logits_bytes = 0, grad_bytes = 0, loss_bytes = 0, ws_bytes = 0;
  int device = -1;
};
HostCache g_cache;

int ensure(void** p, size_t* have, size_t need) {
  if (*have >= need) return CTCG_OK;
  if (*p) cudaFree(*p);
  *p = nullptr;
  *have = 0;
  CTCG_CUDA(cudaMalloc(p, need));
  *have = need;
  return CTCG_OK;
}

}  // namespace

// Frames t >= lens[b] carry nothing (no kernel reads them, the gradient rows are zeros): a length-sorted batch
// is copied as one strided block per run of frames with the same number of live utterances -- frame t holds
// its live utterances as a prefix of the (B, V) row -- so the bus moves sum(lens) * V floats instead of
// T * B * V (20 % less at lens in [0.6 T, T]).  The padding of the device buffer is zero-filled on the device.
extern "C" const char* ctcg_last_error_detail(void) { return g_detail; }

extern "C" int ctcg_upload_logits(const float* logits_host, int32_t T, int32_t B, int32_t V,
                                  const int32_t* lens_host, float* logits_dev, void* stream) {
  if (!logits_host || !lens_host || !logits_dev || T < 1 || B < 1 || V < 1) return CTCG_EINVAL;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const size_t row = (size_t)B * V * 4;
  bool sorted = true;
  long long live = 0;
  for (int b = 0; b < B; ++b) {
    if (lens_host[b] < 0 || lens_host[b] > T) return CTCG_EINVAL;
    if (b > 0 && lens_host[b] > lens_host[b - 1]) sorted = false;
    live += lens_host[b];
  }
  if (!sorted || live == (long long)T * B) {  // no prefix structure / nothing to skip: one flat copy
    CTCG_CUDA(cudaMemcpyAsync(logits_dev, logits_host, row * T, cudaMemcpyHostToDevice, st));
    return CTCG_OK;
  }
  CTCG_CUDA(cudaMemsetAsync(logits_dev, 0, row * T, st));
  // runs of frames [t0, t1) during which exactly n utterances are alive: n = B for t < lens[B-1], then one
  // utterance less after every distinct length
  int t0 = 0;
  for (int n = B; n >= 1 && t0 < T; --n) {
    const int t1 = std::min<int>(lens_host[n - 1], T);  // utterance n-1 (the shortest of the first n) ends here
    if (t1 <= t0) continue;
    CTCG_CUDA(cudaMemcpy2DAsync(logits_dev + (size_t)t0 * B * V, row, logits_host + (size_t)t0 * B * V, row,
                                (size_t)n * V * 4, (size_t)(t1 - t0), cudaMemcpyHostToDevice, st));
    t0 = t1;
  }
  return CTCG_OK;
}

extern "C" int ctcg_loss_host(const float* logits_host, int32_t T, int32_t B, int32_t V,
                              const int32_t* lens_host, ctcg_graph* num, ctcg_graph* den, int32_t normalize,
                              int32_t num_symbols, float* loss_host, float* grad_host) {
  if (!logits_host || !lens_host || !num || !loss_host || T < 1 || B < 1 || V < 1) return CTCG_EINVAL;
  std::lock_guard<std::mutex> lock(g_cache.mu);
  int dev = -1;
  CTCG_CUDA(cudaGetDevice(&dev));
  if (g_cache.device != dev) {
    // scratch belongs to one device; drop it when the caller switches
    if (g_cache.logits) cudaFree(g_cache.logits);
    if (g_cache.grad) cudaFree(g_cache.grad);
    if (g_cache.loss) cudaFree(g_cache.loss);
    if (g_cache.ws) cudaFree(g_cache.ws);
    g_cache.logits = g_cache.grad = g_cache.loss = g_cache.ws = nullptr;
    g_cache.logits_bytes = g_cache.grad_bytes = g_cache.loss_bytes = g_cache.ws_bytes = 0;
    g_cache.device = dev;
  }
  int st_;
  if ((st_ = ctcg_graph_upload(num, nullptr)) != CTCG_OK) return st_;
  if (den && (st_ = ctcg_graph_upload(den, nullptr)) != CTCG_OK) return st_;
  const size_t n = (size_t)T * B * V * 4;
  if ((st_ = ensure(&g_cache.logits, &g_cache.logits_bytes, n)) != CTCG_OK) return st_;
  if (grad_host && (st_ = ensure(&g_cache.grad, &g_cache.grad_bytes, n)) != CTCG_OK) return st_;
  if ((st_ = ensure(&g_cache.loss, &g_cache.loss_bytes, (size_t)B * 4)) != CTCG_OK) return st_;
  const size_t wsb = ctcg_workspace_bytes_norm(T, B, V, num, den, normalize);
  if ((st_ = ensure(&g_cache.ws, &g_cache.ws_bytes, wsb)) != CTCG_OK) return st_;
  cudaStream_t st = nullptr;
  if ((st_ = ctcg_upload_logits(logits_host, T, B, V, lens_host, static_cast<float*>(g_cache.logits), st)) != CTCG_OK)
    return st_;
  st_ = ctcg_loss_fwd_bwd(static_cast<const float*>(g_cache.logits), T, B, V, lens_host, num, den, normalize,
                          num_symbols, nullptr, static_cast<float*>(g_cache.loss),
                          grad_host ? static_cast<float*>(g_cache.grad) : nullptr, g_cache.ws, g_cache.ws_bytes,
                          st);
  if (st_ != CTCG_OK) return st_;
  CTCG_CUDA(cudaMemcpyAsync(loss_host, g_cache.loss, (size_t)B * 4, cudaMemcpyDeviceToHost, st));
  if (grad_host) CTCG_CUDA(cudaMemcpyAsync(grad_host, g_cache.grad, n, cudaMemcpyDeviceToHost, st));
  CTCG_CUDA(cudaStreamSynchronize(st));
  return CTCG_OK;
}
