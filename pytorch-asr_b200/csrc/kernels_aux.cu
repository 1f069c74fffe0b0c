// Tropical (Viterbi) and counting semirings on the same CSR skeleton.
//
// Viterbi replaces path_reduction(red_kind='viterbi') + the gradient arg-min
// read-out of FSTDecoder.decode (att_speech/fst_utils.py:366-397,
// modules/decoders/advanced_decoder.py:546-554) and the dense
// raw_generic_viterbi back-trace (modules/ctc_losses.py:412-457): instead of
// autograd through T cloned alphas it keeps one back-pointer per (frame, state)
// and walks them once.  Ties take the first arc in (src, label, weight) order,
// which is what torch.max over the reference's sorted arc lists returns.
//
// Path counting is the integer counterpart of path_reduction on all-zero
// log-probs: exact modulo 2^64.
#include <mutex>

#include "device_utils.cuh"
#include "kernels.cuh"

namespace ctcg {

namespace {

constexpr int NT = 256;
constexpr int NWARP = NT / 32;

// WIDE (shared graphs with tens of arcs per state, as in kernels_path.cu): one state per thread over the transposed
// arc tables of GraphDev, 1024 threads.  (One state per WARP over the CSR was tried first: 184 warp instructions per
// state, issue-bound, no faster than the 256-thread CSR loop -- profiles/micro/generic_viterbi.py.)
template <bool WIDE, int NT>
__global__ void __launch_bounds__(NT) viterbi_kernel(const float* __restrict__ x, int T, int B, int V,
                                                     const int32_t* __restrict__ lens, GraphDev g,
                                                     int32_t* __restrict__ back_ws, int row,
                                                     float* __restrict__ score, int32_t* __restrict__ frame_labels) {
  constexpr int NWARP = NT / 32;
  extern __shared__ __align__(16) float vbuf[];  // 2*np + NWARP*2 + NWARP ints
  const int np = pad4(g.max_states);
  float* wred = vbuf + 2 * np;
  float* fval = wred + 2 * NWARP;
  int* fidx = reinterpret_cast<int*>(fval + NWARP);
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int b = blockIdx.x;
  const int gi = g.num_graphs == 1 ? 0 : b;
  const int so = g.state_off[gi], n = g.state_off[gi + 1] - so, ro = so + gi;
  const int len = min(lens[b], T);
  const int32_t* in_ptr = g.in_ptr + ro;
  const float* final_w = g.final_w + so;
  int32_t* back = back_ws + (long long)b * T * row;
  float* v_cur = vbuf;
  float* v_nxt = vbuf + np;
  for (int s = tid; s < n; s += NT) v_cur[s] = s == 0 ? 0.f : neg_inf();
  if (tid < 2 * NWARP) wred[tid] = 0.f;
  for (int t = len + tid; t < T; t += NT) frame_labels[(long long)t * B + b] = 0;
  double c = 0.0;
  for (int t = 0; t < len; ++t) {
    __syncthreads();
    float mx = wred[(t & 1) * NWARP];
#pragma unroll
    for (int i = 1; i < NWARP; ++i) mx = fmaxf(mx, wred[(t & 1) * NWARP + i]);
    if (!(mx > neg_inf())) mx = 0.f;
    c += (double)mx;
    const float* xrow = x + ((long long)t * B + b) * V;
    float lm = neg_inf();
    if (WIDE) {
      // transposed tables: slot k of state s at [k * np + s] -- consecutive lanes, consecutive addresses; the
      // back-pointer is the slot (lowest slot among equal scores = lowest arc index of the CSR loop)
      const int enp = g.ell_np, K = g.ell_kin;
      for (int s = tid; s < n; s += NT) {
        float best = neg_inf();
        int arg = -1;
#pragma unroll 4
        for (int k = 0; k < K; ++k) {
          const int i = k * enp + s;
          const float v = g.ell_in_w[i] + (v_cur[g.ell_in_src[i]] - mx) + xrow[g.ell_in_lab[i]];
          if (v > best) {
            best = v;
            arg = k;
          }
        }
        v_nxt[s] = best;
        back[(long long)t * row + s] = arg;
        lm = fmaxf(lm, best);
      }
    } else {
      for (int s = tid; s < n; s += NT) {
        float best = neg_inf();
        int arg = -1;
        for (int e = in_ptr[s]; e < in_ptr[s + 1]; ++e) {
          const float v = g.in_w[e] + (v_cur[g.in_src[e]] - mx) + xrow[g.in_lab[e]];
          if (v > best) {
            best = v;
            arg = e;
          }
        }
        v_nxt[s] = best;
        back[(long long)t * row + s] = arg;
        lm = fmaxf(lm, best);
      }
    }
    lm = warp_max(lm);
    if (lane == 0) wred[((t + 1) & 1) * NWARP + warp] = lm;
    float* tmp = v_cur;
    v_cur = v_nxt;
    v_nxt = tmp;
  }
  __syncthreads();
  float mx = 0.f;
  if (len > 0) {
    mx = wred[(len & 1) * NWARP];
#pragma unroll
    for (int i = 1; i < NWARP; ++i) mx = fmaxf(mx, wred[(len & 1) * NWARP + i]);
    if (!(mx > neg_inf())) mx = 0.f;
  }
  c += (double)mx;
  // best final state, lowest index on ties
  float bv = neg_inf();
  int bi = 0x7fffffff;
  for (int s = tid; s < n; s += NT) {
    const float v = (v_cur[s] - mx) + final_w[s];
    if (v > bv) {
      bv = v;
      bi = s;
    }
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const float ov = __shfl_xor_sync(0xffffffffu, bv, o);
    const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
    if (ov > bv || (ov == bv && oi < bi)) {
      bv = ov;
      bi = oi;
    }
  }
  if (lane == 0) {
    fval[warp] = bv;
    fidx[warp] = bi;
  }
  __syncthreads();
  if (tid == 0) {
    for (int i = 1; i < NWARP; ++i)
      if (fval[i] > bv || (fval[i] == bv && fidx[i] < bi)) {
        bv = fval[i];
        bi = fidx[i];
      }
    if (!(bv > neg_inf())) {
      score[b] = -1e20f;
      for (int t = 0; t < len; ++t) frame_labels[(long long)t * B + b] = 0;
    } else {
      score[b] = (float)(c + (double)bv);
      int s = bi;
      for (int t = len - 1; t >= 0; --t) {
        const int e = back[(long long)t * row + s];
        if (WIDE) {  // e is a slot of state s
          frame_labels[(long long)t * B + b] = g.ell_in_lab[e * g.ell_np + s];
          s = g.ell_in_src[e * g.ell_np + s];
        } else {
          frame_labels[(long long)t * B + b] = g.in_lab[e];
          s = g.in_src[e];
        }
      }
    }
  }
}

// ---------------------------------------------------------------------------
// Viterbi on the bi-char decoding graph in its factorised form (the tropical counterpart of
// kernels_bisweep.cu; fst_utils.py:366-397 on the graph of :765-817).  State (c, l) = (context, last
// letter), id c*S + l.  Every arc into (c, l) carries class cls(c, l) and comes from a state whose next
// context is c -- (c1, c) for every c1 when c != 0, and the blank state (c, 0) -- or is the explicit
// self-loop (allow_nonblank_selfloops, l != 0, l != c):
//     Pmax[c]  = max( max_{c1} v[c1][c]  (c != 0),  v[c][0] )
//     v'[c][l] = x[cls(c, l)] + max( Pmax[c], loop(c, l) ? v[c][l] : -inf )
// O(V) per frame instead of the 51 V arcs of the generic kernel.  Ties take the source with the smallest
// state id, the order torch.max sees in the reference's arc lists (sorted by source state, :285).
// One CTA per utterance; back[t][s] = source state (16 bits); the row of frame t+1 is requested a frame ahead.
constexpr int BV_NT = 512;
constexpr int BV_PER = 8;  // states per thread: V <= 4096

__global__ void __launch_bounds__(BV_NT) biviterbi_kernel(const float* __restrict__ x, int T, int B, int S,
                                                          const int32_t* __restrict__ lens, int loops, int ctx,
                                                          unsigned short* __restrict__ back_ws,
                                                          float* __restrict__ score, int32_t* __restrict__ frame_labels) {
  extern __shared__ __align__(16) float bsm[];  // v[2][V], pmax[S], parg[S] (ints), red
  const int V = S * S;
  float* v0 = bsm;
  float* v1 = bsm + V;
  float* pmax = bsm + 2 * V;
  int* parg = reinterpret_cast<int*>(pmax + 64);
  float* rval = reinterpret_cast<float*>(parg + 64);
  int* ridx = reinterpret_cast<int*>(rval + BV_NT / 32);
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int b = blockIdx.x;
  const int len = min(max(lens[b], 0), T);
  unsigned short* back = back_ws + (long long)b * T * V;
  for (int s = tid; s < V; s += BV_NT) v0[s] = (s == 0) ? 0.f : neg_inf();
  for (int t = len + tid; t < T; t += BV_NT) frame_labels[(long long)t * B + b] = 0;
  float xr[BV_PER];
  auto load_row = [&](int t) {
    const float* row = x + ((long long)t * B + b) * V;
#pragma unroll
    for (int k = 0; k < BV_PER; ++k) {
      const int s = tid + k * BV_NT;
      if (s < V) {
        const int c = s / S, l = s - c * S;
        const int cls = (l != 0 || ctx) ? s : 0;  // all (c, 0) share class 0 unless blanks are contextual
        xr[k] = __ldg(row + cls);
        (void)c;
      }
    }
  };
  if (len > 0) load_row(0);
  float* vc = v0;
  float* vn = v1;
  __syncthreads();
  for (int t = 0; t < len; ++t) {
    // phase 1: Pmax[c] and its source, smallest state id among the maxima
    for (int c = warp; c < S; c += BV_NT / 32) {
      float best = neg_inf();
      int arg = 0x7fffffff;
      if (c != 0)
        for (int c1 = lane; c1 < S; c1 += 32) {
          const float val = vc[c1 * S + c];
          if (val > best) {  // ids ascend with c1: strict > keeps the smallest
            best = val;
            arg = c1 * S + c;
          }
        }
      if (lane == 0) {  // the blank state (c, 0)
        const float val = vc[c * S];
        const int id = c * S;
        if (val > best || (val == best && id < arg)) {
          best = val;
          arg = id;
        }
      }
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) {
        const float ov = __shfl_xor_sync(0xffffffffu, best, o);
        const int oi = __shfl_xor_sync(0xffffffffu, arg, o);
        if (ov > best || (ov == best && oi < arg)) {
          best = ov;
          arg = oi;
        }
      }
      if (lane == 0) {
        pmax[c] = best;
        parg[c] = arg;
      }
    }
    __syncthreads();
    // phase 2: every state from its context's best source or its own self-loop
    float xcur[BV_PER];
#pragma unroll
    for (int k = 0; k < BV_PER; ++k) xcur[k] = xr[k];
    if (t + 1 < len) load_row(t + 1);
    unsigned short* bk = back + (long long)t * V;
#pragma unroll
    for (int k = 0; k < BV_PER; ++k) {
      const int s = tid + k * BV_NT;
      if (s < V) {
        const int c = s / S, l = s - c * S;
        float best = pmax[c];
        int arg = parg[c];
        if (loops && l != 0 && l != c) {
          const float sv = vc[s];
          if (sv > best || (sv == best && s < arg)) {
            best = sv;
            arg = s;
          }
        }
        const bool alive = best > neg_inf();
        vn[s] = alive ? best + xcur[k] : neg_inf();
        bk[s] = (unsigned short)(alive ? arg : 0);
      }
    }
    __syncthreads();
    float* tmp = vc;
    vc = vn;
    vn = tmp;
  }
  // best final state (all states are final), lowest id on ties
  float bv = neg_inf();
  int bi = 0x7fffffff;
  for (int s = tid; s < V; s += BV_NT) {
    const float val = vc[s];
    if (val > bv) {
      bv = val;
      bi = s;
    }
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const float ov = __shfl_xor_sync(0xffffffffu, bv, o);
    const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
    if (ov > bv || (ov == bv && oi < bi)) {
      bv = ov;
      bi = oi;
    }
  }
  if (lane == 0) {
    rval[warp] = bv;
    ridx[warp] = bi;
  }
  __syncthreads();
  if (tid == 0) {
    for (int i = 1; i < BV_NT / 32; ++i)
      if (rval[i] > bv || (rval[i] == bv && ridx[i] < bi)) {
        bv = rval[i];
        bi = ridx[i];
      }
    if (!(bv > neg_inf())) {
      score[b] = -1e20f;
      for (int t = 0; t < len; ++t) frame_labels[(long long)t * B + b] = 0;
    } else {
      score[b] = bv;
      int s = bi;
      for (int t = len - 1; t >= 0; --t) {
        const int l = s % S;
        frame_labels[(long long)t * B + b] = (l != 0 || ctx) ? s : 0;  // the class on every arc into s
        s = back[(long long)t * V + s];
      }
    }
  }
}

__global__ void __launch_bounds__(NT) count_kernel(int B, const int32_t* __restrict__ lens, GraphDev g,
                                                   unsigned long long* __restrict__ counts) {
  extern __shared__ __align__(16) unsigned long long cbuf[];  // 2*n + NT
  const int tid = threadIdx.x;
  const int b = blockIdx.x;
  const int gi = g.num_graphs == 1 ? 0 : b;
  const int so = g.state_off[gi], n = g.state_off[gi + 1] - so, ro = so + gi;
  const int32_t* in_ptr = g.in_ptr + ro;
  unsigned long long* cur = cbuf;
  unsigned long long* nxt = cbuf + g.max_states;
  unsigned long long* red = cbuf + 2 * g.max_states;
  for (int s = tid; s < n; s += NT) cur[s] = s == 0 ? 1ull : 0ull;
  const int len = lens[b];
  for (int t = 0; t < len; ++t) {
    __syncthreads();
    for (int s = tid; s < n; s += NT) {
      unsigned long long acc = 0;
      for (int e = in_ptr[s]; e < in_ptr[s + 1]; ++e) acc += cur[g.in_src[e]];
      nxt[s] = acc;
    }
    unsigned long long* tmp = cur;
    cur = nxt;
    nxt = tmp;
  }
  __syncthreads();
  unsigned long long acc = 0;
  for (int s = tid; s < n; s += NT)
    if (g.final_w[so + s] > kDeadWeight) acc += cur[s];
  red[tid] = acc;
  __syncthreads();
  if (tid == 0) {
    unsigned long long tot = 0;
    for (int i = 0; i < NT; ++i) tot += red[i];
    counts[b] = tot;
  }
}

}  // namespace

cudaError_t launch_viterbi(const float* x, int T, int B, int V, const int32_t* lens, const GraphDev& g,
                           int32_t* back_ws, int row, float* score, int32_t* frame_labels, cudaStream_t st) {
  static const bool narrow_only = [] {
    const char* e = getenv("CTCG_PATH_NARROW");
    return e && e[0] == '1';
  }();
  const bool wide = !narrow_only && g.ell_in_src != nullptr;  // transposed arc tables (wide shared graphs)
  constexpr int NTW = 1024;
  const size_t smem = (size_t)(2 * pad4(g.max_states) + 4 * ((wide ? NTW : NT) / 32)) * 4;
  if (smem > 200 * 1024) return cudaErrorInvalidConfiguration;
  cudaError_t e = wide ? cudaFuncSetAttribute(viterbi_kernel<true, NTW>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)
                       : cudaFuncSetAttribute(viterbi_kernel<false, NT>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  if (e != cudaSuccess) return e;
  ProfScope prof("viterbi", st);
  if (wide) viterbi_kernel<true, NTW><<<B, NTW, smem, st>>>(x, T, B, V, lens, g, back_ws, row, score, frame_labels);
  else viterbi_kernel<false, NT><<<B, NT, smem, st>>>(x, T, B, V, lens, g, back_ws, row, score, frame_labels);
  ++g_launches;
  return cudaGetLastError();
}

cudaError_t launch_biviterbi(const float* x, int T, int B, int S, const int32_t* lens, int loops, int ctx,
                             unsigned short* back_ws, float* score, int32_t* frame_labels, cudaStream_t st) {
  if (S < 2 || S > 64) return cudaErrorInvalidConfiguration;
  const size_t smem = (size_t)(2 * S * S + 128 + 2 * (BV_NT / 32)) * 4;
  cudaError_t e = cudaFuncSetAttribute(biviterbi_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  if (e != cudaSuccess) return e;
  ProfScope prof("biviterbi", st);
  biviterbi_kernel<<<B, BV_NT, smem, st>>>(x, T, B, S, lens, loops, ctx, back_ws, score, frame_labels);
  ++g_launches;
  return cudaGetLastError();
}

cudaError_t launch_path_count(int B, const int32_t* lens, const GraphDev& g, uint64_t* counts, cudaStream_t st) {
  const size_t smem = (size_t)(2 * g.max_states + NT) * 8;
  if (smem > 200 * 1024) return cudaErrorInvalidConfiguration;
  cudaError_t e = cudaFuncSetAttribute(count_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  if (e != cudaSuccess) return e;
  ProfScope prof("path_count", st);
  count_kernel<<<B, NT, smem, st>>>(B, lens, g, reinterpret_cast<unsigned long long*>(counts));
  ++g_launches;
  return cudaGetLastError();
}

}  // namespace ctcg

// ---------------------------------------------------------------------------
// ctcg_viterbi_static: Viterbi with NO allocation anywhere on the call path (no counterpart in the
// reference; added for callers that cannot allocate: CUDA graph
// capture, tight serving loops).  The scratch of viterbi_kernel -- the utterance lengths and one
// back-pointer per (frame, state) -- lives in a file-scope device array of CTCG_STATIC_SCRATCH_BYTES
// that every GPU which loads the library holds until exit.  In host code the array's name is only a
// stand-in: its device address is looked up with cudaGetSymbolAddress, once per device with that
// device current, and that pointer is what the kernels get.
//
// Every call shares the one array, which keeps the previous call's contents.  Nothing in it has to
// start at zero: the lengths are rewritten by clamp_lens_kernel and viterbi_kernel writes every
// back-pointer (t < len, s < n) before the back-trace reads it.  Calls that overlap on the GPU
// (different streams, or several host threads on one device) race on the array: serialise them.
static __device__ __align__(256) unsigned char ctcg_static_scratch[CTCG_STATIC_SCRATCH_BYTES];

namespace {

__global__ void clamp_lens_kernel(const int32_t* __restrict__ in, int B, int T, int32_t* __restrict__ out) {
  const int b = blockIdx.x * blockDim.x + threadIdx.x;
  if (b < B) out[b] = min(max(in[b], 0), T);
}

constexpr int kMaxDevices = 64;
unsigned char* g_static_ptr[kMaxDevices];  // device address of ctcg_static_scratch, per device
bool g_static_ok[kMaxDevices];
std::mutex g_static_mu;

// Device address of the scratch on the current device (set-up on the first call per device).
int static_scratch(unsigned char** out) {
  int dev = -1;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= kMaxDevices) {
    (void)cudaGetLastError();
    return CTCG_ECUDA;
  }
  std::lock_guard<std::mutex> lock(g_static_mu);
  if (!g_static_ok[dev]) {
    void* p = nullptr;
    if (cudaGetSymbolAddress(&p, ctcg_static_scratch) != cudaSuccess || !p) {
      (void)cudaGetLastError();
      return CTCG_ECUDA;
    }
    g_static_ptr[dev] = static_cast<unsigned char*>(p);
    g_static_ok[dev] = true;
  }
  *out = g_static_ptr[dev];
  return CTCG_OK;
}

}  // namespace

extern "C" int ctcg_viterbi_static(const float* log_probs, int32_t T, int32_t B, int32_t V, const int32_t* lens_dev,
                                   const ctcg_graph* graph, float* score, int32_t* frame_labels, void* stream) {
  using namespace ctcg;
  if (!log_probs || !lens_dev || !graph || !score || !frame_labels || T < 1 || B < 1 || V < 1) return CTCG_EINVAL;
  if (graph->num_graphs != 1 && graph->num_graphs != B) return CTCG_EINVAL;
  if (graph->max_label >= V) return CTCG_EINVAL;
  int dev = -1;
  if (cudaGetDevice(&dev) != cudaSuccess) {
    (void)cudaGetLastError();
    return CTCG_ECUDA;
  }
  if (!graph->dev_blob || graph->device != dev) return CTCG_ENOTUPLOADED;  // upload is set-up, done by the caller beforehand
  const int row = pad4(graph->max_states);
  const size_t lens_bytes = ((size_t)B * 4 + 255) / 256 * 256;
  const size_t need = lens_bytes + (size_t)B * (size_t)T * (size_t)row * 4;
  if (need > (size_t)CTCG_STATIC_SCRATCH_BYTES) return CTCG_ESCRATCH;
  const size_t smem = (size_t)(2 * pad4(graph->max_states) + 4 * NWARP) * 4;
  if (smem > 200 * 1024) return CTCG_EUNSUPPORTED;
  unsigned char* scratch = nullptr;
  const int st_ = static_scratch(&scratch);
  if (st_ != CTCG_OK) return st_;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  int32_t* lens = reinterpret_cast<int32_t*>(scratch);
  int32_t* back = reinterpret_cast<int32_t*>(scratch + lens_bytes);
  clamp_lens_kernel<<<(B + 127) / 128, 128, 0, st>>>(lens_dev, B, T, lens);
  if (cudaFuncSetAttribute(viterbi_kernel<false, NT>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) != cudaSuccess) {
    (void)cudaGetLastError();
    return CTCG_ECUDA;
  }
  viterbi_kernel<false, NT><<<B, NT, smem, st>>>(log_probs, T, B, V, lens, graph->dev, back, row, score, frame_labels);
  g_launches += 2;
  if (cudaGetLastError() != cudaSuccess) return CTCG_ECUDA;
  return CTCG_OK;
}
