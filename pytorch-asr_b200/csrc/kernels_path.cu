// Generic log-semiring forward-backward over CSR graphs, one CTA per utterance.
//
// Replaces PathLogSumExp.forward/backward (att_speech/fst_utils.py:400-488):
//   alpha_{t+1}[d] = LSE_{arcs into d}(w + alpha_t[src] + x[t, label])      :434-443
//   cost           = LSE_n(alpha_len[n] + final[n])                          :445
//   beta_t[s]      = LSE_{arcs out of s}(w + beta_{t+1}[dst] + x[t, label])  :456-461
//   grad[t, label] += exp(w + beta_{t+1}[dst] + x + alpha_t[s] - cost)       :463-470
// with three differences that matter for speed and accuracy on B200:
//   * CSR instead of padded (N,K) matrices: no padding arcs, no (B,N,K) temporaries;
//   * alpha/beta are renormalised every frame (alpha_hat = alpha - c_t, the
//     offsets c_t accumulated in fp64), so fp32 keeps ~1e-7 relative precision
//     on the posteriors where the reference's un-normalised fp32 loses 1e-4;
//   * everything is in the log2 domain: one MUFU.EX2 per arc, one MUFU.LG2 per state.
// The time recursion is serial per utterance; states are spread over the CTA,
// alpha lives in shared memory (double-buffered, one barrier per frame), the
// frame's scores arrive through an 8-deep cp.async ring in compact mode.
#include <cstdlib>
#include <cstring>
#include <vector>

#include "device_utils.cuh"
#include "kernels.cuh"

namespace ctcg {

unsigned long long g_launches = 0;

namespace {
struct ProfRec {
  const char* name;
  cudaEvent_t a, b;
};
bool g_prof_on = false;
std::vector<ProfRec> g_prof;
}  // namespace

ProfScope::ProfScope(const char* name, cudaStream_t st) : st_(st), idx_(-1) {
  if (!g_prof_on) return;
  ProfRec r{name, nullptr, nullptr};
  if (cudaEventCreate(&r.a) != cudaSuccess || cudaEventCreate(&r.b) != cudaSuccess) return;
  cudaEventRecord(r.a, st);
  g_prof.push_back(r);
  idx_ = (int)g_prof.size() - 1;
}
ProfScope::~ProfScope() {
  if (idx_ >= 0) cudaEventRecord(g_prof[idx_].b, st_);
}
void profile_enable(bool on) { g_prof_on = on; }
bool profile_is_on() { return g_prof_on; }
int profile_collect(char* names, size_t cap, float* ms, int max_records) {
  int n = 0;
  size_t off = 0;
  if (names && cap) names[0] = 0;
  for (ProfRec& r : g_prof) {
    cudaEventSynchronize(r.b);
    float t = 0.f;
    cudaEventElapsedTime(&t, r.a, r.b);
    if (n < max_records) {
      ms[n] = t;
      const size_t l = strlen(r.name);
      if (names && off + l + 2 < cap) {
        memcpy(names + off, r.name, l);
        names[off + l] = ';';
        names[off + l + 1] = 0;
        off += l + 1;
      }
      ++n;
    }
    cudaEventDestroy(r.a);
    cudaEventDestroy(r.b);
  }
  g_prof.clear();
  return n;
}

namespace {

constexpr int NT = 256;     // narrow graphs (numerators: <= 4 arcs per state): one state per thread
constexpr int NTWIDE = 1024;  // wide graphs (shared denominators: tens of arcs per state): one state per WARP
constexpr int RING = 8;

struct SmemPlan {
  int np, gw, xw, ar, arcs;
  size_t bytes;
};

__host__ __device__ inline SmemPlan plan_smem(int max_states, int width, int max_labels, int alpha_row,
                                              int stage_arcs, bool compact, int NWARP, bool wide = false) {
  SmemPlan p;
  p.np = pad4(max_states);
  p.gw = pad4(width);
  p.xw = compact ? pad4(max_labels) : 0;
  p.ar = compact ? alpha_row : 0;
  p.arcs = stage_arcs;
  size_t f = 2 * (size_t)p.np + 2 * (size_t)p.gw + 4 * NWARP + (size_t)RING * p.xw + (size_t)RING * p.ar;
  p.bytes = f * 4 + (size_t)stage_arcs * 24;
  if (wide) p.bytes += 2 * (size_t)(p.np + 4) * 4;  // the two CSR row-pointer arrays (one L2 round trip less per state)
  return p;
}

// Block-wide max over per-warp partials written to red[0..NWARP).
template <int NWARP>
__device__ __forceinline__ float read_block_max(const float* red) {
  float m = red[0];
#pragma unroll
  for (int i = 1; i < NWARP; ++i) m = fmaxf(m, red[i]);
  return m;
}

// WIDE: a warp owns a state and its lanes the state's arcs -- the CSR keeps them contiguous, so the arc tables
// are read in coalesced 128-byte lines and each lane keeps its first two arc scores in registers between the
// max and the sum (with one state per thread, the 51 arcs of a bi-char state are 51 strided loads per lane,
// twice).  NT threads: 256 for numerators, 1024 for the wide graphs, whose frames are latency-bound.
// MODE 2 (shared graphs with >= 512 states and transposed arc tables, GraphDev::ell_*): one state per THREAD over
// the tables -- slot k of consecutive states at consecutive addresses, a tenth of MODE 1's warp instructions (ncu:
// MODE 1 is issue-bound at ~180 warp instructions per state) -- and, where every arc into a state carries the same
// label, the gradient from the per-STATE occupancy alpha_{t+1}[d] beta_{t+1}[d] (one shared-memory atomic per
// state and frame instead of one per arc); the forward pass spills one more alpha row (row len) for it.
template <bool COMPACT, int MODE, int NT>
__global__ void __launch_bounds__(NT) path_fwd_bwd_kernel(const PathArgs a, const int stage_arcs) {
  constexpr bool WIDE = MODE == 1, ELL = MODE == 2;
  constexpr int NWARP = NT / 32;
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const GraphDev& g = a.g;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int b = blockIdx.x;
  const int gi = (g.num_graphs == 1) ? 0 : b;
  const int so = g.state_off[gi];
  const int n = g.state_off[gi + 1] - so;
  const int ro = so + gi;
  const int len = min(a.lens[b], a.T);
  const SmemPlan P = plan_smem(g.max_states, a.g_width, g.max_labels, a.alpha_row, stage_arcs, COMPACT, NWARP, WIDE);

  float* abuf = reinterpret_cast<float*>(smem_raw);
  float* grow = abuf + 2 * P.np;
  float* wred = grow + 2 * P.gw;
  float* wsum = wred + 2 * NWARP;  // per-warp partial sums of the frame's posteriors
  float* xring = wsum + 2 * NWARP;
  float* aring = xring + RING * P.xw;
  int32_t* s_in_src = reinterpret_cast<int32_t*>(aring + RING * P.ar);
  int32_t* s_in_lab = s_in_src + stage_arcs;
  float* s_in_w = reinterpret_cast<float*>(s_in_lab + stage_arcs);
  int32_t* s_out_dst = reinterpret_cast<int32_t*>(s_in_w + stage_arcs);
  int32_t* s_out_lab = s_out_dst + stage_arcs;
  float* s_out_w = reinterpret_cast<float*>(s_out_lab + stage_arcs);

  const int32_t* in_ptr = g.in_ptr + ro;
  const int32_t* out_ptr = g.out_ptr + ro;
  const int arc0 = in_ptr[0];
  const int narcs = in_ptr[n] - arc0;
  if (WIDE) {  // row pointers in shared memory (behind the arc staging area, which wide graphs do not use)
    int32_t* sp = reinterpret_cast<int32_t*>(s_out_w + stage_arcs);
    for (int i = tid; i <= n; i += NT) {
      sp[i] = in_ptr[i];
      sp[P.np + 4 + i] = out_ptr[i];
    }
    in_ptr = sp;
    out_ptr = sp + P.np + 4;
    __syncthreads();
  }
  const bool staged = stage_arcs > 0 && narcs <= stage_arcs;

  // Arc tables: shared-memory copies when they fit (numerator graphs), else global.
  const int32_t* in_src;
  const int32_t* in_lab;
  const float* in_w;
  const int32_t* out_dst;
  const int32_t* out_lab;
  const float* out_w;
  float wscale;
  if (staged) {
    for (int i = tid; i < narcs; i += NT) {
      s_in_src[i] = g.in_src[arc0 + i];
      s_in_lab[i] = COMPACT ? g.in_clab[arc0 + i] : g.in_lab[arc0 + i];
      s_in_w[i] = g.in_w[arc0 + i] * kLog2e;
      s_out_dst[i] = g.out_dst[arc0 + i];
      s_out_lab[i] = COMPACT ? g.out_clab[arc0 + i] : g.out_lab[arc0 + i];
      s_out_w[i] = g.out_w[arc0 + i] * kLog2e;
    }
    in_src = s_in_src - arc0;
    in_lab = s_in_lab - arc0;
    in_w = s_in_w - arc0;
    out_dst = s_out_dst - arc0;
    out_lab = s_out_lab - arc0;
    out_w = s_out_w - arc0;
    wscale = 1.f;
  } else {
    in_src = g.in_src;
    in_lab = COMPACT ? g.in_clab : g.in_lab;
    in_w = g.in_w;
    out_dst = g.out_dst;
    out_lab = COMPACT ? g.out_clab : g.out_lab;
    out_w = g.out_w;
    wscale = kLog2e;
  }
  const bool unit = g.unit_weights != 0;  // every arc weight is One: no weight loads
  const float* final_w = g.final_w + so;
  const float* xb = a.x + (long long)b * a.x_stride_b;
  float* arow_base = a.alpha_ws + (long long)b * (a.T + (ELL ? 1 : 0)) * a.alpha_row;
  const int nl = COMPACT ? (g.lab_off[gi + 1] - g.lab_off[gi]) : a.g_width;
  const int xcopy = COMPACT ? pad4(nl) / 4 : 0;  // 16-byte chunks per score row

  // ------------------------------------------------------------------ forward
  float* a_cur = abuf;
  float* a_nxt = abuf + P.np;
  for (int s = tid; s < n; s += NT) a_cur[s] = (s == 0) ? 0.f : neg_inf();
  for (int j = tid; j < 2 * P.gw; j += NT) grow[j] = 0.f;
  if (tid < 2 * NWARP) wred[tid] = 0.f;

  auto issue_fwd = [&](int i) {
    if (COMPACT && i < len) {
      const float* src = xb + (long long)i * a.x_stride_t;
      float* dst = xring + (i % RING) * P.xw;
      for (int c = tid; c < xcopy; c += NT) cp_async16(dst + 4 * c, src + 4 * c);
    }
    cp_async_commit();
  };
  if (COMPACT)
    for (int i = 0; i < RING - 1; ++i) issue_fwd(i);

  double c = 0.0;  // alpha_t = alpha_hat_t + c (log2 units); identical in every thread
  for (int t = 0; t < len; ++t) {
    if (COMPACT) cp_async_wait<RING - 2>();
    __syncthreads();
    if (COMPACT) issue_fwd(t + RING - 1);
    float mx = read_block_max<NWARP>(wred + (t & 1) * NWARP);
    if (!(mx > neg_inf())) mx = 0.f;
    c += (double)mx;
    const float* xrow = COMPACT ? (xring + (t % RING) * P.xw) : (xb + (long long)t * a.x_stride_t);
    float* arow = arow_base + (long long)t * a.alpha_row;
    if (tid == 0) {
      const float hi = (float)c;
      arow[a.alpha_row - 2] = hi;
      arow[a.alpha_row - 1] = (float)(c - (double)hi);
    }
    float lm = neg_inf();
    auto fscore = [&](int e) {
      const float w = unit ? 0.f : in_w[e] * wscale;
      return w + (a_cur[in_src[e]] - mx) + xrow[in_lab[e]] * a.x_scale;
    };
    if (ELL) {
      const int enp = g.ell_np, K = g.ell_kin;
      for (int s = tid; s < n; s += NT) {
        arow[s] = a_cur[s] - mx;
        float M = neg_inf();
#pragma unroll 4
        for (int k = 0; k < K; ++k) {
          const int i = k * enp + s;
          M = fmaxf(M, g.ell_in_w[i] * kLog2e + (a_cur[g.ell_in_src[i]] - mx) + xrow[g.ell_in_lab[i]] * a.x_scale);
        }
        float r = neg_inf();
        if (M > neg_inf()) {
          float acc = 0.f;
#pragma unroll 4
          for (int k = 0; k < K; ++k) {
            const int i = k * enp + s;
            acc += ex2(g.ell_in_w[i] * kLog2e + (a_cur[g.ell_in_src[i]] - mx) + xrow[g.ell_in_lab[i]] * a.x_scale - M);
          }
          r = M + lg2(acc);
        }
        a_nxt[s] = r;
        lm = fmaxf(lm, r);
      }
    } else if (WIDE) {
      // two states per iteration: their (L2-latency) pointer and arc loads overlap
      for (int s = warp; s < n; s += 2 * NWARP) {
        const int s2 = s + NWARP;
        const bool two = s2 < n;
        const int lo = in_ptr[s], hi = in_ptr[s + 1];
        const int lo2 = two ? in_ptr[s2] : 0, hi2 = two ? in_ptr[s2 + 1] : 0;
        if (lane == 0) {
          arow[s] = a_cur[s] - mx;
          if (two) arow[s2] = a_cur[s2] - mx;
        }
        const int e0 = lo + lane, f0 = lo2 + lane;
        const float v0 = e0 < hi ? fscore(e0) : neg_inf();
        const float u0 = f0 < hi2 ? fscore(f0) : neg_inf();
        const float v1 = e0 + 32 < hi ? fscore(e0 + 32) : neg_inf();
        const float u1 = f0 + 32 < hi2 ? fscore(f0 + 32) : neg_inf();
        float M = fmaxf(v0, v1), M2 = fmaxf(u0, u1);
        for (int e = e0 + 64; e < hi; e += 32) M = fmaxf(M, fscore(e));
        for (int e = f0 + 64; e < hi2; e += 32) M2 = fmaxf(M2, fscore(e));
        M = warp_max(M);
        M2 = warp_max(M2);
        float acc = 0.f, acc2 = 0.f;
        if (M > neg_inf()) {
          acc = ex2(v0 - M) + ex2(v1 - M);
          for (int e = e0 + 64; e < hi; e += 32) acc += ex2(fscore(e) - M);
        }
        if (M2 > neg_inf()) {
          acc2 = ex2(u0 - M2) + ex2(u1 - M2);
          for (int e = f0 + 64; e < hi2; e += 32) acc2 += ex2(fscore(e) - M2);
        }
        acc = warp_sum(acc);
        acc2 = warp_sum(acc2);
        const float r = M > neg_inf() ? M + lg2(acc) : neg_inf();
        const float r2 = M2 > neg_inf() ? M2 + lg2(acc2) : neg_inf();
        if (lane == 0) {
          a_nxt[s] = r;
          if (two) a_nxt[s2] = r2;
        }
        lm = fmaxf(lm, fmaxf(r, r2));
      }
    } else {
      for (int s = tid; s < n; s += NT) {
        arow[s] = a_cur[s] - mx;
        const int lo = in_ptr[s], hi = in_ptr[s + 1];
        float M = neg_inf();
        for (int e = lo; e < hi; ++e) M = fmaxf(M, fscore(e));
        float r = neg_inf();
        if (M > neg_inf()) {
          float acc = 0.f;
          for (int e = lo; e < hi; ++e) acc += ex2(fscore(e) - M);
          r = M + lg2(acc);
        }
        a_nxt[s] = r;
        lm = fmaxf(lm, r);
      }
    }
    lm = warp_max(lm);
    if (lane == 0) wred[((t + 1) & 1) * NWARP + warp] = lm;
    float* tmp = a_cur;
    a_cur = a_nxt;
    a_nxt = tmp;
  }
  if (COMPACT) cp_async_wait<0>();
  __syncthreads();
  // terminal: cost = c_len + LSE_n(alpha_hat_len[n] + final[n])
  float mx = (len > 0) ? read_block_max<NWARP>(wred + (len & 1) * NWARP) : 0.f;
  if (!(mx > neg_inf())) mx = 0.f;
  c += (double)mx;
  if (ELL) {  // alpha_hat_len and its offset: the occupancy of the last frame needs them
    float* arow = arow_base + (long long)len * a.alpha_row;
    for (int s = tid; s < n; s += NT) arow[s] = a_cur[s] - mx;
    if (tid == 0) {
      const float hi = (float)c;
      arow[a.alpha_row - 2] = hi;
      arow[a.alpha_row - 1] = (float)(c - (double)hi);
    }
  }
  float* red2 = wred;  // both parities are free between the passes
  float tm = neg_inf();
  for (int s = tid; s < n; s += NT) tm = fmaxf(tm, (a_cur[s] - mx) + final_w[s] * kLog2e);
  tm = warp_max(tm);
  __syncthreads();
  if (lane == 0) red2[warp] = tm;
  __syncthreads();
  float TM = read_block_max<NWARP>(red2);
  __syncthreads();
  float ts = 0.f;
  if (TM > neg_inf())
    for (int s = tid; s < n; s += NT) ts += ex2((a_cur[s] - mx) + final_w[s] * kLog2e - TM);
  ts = warp_sum(ts);
  if (lane == 0) red2[warp] = ts;
  __syncthreads();
  float TS = 0.f;
#pragma unroll
  for (int i = 0; i < NWARP; ++i) TS += red2[i];
  __syncthreads();
  if (tid < NWARP) red2[tid] = 0.f;
  const bool feasible = TM > neg_inf() && TS > 0.f;
  const double cost2 = feasible ? (c + (double)TM + (double)lg2(TS)) : -1e20 / kLn2;
  if (tid == 0) a.cost[b] = feasible ? cost2 * kLn2 : -1e20;
  if (!a.write_grad) return;

  // ----------------------------------------------------------------- backward
  const float gmul = a.g_sign * (a.g_scale ? a.g_scale[b] : 1.f);
  float* gb = a.gout + (long long)b * a.g_stride_b;
  if (!COMPACT) {
    // dense output: rows at or beyond the utterance's end are zero
    for (int t = len; t < a.T; ++t)
      for (int j = tid; j < a.g_width; j += NT) gb[(long long)t * a.g_stride_t + j] = 0.f;
  }
  if (!feasible) {
    for (int t = 0; t < len; ++t)
      for (int j = tid; j < nl; j += NT) gb[(long long)t * a.g_stride_t + j] = 0.f;
    return;
  }
  float* b_cur = a_nxt;  // the buffer not holding alpha_len
  float* b_nxt = a_cur;
  __syncthreads();
  {
    float lm = neg_inf();
    for (int s = tid; s < n; s += NT) {
      const float v = final_w[s] * kLog2e;
      b_cur[s] = v;
      lm = fmaxf(lm, v);
    }
    lm = warp_max(lm);
    if (lane == 0) wred[warp] = lm;  // parity 0 == iteration 0
  }
  auto issue_bwd = [&](int i) {
    if (COMPACT && i < len) {
      const int t = len - 1 - i;
      const float* src = xb + (long long)t * a.x_stride_t;
      float* dst = xring + (i % RING) * P.xw;
      for (int cidx = tid; cidx < xcopy; cidx += NT) cp_async16(dst + 4 * cidx, src + 4 * cidx);
      const float* asrc = arow_base + (long long)t * a.alpha_row;
      float* adst = aring + (i % RING) * P.ar;
      for (int cidx = tid; cidx < a.alpha_row / 4; cidx += NT) cp_async16(adst + 4 * cidx, asrc + 4 * cidx);
    }
    cp_async_commit();
  };
  if (COMPACT)
    for (int i = 0; i < RING - 1; ++i) issue_bwd(i);

  double d = 0.0;  // beta_{t+1} = beta_hat + d
  for (int i = 0; i < len; ++i) {
    const int t = len - 1 - i;
    if (COMPACT) cp_async_wait<RING - 2>();
    __syncthreads();
    if (COMPACT) issue_bwd(i + RING - 1);
    // flush the posterior row of the previous iteration (frame t+1), double-buffered
    if (i > 0) {
      float* gr = grow + ((i - 1) & 1) * P.gw;
      float* dst = gb + (long long)(t + 1) * a.g_stride_t;
      // Every accepting path crosses exactly one arc per frame, so the frame's arc
      // posteriors sum to one: dividing by their computed sum removes the rounding
      // drift that alpha_hat/beta_hat/cost accumulate over hundreds of frames.
      float tot = 0.f;
#pragma unroll
      for (int w = 0; w < NWARP; ++w) tot += wsum[((i - 1) & 1) * NWARP + w];
      const float sc = tot > 0.f ? gmul / tot : 0.f;
      for (int j = tid; j < nl; j += NT) {
        dst[j] = gr[j] * sc;
        gr[j] = 0.f;
      }
    }
    float mxb = read_block_max<NWARP>(wred + (i & 1) * NWARP);
    if (!(mxb > neg_inf())) mxb = 0.f;
    d += (double)mxb;
    const float* xrow = COMPACT ? (xring + (i % RING) * P.xw) : (xb + (long long)t * a.x_stride_t);
    const float* arow = COMPACT ? (aring + (i % RING) * P.ar) : (arow_base + (long long)t * a.alpha_row);
    const double ct = (double)arow[a.alpha_row - 2] + (double)arow[a.alpha_row - 1];
    const float Kt = (float)(ct + d - cost2);
    float* gr = grow + (i & 1) * P.gw;
    float lm = neg_inf();
    float gsum = 0.f;
    auto bscore = [&](int e) {
      const float w = unit ? 0.f : out_w[e] * wscale;
      return w + (b_cur[out_dst[e]] - mxb) + xrow[out_lab[e]] * a.x_scale;
    };
    if (ELL) {
      const int enp = g.ell_np, K = g.ell_kout;
      const bool uniform = g.in_label_uniform != 0;
      if (uniform) {
        // posterior of frame t by state: gamma_t[label into d] += alpha_{t+1}[d] beta_{t+1}[d] / Z
        const float* arow1 = arow_base + (long long)(t + 1) * a.alpha_row;
        const double c1 = (double)arow1[a.alpha_row - 2] + (double)arow1[a.alpha_row - 1];
        const float K1 = (float)(c1 + d - cost2);
        for (int s = tid; s < n; s += NT) {
          const float occ = ex2(arow1[s] + (b_cur[s] - mxb) + K1);
          if (occ > 0.f) {
            atomicAdd(&gr[g.ell_in_lab[s]], occ);  // slot 0: the label every arc into s carries
            gsum += occ;
          }
        }
      }
      for (int s = tid; s < n; s += NT) {
        float M = neg_inf();
#pragma unroll 4
        for (int k = 0; k < K; ++k) {
          const int i = k * enp + s;
          M = fmaxf(M, g.ell_out_w[i] * kLog2e + (b_cur[g.ell_out_dst[i]] - mxb) + xrow[g.ell_out_lab[i]] * a.x_scale);
        }
        float r = neg_inf();
        if (M > neg_inf()) {
          const float f = uniform ? 0.f : ex2(M + arow[s] + Kt);
          float acc = 0.f;
#pragma unroll 4
          for (int k = 0; k < K; ++k) {
            const int i = k * enp + s;
            const int lab = g.ell_out_lab[i];
            const float ev = ex2(g.ell_out_w[i] * kLog2e + (b_cur[g.ell_out_dst[i]] - mxb) + xrow[lab] * a.x_scale - M);
            acc += ev;
            if (f > 0.f && ev > 0.f) {
              atomicAdd(&gr[lab], ev * f);
              gsum += ev * f;
            }
          }
          r = M + lg2(acc);
        }
        b_nxt[s] = r;
        lm = fmaxf(lm, r);
      }
    } else if (WIDE) {
      for (int s = warp; s < n; s += NWARP) {
        const int lo = out_ptr[s], hi = out_ptr[s + 1];
        const int e0 = lo + lane;
        const float v0 = e0 < hi ? bscore(e0) : neg_inf();
        const float v1 = e0 + 32 < hi ? bscore(e0 + 32) : neg_inf();
        float M = fmaxf(v0, v1);
        for (int e = e0 + 64; e < hi; e += 32) M = fmaxf(M, bscore(e));
        M = warp_max(M);
        float r = neg_inf();
        if (M > neg_inf()) {
          const float f = ex2(M + arow[s] + Kt);  // alpha_hat_t[s] = -inf -> 0
          const float ev0 = ex2(v0 - M), ev1 = ex2(v1 - M);
          float acc = ev0 + ev1;
          if (f > 0.f) {
            if (e0 < hi) atomicAdd(&gr[out_lab[e0]], ev0 * f);
            if (e0 + 32 < hi) atomicAdd(&gr[out_lab[e0 + 32]], ev1 * f);
            gsum += acc * f;
          }
          for (int e = e0 + 64; e < hi; e += 32) {
            const float ev = ex2(bscore(e) - M);
            acc += ev;
            if (f > 0.f) {
              atomicAdd(&gr[out_lab[e]], ev * f);
              gsum += ev * f;
            }
          }
          r = M + lg2(warp_sum(acc));
        }
        if (lane == 0) b_nxt[s] = r;
        lm = fmaxf(lm, r);
      }
    } else {
      for (int s = tid; s < n; s += NT) {
        const int lo = out_ptr[s], hi = out_ptr[s + 1];
        float M = neg_inf();
        for (int e = lo; e < hi; ++e) M = fmaxf(M, bscore(e));
        float r = neg_inf();
        if (M > neg_inf()) {
          const float f = ex2(M + arow[s] + Kt);  // alpha_hat_t[s] = -inf -> 0
          float acc = 0.f;
          for (int e = lo; e < hi; ++e) {
            const float ev = ex2(bscore(e) - M);
            acc += ev;
            if (f > 0.f) {
              atomicAdd(&gr[out_lab[e]], ev * f);
              gsum += ev * f;
            }
          }
          r = M + lg2(acc);
        }
        b_nxt[s] = r;
        lm = fmaxf(lm, r);
      }
    }
    lm = warp_max(lm);
    gsum = warp_sum(gsum);
    if (lane == 0) {
      wred[((i + 1) & 1) * NWARP + warp] = lm;
      wsum[(i & 1) * NWARP + warp] = gsum;
    }
    float* tmp = b_cur;
    b_cur = b_nxt;
    b_nxt = tmp;
  }
  if (COMPACT) cp_async_wait<0>();
  __syncthreads();
  if (len > 0) {
    float* gr = grow + ((len - 1) & 1) * P.gw;
    float tot = 0.f;
#pragma unroll
    for (int w = 0; w < NWARP; ++w) tot += wsum[((len - 1) & 1) * NWARP + w];
    const float sc = tot > 0.f ? gmul / tot : 0.f;
    for (int j = tid; j < nl; j += NT) gb[j] = gr[j] * sc;
  }
}

}  // namespace

template <bool COMPACT, int MODE, int NTT>
cudaError_t launch_path_variant(const PathArgs& a, int stage, size_t bytes, cudaStream_t st) {
  cudaError_t e = cudaFuncSetAttribute(path_fwd_bwd_kernel<COMPACT, MODE, NTT>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       (int)bytes);
  if (e != cudaSuccess) return e;
  path_fwd_bwd_kernel<COMPACT, MODE, NTT><<<a.B, NTT, bytes, st>>>(a, stage);
  return cudaSuccess;
}

cudaError_t launch_path_fwd_bwd(const PathArgs& a, bool compact, cudaStream_t st) {
  // wide graphs (a shared denominator with tens of arcs per state): one warp per state, 1024 threads
  static const bool narrow_only = [] {   // CTCG_PATH_NARROW=1: the one-state-per-thread kernel for every graph (cross-checks)
    const char* e = getenv("CTCG_PATH_NARROW");
    return e && e[0] == '1';
  }();
  const bool wide = !narrow_only && (long long)a.g.max_arcs >= 16LL * (long long)max(a.g.max_states, 1);
  const int nwarp = (wide ? NTWIDE : NT) / 32;
  // stage arc tables in shared memory when every graph's arcs fit in 48 KB
  int stage = 0;
  const SmemPlan base = plan_smem(a.g.max_states, a.g_width, a.g.max_labels, a.alpha_row, 0, compact, nwarp, wide);
  const size_t budget = 200 * 1024;
  if (base.bytes > budget) return cudaErrorInvalidConfiguration;
  // a.g.max_in_degree * max_states bounds the arc count of one graph
  long long bound = (long long)a.g.max_states * (long long)max(a.g.max_in_degree, 1);
  if (bound * 24 + (long long)base.bytes <= (long long)budget && bound <= 8192) stage = (int)bound;
  const SmemPlan P = plan_smem(a.g.max_states, a.g_width, a.g.max_labels, a.alpha_row, stage, compact, nwarp, wide);
  cudaError_t e;
  ProfScope prof(compact ? "path_fwd_bwd<compact>" : "path_fwd_bwd<dense>", st);
  // transposed arc tables and enough states to fill the CTA: one state per thread over them (dense mode: the
  // shared graphs; the caller's alpha workspace holds T + 1 rows per utterance, a.alpha_rows_plus1)
  const bool ell = wide && !compact && a.g.ell_in_src != nullptr && a.g.max_states >= 512 && a.alpha_rows_plus1;
  if (compact) e = wide ? launch_path_variant<true, 1, NTWIDE>(a, stage, P.bytes, st) : launch_path_variant<true, 0, NT>(a, stage, P.bytes, st);
  else if (ell) e = launch_path_variant<false, 2, NTWIDE>(a, stage, P.bytes, st);
  else e = wide ? launch_path_variant<false, 1, NTWIDE>(a, stage, P.bytes, st) : launch_path_variant<false, 0, NT>(a, stage, P.bytes, st);
  if (e != cudaSuccess) return e;
  ++g_launches;
  return cudaGetLastError();
}

}  // namespace ctcg
