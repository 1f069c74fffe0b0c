// Numerator (per-utterance training graph) forward-backward on the compact label columns,
// log2 domain, sm_100a: the alpha and the beta recursion on separate CTAs at the same time
// (num_sweep_kernel), then a time-parallel posterior kernel (num_combine_kernel).
//
// Reference semantics: PathLogSumExp.forward/backward (att_speech/fst_utils.py:400-488)
// on the graphs of BaseGraphGen.get_training_matrices (fst_utils.py:647-660): N = 2U+1
// states with at most 3 arcs into / out of a state.
//
// Why the log domain here and plain scaled numbers in the denominator kernels: a
// training graph is a chain, so alpha_t of the states that carry the posterior mass can
// sit hundreds of binary orders below max_s alpha_t[s] (measured: below 2^-126 at T=400
// on random logits) -- a per-frame rescale cannot keep alpha * beta representable.
// In the decoding graph every state reaches every other within two frames, which
// bounds that ratio by the frame's logit spread.
//
// Input: the compact label columns xg[b][t][j] = logits[t][b][lab_list_b[j]] (natural log;
// multiplied by log2(e) inside an fma so that the input is rounded once) written by
// gather_kernel; output: cost[b] (natural log) and either the compact posterior rows
// gout[b][t][j] (plain CTC: the row kernels fold them into the dense gradient) or the
// subtraction from the dense gradient rows directly (with a factorised denominator).
#include <algorithm>
#include <cstdlib>

#include "device_utils.cuh"
#include "kernels.cuh"

namespace ctcg {
namespace {

constexpr int NS = 256;                 // state threads
constexpr int W_LAB = NS / 32;          // first warp after the state warps (the io warp)
constexpr int XR = 8;          // cp.async ring depth (frames)
constexpr int DEG = 4;         // arcs packed per state
constexpr int HEAVY = 12;      // labels with more arcs are summed by the whole warp
constexpr unsigned NOARC = 0xffffffffu;
constexpr int IMIN = -2147483647 - 1;

// 2^(a + b + k) with the rounding error of a + b carried along (two-sum): a and b sit
// tens of binary orders below zero where one fp32 ulp is ~1e-5, k brings the sum back
// to <= 0.
__device__ __forceinline__ float post_term(float a, float b, float k) {
  const float s = a + b;
  const float bb = s - a;
  float err = (a - (s - bb)) + (b - bb);
  err = (s > neg_inf()) ? err : 0.f;
  return ex2((s + k) + err);
}


// floor(max over the warp) as an integer; -inf saturates to IMIN
__device__ __forceinline__ int warp_floor_max(float v) {
  return __reduce_max_sync(0xffffffffu, __float2int_rd(v));
}
__device__ __forceinline__ float norm_of(int m) { return (m == IMIN) ? 0.f : (float)m; }
__device__ __forceinline__ int next3(int r) { return r == 2 ? 0 : r + 1; }

__device__ __forceinline__ float lds_f(unsigned a) {
  float v;
  asm volatile("ld.shared.f32 %0, [%1];" : "=f"(v) : "r"(a));
  return v;
}
__device__ __forceinline__ void sts_f(unsigned a, float v) { asm volatile("st.shared.f32 [%0], %1;" ::"r"(a), "f"(v)); }
__device__ __forceinline__ unsigned sm_u32(const void* p) { return static_cast<unsigned>(__cvta_generic_to_shared(p)); }

// =====================================================================================
// The alpha and the beta recursion run AT THE SAME TIME on separate CTAs (grid (B, 2)), each
// spilling its normalised rows; the posteriors are then formed by a time-parallel combine
// kernel: the serial chain is T frames instead of 2T and the posterior work is off it.
//
//   num_sweep_kernel    DIR 0: alpha_hat_t rows (+ fp64 offset c_t) and the cost, arcs by destination
//                       DIR 1: beta_hat_{t+1} rows (+ offset d_t), arcs by source, frames descending
//   num_combine_kernel  one warp per frame: gamma_t[j] = sum_{arcs e labelled j}
//                       2^(alpha_hat_t[src] + beta_hat_{t+1}[dst] + x_t[j] + w_e + c_t + d_t - cost),
//                       fixed summation order (arcs grouped by label on the host), row normalised to one
// =====================================================================================
constexpr int SW_TH = NS + 32;  // state warps + io warp

struct SweepPlan {
  int pk, w, a0, a1, fin, red, xring, bytes;
};
__host__ __device__ inline SweepPlan sweep_plan(int maxN, int nlp, bool unit) {
  SweepPlan p;
  int o = 0;
  auto take = [&](int bytes) {
    const int at = o;
    o += (bytes + 15) & ~15;
    return at;
  };
  p.pk = take(maxN * DEG * 4);
  p.w = unit ? 0 : take(maxN * DEG * 4);
  p.a0 = take((maxN + 1) * 4);
  p.a1 = take((maxN + 1) * 4);
  p.fin = take(maxN * 4);
  p.red = take((16 + 2 * (SW_TH / 32)) * 4);
  p.xring = take(XR * nlp * 4);
  p.bytes = o;
  return p;
}

// DK: arcs per state that exist anywhere in the batch (3 for the CTC training graphs, fst_utils.py:592-605).
template <int DK>
__device__ __forceinline__ float lseK(const float (&v)[DK]) {
  float M = v[0];
#pragma unroll
  for (int k = 1; k < DK; ++k) M = fmaxf(M, v[k]);
  const float Ms = (M > neg_inf()) ? M : 0.f;
  float s = 0.f;
#pragma unroll
  for (int k = 0; k < DK; ++k) s += ex2(v[k] - Ms);
  return Ms + lg2(s);  // lg2(0) = -inf
}

template <bool UNIT, int SPT, int DK>
__global__ void __launch_bounds__(SW_TH) num_sweep_kernel(const PathArgs a, const int maxN) {
  extern __shared__ __align__(16) unsigned char sm[];
  const GraphDev& g = a.g;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int b = blockIdx.x;
  const int DIR = blockIdx.y;
  const int gi = (g.num_graphs == 1) ? 0 : b;
  const int so = g.state_off[gi];
  const int n = g.state_off[gi + 1] - so;
  const int ro = so + gi;
  const int len = min(a.lens[b], a.T);
  const int nlp = (int)a.x_stride_t, arow = a.alpha_row;
  const SweepPlan P = sweep_plan(maxN, nlp, UNIT);
  uint4* pkv = reinterpret_cast<uint4*>(sm + P.pk);
  float4* wv = reinterpret_cast<float4*>(sm + P.w);
  float* abuf0 = reinterpret_cast<float*>(sm + P.a0);
  float* abuf1 = reinterpret_cast<float*>(sm + P.a1);
  float* fin = reinterpret_cast<float*>(sm + P.fin);
  int* wmax = reinterpret_cast<int*>(sm + P.red);
  float* scr = reinterpret_cast<float*>(wmax + 16);
  float* xring = reinterpret_cast<float*>(sm + P.xring);
  constexpr int NW = SW_TH / 32;

  const int32_t* ptr = DIR ? g.out_ptr : g.in_ptr;
  const int32_t* oth = DIR ? g.out_dst : g.in_src;
  const int32_t* clab = DIR ? g.out_clab : g.in_clab;
  const float* wgt = DIR ? g.out_w : g.in_w;
  for (int s = tid; s < n; s += SW_TH) {
    unsigned pi[DEG];
    float wi[DEG];
#pragma unroll
    for (int k = 0; k < DEG; ++k) {
      const int e = ptr[ro + s] + k;
      const bool v = e < ptr[ro + s + 1];
      pi[k] = v ? ((unsigned)oth[e] | ((unsigned)clab[e] << 16)) : NOARC;
      wi[k] = (v && !UNIT) ? wgt[e] * kLog2e : 0.f;
    }
    pkv[s] = make_uint4(pi[0], pi[1], pi[2], pi[3]);
    if (!UNIT) wv[s] = make_float4(wi[0], wi[1], wi[2], wi[3]);
    const float f = g.final_w[so + s];
    fin[s] = (f > kDeadWeight) ? f * kLog2e : neg_inf();
    abuf0[s] = DIR ? fin[s] : ((s == 0) ? 0.f : neg_inf());
  }
  if (tid < 3) wmax[tid] = (tid == 0) ? 0 : IMIN;  // the vector entering the first frame is not normalised
  if (tid == 0) abuf0[maxN] = abuf1[maxN] = neg_inf();
  const float* xb = a.x + (long long)b * a.x_stride_b;
  float* rows = (DIR ? a.beta_ws : a.alpha_ws) + (long long)b * a.T * arow;
  const int xchunks = nlp / 4;
  __syncthreads();
  auto frame_t = [&](int i) { return DIR ? len - 1 - i : i; };
  const long long rstep = DIR ? -(long long)arow : (long long)arow;

  float* a_cur = abuf0;
  float* a_nxt = abuf1;
  double c = 0.0;  // io warp: vector = normalised vector + c (log2 units)
  if (SPT > 0 && warp < W_LAB) {
    constexpr int NSP = SPT > 0 ? SPT : 1;
    unsigned so4[NSP][DK], lo4[NSP][DK];
    float wi[NSP][DK];
    float* arp[NSP];
#pragma unroll
    for (int i = 0; i < NSP; ++i) {
      const int s = tid + i * NS;
      const uint4 pk = (s < n) ? pkv[s] : make_uint4(NOARC, NOARC, NOARC, NOARC);
      const unsigned pkk[DEG] = {pk.x, pk.y, pk.z, pk.w};
      const float4 w = (!UNIT && s < n) ? wv[s] : make_float4(0.f, 0.f, 0.f, 0.f);
      const float ww[DEG] = {w.x, w.y, w.z, w.w};
#pragma unroll
      for (int k = 0; k < DK; ++k) {
        so4[i][k] = (pkk[k] == NOARC) ? (unsigned)maxN * 4u : (pkk[k] & 0xffffu) * 4u;
        lo4[i][k] = (pkk[k] == NOARC) ? 0u : (pkk[k] >> 16) * 4u;
        wi[i][k] = ww[k];
      }
      arp[i] = rows + (long long)frame_t(0) * arow + (s < n ? s : 0);
    }
    unsigned ac = sm_u32(a_cur), an = sm_u32(a_nxt);
    const unsigned xr0 = sm_u32(xring), wm_a = sm_u32(wmax);
    int r = 0, slot = 0;
    for (int t = 0; t < len; ++t) {
      __syncthreads();
      int wmi;
      asm volatile("ld.shared.s32 %0, [%1];" : "=r"(wmi) : "r"(wm_a + r * 4));
      const float mx = norm_of(wmi);
      r = next3(r);
      const unsigned xa = xr0 + slot * nlp * 4;
      slot = (slot + 1 == XR) ? 0 : slot + 1;
      float lm = neg_inf();
#pragma unroll
      for (int i = 0; i < NSP; ++i) {
        const int s = tid + i * NS;
        float v[DK];
#pragma unroll
        for (int k = 0; k < DK; ++k) {
          v[k] = fmaf(lds_f(xa + lo4[i][k]), kLog2e, lds_f(ac + so4[i][k]) - mx);
          if (!UNIT) v[k] += wi[i][k];
        }
        const float res = lseK<DK>(v);
        if (s < n) {
          *arp[i] = lds_f(ac + s * 4) - mx;
          sts_f(an + s * 4, res);
        }
        arp[i] += rstep;
        lm = fmaxf(lm, res);
      }
      const int wm = warp_floor_max(lm);
      if (lane == 0) atomicMax(&wmax[r], wm);
      const unsigned tmp = ac;
      ac = an;
      an = tmp;
    }
  } else if (warp < W_LAB) {
    int r = 0;
    for (int t = 0; t < len; ++t) {
      __syncthreads();
      const float mx = norm_of(wmax[r]);
      r = next3(r);
      const float* xrow = xring + (t % XR) * nlp;
      float* ar = rows + (long long)frame_t(t) * arow;
      float lm = neg_inf();
      for (int s = tid; s < n; s += NS) {
        ar[s] = a_cur[s] - mx;
        const uint4 pk = pkv[s];
        const unsigned pkk[DEG] = {pk.x, pk.y, pk.z, pk.w};
        float v[DEG];
#pragma unroll
        for (int k = 0; k < DEG; ++k) {
          const unsigned q = (pkk[k] == NOARC) ? 0u : pkk[k];
          const float val = fmaf(xrow[q >> 16], kLog2e, a_cur[q & 0xffffu] - mx);
          v[k] = (pkk[k] == NOARC) ? neg_inf() : val;
        }
        if (!UNIT) {
          const float4 w = wv[s];
          v[0] += w.x;
          v[1] += w.y;
          v[2] += w.z;
          v[3] += w.w;
        }
        const float res = lseK<DEG>(v);
        a_nxt[s] = res;
        lm = fmaxf(lm, res);
      }
      const int wm = warp_floor_max(lm);
      if (lane == 0) atomicMax(&wmax[r], wm);
      float* tmp = a_cur;
      a_cur = a_nxt;
      a_nxt = tmp;
    }
  } else {  // io warp
    auto issue_x = [&](int i, int slot) {
      if (i >= 0 && i < len) {
        const float* src = xb + (long long)frame_t(i) * nlp;
        float* dst = xring + slot * nlp;
        for (int cc = lane; cc < xchunks; cc += 32) cp_async16(dst + 4 * cc, src + 4 * cc);
      }
      cp_async_commit();
    };
    for (int i = 0; i < XR - 1; ++i) issue_x(i, i);
    cp_async_wait<XR - 2>();
    int r = 0;
    for (int t = 0; t < len; ++t) {
      __syncthreads();
      issue_x(t + XR - 1, (t + XR - 1) % XR);
      c += (double)norm_of(wmax[r]);
      r = next3(r);
      if (lane == 0) {
        float* ar = rows + (long long)frame_t(t) * arow;
        const float hi = (float)c;
        ar[arow - 2] = hi;
        ar[arow - 1] = (float)(c - (double)hi);
        wmax[next3(r)] = IMIN;  // the slot the state warps fill during the next frame
      }
      cp_async_wait<XR - 2>();
    }
    cp_async_wait<0>();
  }
  if (DIR) return;
  a_cur = (len & 1) ? abuf1 : abuf0;  // alpha_len (the register path swapped addresses, not these pointers)
  __syncthreads();
  // terminal: cost = c_len + LSE_n(alpha_hat_len[n] + final[n])
  float tm = neg_inf();
  for (int s = tid; s < n; s += SW_TH) tm = fmaxf(tm, a_cur[s] + fin[s]);
  tm = warp_max(tm);
  if (lane == 0) scr[warp] = tm;
  __syncthreads();
  float TM = scr[0];
#pragma unroll
  for (int w = 1; w < NW; ++w) TM = fmaxf(TM, scr[w]);
  float ts = 0.f;
  if (TM > neg_inf())
    for (int s = tid; s < n; s += SW_TH) ts += ex2(a_cur[s] + fin[s] - TM);
  ts = warp_sum(ts);
  if (lane == 0) scr[NW + warp] = ts;
  __syncthreads();
  float TS = 0.f;
#pragma unroll
  for (int w = 0; w < NW; ++w) TS += scr[NW + w];
  const bool feasible = TM > neg_inf() && TS > 0.f;
  if (warp == W_LAB && lane == 0) {  // the io warp holds c
    const double cost2 = c + (double)TM + (double)lg2(TS);
    a.cost[b] = feasible ? cost2 * kLn2 : -1e20;
  }
}

constexpr int CB_W = 8;             // warps per combine CTA, one frame at a time each
constexpr int CB_FPW = 4;           // frames per warp: the graph tables are staged once per CB_W * CB_FPW frames
constexpr int CB_FR = CB_W * CB_FPW;
constexpr int CB_TH = CB_W * 32;

// With a.scatter_grad set the normalised posteriors are subtracted from the dense gradient rows the
// denominator wrote (grad[t][b][lab_list[j]] -= scale_b * gamma_t[j], the scatter_sub step fused in:
// every (t, b, label) is touched by exactly one lane, no atomics); otherwise the compact rows
// gout[b][t][j] are written for the row kernels of plain CTC.
template <bool UNIT>
__global__ void __launch_bounds__(CB_TH, 3) num_combine_kernel(const PathArgs a, const int maxA, const int maxL) {
  extern __shared__ __align__(16) unsigned char sm[];
  const GraphDev& g = a.g;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  // utterances in descending order: with the scatter fused in, the gradient rows written last by the
  // denominator's combine kernel (the highest b) are the ones most likely to be L2 resident still
  const int b = a.B - 1 - (int)blockIdx.y;
  const int len = min(a.lens[b], a.T);
  if (blockIdx.x * CB_FR >= len) return;
  const int gi = (g.num_graphs == 1) ? 0 : b;
  const int ro = g.state_off[gi] + gi;
  const int n = g.state_off[gi + 1] - g.state_off[gi];
  const int lo = g.lab_off[gi];
  const int nl = g.lab_off[gi + 1] - lo;
  const int nlp = (int)a.x_stride_t, arow = a.alpha_row;
  int o = 0;
  auto take = [&](int bytes) {
    const int at = o;
    o += (bytes + 15) & ~15;
    return at;
  };
  int* lab_ptr = reinterpret_cast<int*>(sm + take((maxL + 1) * 4));
  unsigned* lab_pk = reinterpret_cast<unsigned*>(sm + take(maxA * 4));
  float* lab_w = reinterpret_cast<float*>(sm + (UNIT ? 0 : take(maxA * 4)));
  // per warp: two staging buffers {alpha row, beta row, x row} (the next frame's rows are copied in with
  // cp.async while this frame's posteriors are summed) and one posterior row
  const int stg = 2 * arow + nlp;
  float* rows_s = reinterpret_cast<float*>(sm + take(CB_W * (2 * stg + nlp) * 4));
  const int arc0 = g.in_ptr[ro];
  const int narcs = g.in_ptr[ro + n] - arc0;
  for (int j = tid; j <= nl; j += CB_TH) lab_ptr[j] = g.lab_ptr[lo + gi + j] - arc0;
  for (int e = tid; e < narcs; e += CB_TH) {
    lab_pk[e] = (unsigned)g.lab_src[arc0 + e] | ((unsigned)g.lab_dst[arc0 + e] << 16);
    if (!UNIT) lab_w[e] = g.lab_w[arc0 + e] * kLog2e;
  }
  __syncthreads();
  float* wbase = rows_s + warp * (2 * stg + nlp);
  float* gt = wbase + 2 * stg;
  const double costn = a.cost[b];
  const bool feasible = costn > -1e19;
  const float gmul = a.g_sign * (a.g_scale ? a.g_scale[b] : 1.f);
  const int V = a.scatter_V;
  auto stage = [&](int q) {  // always commits a group (possibly empty), so that the wait counts stay uniform
    const int t = blockIdx.x * CB_FR + q * CB_W + warp;
    if (q < CB_FPW && t < len && feasible) {
      float* d = wbase + (q & 1) * stg;
      const float* as = a.alpha_ws + ((long long)b * a.T + t) * arow;
      const float* bs = a.beta_ws + ((long long)b * a.T + t) * arow;
      const float* xs = a.x + (long long)b * a.x_stride_b + (long long)t * nlp;
      for (int i = lane; i < arow / 4; i += 32) {
        cp_async16(d + 4 * i, as + 4 * i);
        cp_async16(d + arow + 4 * i, bs + 4 * i);
      }
      for (int i = lane; i < nlp / 4; i += 32) cp_async16(d + 2 * arow + 4 * i, xs + 4 * i);
    }
    cp_async_commit();
  };
  stage(0);
  for (int q = 0; q < CB_FPW; ++q) {
    const int t = blockIdx.x * CB_FR + q * CB_W + warp;
    if (t >= len) break;
    float* dst = a.gout + (long long)b * a.g_stride_b + (long long)t * a.g_stride_t;
    if (!feasible) {  // no accepting path: zero posterior
      if (!a.scatter_grad)
        for (int j = lane; j < nl; j += 32) dst[j] = 0.f;
      continue;
    }
    __syncwarp();       // the previous frame's reads of the buffer about to be refilled are done
    stage(q + 1);
    cp_async_wait<1>();
    __syncwarp();
    const float* at = wbase + (q & 1) * stg;
    const float* bt = at + arow;
    const float* xt = bt + arow;
    const double ct = (double)at[arow - 2] + (double)at[arow - 1];
    const double dt = (double)bt[arow - 2] + (double)bt[arow - 1];
    const float Kt = (float)(ct + dt - costn / kLn2);
    // scatter mode: the gradient entries this frame will touch are requested now, so that their DRAM
    // latency runs behind the posterior sums (first 256 labels; the labels of an utterance are distinct)
    float* out = a.scatter_grad ? a.scatter_grad + ((long long)t * a.B + b) * V : nullptr;
    float* ptr[8];
    float cur[8];
    if (out) {
#pragma unroll
      for (int u = 0; u < 8; ++u) {
        const int j = lane + 32 * u;
        ptr[u] = out + ((j < nl) ? g.lab_list[lo + j] : 0);
        cur[u] = 0.f;
        if (j < nl) cur[u] = __ldcg(ptr[u]);
      }
    }
    float ls = 0.f;
    for (int base = 0; base < nl; base += 32) {
      const int j = base + lane;
      int e0 = 0, e1 = 0;
      if (j < nl) {
        e0 = lab_ptr[j];
        e1 = lab_ptr[j + 1];
      }
      const bool heavy = (e1 - e0) > HEAVY;
      float acc = 0.f;
      if (j < nl && !heavy) {
        const float kx = fmaf(xt[j], kLog2e, Kt);
        for (int e = e0; e < e1; ++e) {
          const unsigned pk = lab_pk[e];
          acc += post_term(at[pk & 0xffffu], bt[pk >> 16], UNIT ? kx : kx + lab_w[e]);
        }
      }
      unsigned hm = __ballot_sync(0xffffffffu, heavy);
      while (hm) {  // labels with many arcs (the blank): the whole warp strides over the arcs
        const int src_lane = __ffs(hm) - 1;
        hm &= hm - 1;
        const int jh = base + src_lane;
        const int h0 = __shfl_sync(0xffffffffu, e0, src_lane), h1 = __shfl_sync(0xffffffffu, e1, src_lane);
        const float kx = fmaf(xt[jh], kLog2e, Kt);
        float part = 0.f;
        for (int e = h0 + lane; e < h1; e += 32) {
          const unsigned pk = lab_pk[e];
          part += post_term(at[pk & 0xffffu], bt[pk >> 16], UNIT ? kx : kx + lab_w[e]);
        }
        part = warp_sum(part);
        if (lane == src_lane) acc = part;
      }
      if (j < nl) gt[j] = acc;
      ls += acc;
    }
    const float tot = warp_sum(ls);
    const float sc = tot > 0.f ? gmul / tot : 0.f;
    __syncwarp();
    if (a.scatter_grad) {
#pragma unroll
      for (int u = 0; u < 8; ++u) {
        const int j = lane + 32 * u;
        if (j < nl) *ptr[u] = cur[u] - gt[j] * sc;
      }
      for (int j0 = 256; j0 < nl; j0 += 256) {  // more than 256 distinct labels: batches of 8 loads, then 8 stores
        float* p2[8];
        float c2[8];
#pragma unroll
        for (int u = 0; u < 8; ++u) {
          const int j = j0 + lane + 32 * u;
          p2[u] = out + ((j < nl) ? g.lab_list[lo + j] : 0);
          if (j < nl) c2[u] = *p2[u];
        }
#pragma unroll
        for (int u = 0; u < 8; ++u) {
          const int j = j0 + lane + 32 * u;
          if (j < nl) *p2[u] = c2[u] - gt[j] * sc;
        }
      }
    } else {
      for (int j = lane; j < nl; j += 32) dst[j] = gt[j] * sc;
    }
  }
}

}  // namespace

template <bool UNIT, int SPT, int DK>
static cudaError_t launch_sweep_dk(const PathArgs& a, int bytes, cudaStream_t st) {
  cudaError_t e = cudaFuncSetAttribute(num_sweep_kernel<UNIT, SPT, DK>, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
  if (e != cudaSuccess) return e;
  num_sweep_kernel<UNIT, SPT, DK><<<dim3(a.B, a.write_grad ? 2 : 1), SW_TH, bytes, st>>>(a, a.g.max_states);
  return cudaGetLastError();
}
template <bool UNIT, int SPT>
static cudaError_t launch_sweep(const PathArgs& a, int bytes, cudaStream_t st) {
  if (SPT > 0 && a.g.max_in_degree <= 3 && a.g.max_out_degree <= 3) return launch_sweep_dk<UNIT, SPT, 3>(a, bytes, st);
  return launch_sweep_dk<UNIT, SPT, DEG>(a, bytes, st);
}

bool num_split_supported(const PathArgs& a) {
  if (!a.beta_ws) return false;
  if (a.g.max_states > 65535 || a.g.max_labels > 65535 || a.g.max_arcs < 1) return false;
  if (a.g.max_in_degree > DEG || a.g.max_out_degree > DEG) return false;
  if (sweep_plan(a.g.max_states, (int)a.x_stride_t, a.g.unit_weights != 0).bytes > 200 * 1024) return false;
  const int cb = ((a.g.max_labels + 1) * 4 + 16) + (a.g.max_arcs * 4 + 16) * 2 +
                 CB_W * (2 * (2 * a.alpha_row + (int)a.x_stride_t) + (int)a.x_stride_t) * 4 + 64;
  return cb <= 200 * 1024;
}

cudaError_t launch_num_sweeps(const PathArgs& a, cudaStream_t st) {
  const bool unit = a.g.unit_weights != 0;
  cudaError_t e;
  ProfScope prof("num_sweep", st);
  const SweepPlan P = sweep_plan(a.g.max_states, (int)a.x_stride_t, unit);
  const int spt = (a.g.max_states + NS - 1) / NS;
  if (spt <= 1) e = unit ? launch_sweep<true, 1>(a, P.bytes, st) : launch_sweep<false, 1>(a, P.bytes, st);
  else if (spt == 2) e = unit ? launch_sweep<true, 2>(a, P.bytes, st) : launch_sweep<false, 2>(a, P.bytes, st);
  else e = unit ? launch_sweep<true, 0>(a, P.bytes, st) : launch_sweep<false, 0>(a, P.bytes, st);
  ++g_launches;
  return e;
}

cudaError_t launch_num_combine(const PathArgs& a, cudaStream_t st) {
  if (!a.write_grad) return cudaSuccess;
  const bool unit = a.g.unit_weights != 0;
  cudaError_t e;
  ProfScope prof("num_combine", st);
  const int bytes = ((a.g.max_labels + 1) * 4 + 16) + (a.g.max_arcs * 4 + 16) * (unit ? 1 : 2) +
                    CB_W * (2 * (2 * a.alpha_row + (int)a.x_stride_t) + (int)a.x_stride_t) * 4 + 64;
  const dim3 grid((a.T + CB_FR - 1) / CB_FR, a.B);
  if (unit) {
    e = cudaFuncSetAttribute(num_combine_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
    if (e != cudaSuccess) return e;
    num_combine_kernel<true><<<grid, CB_TH, bytes, st>>>(a, a.g.max_arcs, a.g.max_labels);
  } else {
    e = cudaFuncSetAttribute(num_combine_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
    if (e != cudaSuccess) return e;
    num_combine_kernel<false><<<grid, CB_TH, bytes, st>>>(a, a.g.max_arcs, a.g.max_labels);
  }
  ++g_launches;
  return cudaGetLastError();
}

}  // namespace ctcg
