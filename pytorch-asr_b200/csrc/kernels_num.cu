// Numerator (per-utterance training graph) forward-backward: one CTA per utterance
// (8 recursion warps + 2 helper warps), log2 domain, sm_100a.
//
// Reference semantics: PathLogSumExp.forward/backward (att_speech/fst_utils.py:400-488)
// on the graphs of BaseGraphGen.get_training_matrices (fst_utils.py:647-660): N = 2U+1
// states with at most 3 arcs into / out of a state.
//
// Why the log domain here and plain scaled numbers in the denominator kernel: a
// training graph is a chain, so alpha_t of the states that carry the posterior mass can
// sit hundreds of binary orders below max_s alpha_t[s] (measured: below 2^-126 at T=400
// on random logits) -- a per-frame rescale cannot keep alpha * beta representable.
// In the decoding graph every state reaches every other within two frames, which
// bounds that ratio by the frame's logit spread.
//
// The kernel is latency bound (one short dependent chain per frame, 2T frames), so the
// CTA is cut into roles that each keep their chain short, one barrier per frame:
//   state warps (8)  one state per thread: 4-way log-sum-exp over the packed arcs, then
//                    floor(max) of the new values through REDUX + one shared atomicMax --
//                    an INTEGER normaliser, so subtracting it next frame is exact;
//   label warps (4)  backward only: posterior of one label per thread, arcs grouped by
//                    label on the host (fixed order, no atomics);
//   heavy warps (4)  backward only: labels with many arcs (the blank), one arc per thread;
//   io warp          cp.async ring, the fp64 running offsets (c, d) and the per-frame
//                    scalar 2^Kt the posteriors are scaled by, one frame ahead;
//   row warp         sums the previous frame's posterior row, normalises it to sum one
//                    and writes it out.
//
// Input: the compact label columns xg[b][t][j] = logits[t][b][lab_list_b[j]] (natural log;
// multiplied by log2(e) inside an fma so that the input is rounded once) written by
// gather_kernel; output: cost[b] (natural log) and the compact posterior rows
// gout[b][t][j], which the row / scatter kernels fold into the dense gradient.
#include <algorithm>
#include <cstdlib>

#include "device_utils.cuh"
#include "kernels.cuh"

namespace ctcg {
namespace {

constexpr int NS = 256;                 // state threads
constexpr int NLB = 128;                // label threads
constexpr int NHV = 128;                // heavy-label threads
constexpr int W_LAB = NS / 32;          // first label warp
constexpr int W_HV = (NS + NLB) / 32;   // first heavy warp
constexpr int W_IO = (NS + NLB + NHV) / 32;
constexpr int W_ROW = W_IO + 1;
constexpr int NTH = (W_ROW + 1) * 32;
constexpr int NWARPS = NTH / 32;
constexpr int NFWD = NS + 32;  // threads that take part in the forward sweep (state warps + io)
constexpr int XR = 8;          // cp.async ring depth (frames)
constexpr int DEG = 4;         // arcs packed per state
constexpr int HEAVY = 12;      // labels with more arcs go to the heavy warps
constexpr int NHMAX = 4;       // at most this many labels take that route; the rest loop
constexpr unsigned NOARC = 0xffffffffu;
constexpr int IMIN = -2147483647 - 1;

struct NumPlan {
  int in_pk, out_pk, in_w, out_w, lab_ptr, lab_pk, lab_w, a0, a1, fin, gr, hl, hx, hp, red, xring, aring;
  int bytes;
};

__host__ __device__ inline NumPlan num_plan(int maxN, int maxA, int maxL, int nlp, int arow, bool unit) {
  NumPlan p;
  int o = 0;
  auto take = [&](int bytes) {
    const int at = o;
    o += (bytes + 15) & ~15;
    return at;
  };
  p.in_pk = take(maxN * DEG * 4);
  p.out_pk = take(maxN * DEG * 4);
  p.in_w = unit ? 0 : take(maxN * DEG * 4);
  p.out_w = unit ? 0 : take(maxN * DEG * 4);
  p.lab_ptr = take((maxL + 1) * 4);
  p.lab_pk = take(maxA * 4);
  p.lab_w = unit ? 0 : take(maxA * 4);
  p.a0 = take((maxN + 1) * 4);  // + one slot that stays -inf: the far end of an absent arc
  p.a1 = take((maxN + 1) * 4);
  p.fin = take(maxN * 4);
  p.gr = take(2 * maxL * 4);
  p.hl = take(NHMAX * 4);
  p.hx = take(maxL * 4);
  p.hp = take(2 * NHMAX * NHV * 4);
  p.red = take((16 + 2 * NWARPS) * 4);
  p.xring = take(XR * nlp * 4);
  p.aring = take(XR * arow * 4);
  p.bytes = o;
  return p;
}

// log2(sum_k 2^v[k]) over DEG terms (-inf = absent), branch free
__device__ __forceinline__ float lse4(const float (&v)[DEG]) {
  const float M = fmaxf(fmaxf(v[0], v[1]), fmaxf(v[2], v[3]));
  const float Ms = (M > neg_inf()) ? M : 0.f;
  const float s = (ex2(v[0] - Ms) + ex2(v[1] - Ms)) + (ex2(v[2] - Ms) + ex2(v[3] - Ms));
  return Ms + lg2(s);  // lg2(0) = -inf
}

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

__device__ __forceinline__ void bar_all() { asm volatile("bar.sync 0;" ::: "memory"); }
__device__ __forceinline__ void bar_fwd() { asm volatile("bar.sync 1, %0;" ::"n"(NFWD) : "memory"); }

// floor(max over the warp) as an integer; -inf saturates to IMIN
__device__ __forceinline__ int warp_floor_max(float v) {
  return __reduce_max_sync(0xffffffffu, __float2int_rd(v));
}
__device__ __forceinline__ float norm_of(int m) { return (m == IMIN) ? 0.f : (float)m; }
__device__ __forceinline__ int next3(int r) { return r == 2 ? 0 : r + 1; }

// SPT > 0: every state thread owns states tid, tid + NS, ... (at most SPT) and every label thread
// labels j0, j0 + NLB, ... (at most SPT) for the whole utterance and keeps their arcs in
// registers as shared-memory byte offsets (no per-frame unpacking); SPT = 0: generic loops.
constexpr int LARC = 4;  // arcs of a label kept in registers; a label with more takes the loop

__device__ __forceinline__ float lds_f(unsigned a) {
  float v;
  asm volatile("ld.shared.f32 %0, [%1];" : "=f"(v) : "r"(a));
  return v;
}
__device__ __forceinline__ void sts_f(unsigned a, float v) { asm volatile("st.shared.f32 [%0], %1;" ::"r"(a), "f"(v)); }
__device__ __forceinline__ unsigned sm_u32(const void* p) { return static_cast<unsigned>(__cvta_generic_to_shared(p)); }

template <bool UNIT, int SPT>
__global__ void __launch_bounds__(NTH) num_log_kernel(const PathArgs a, const int maxN, const int maxA, const int maxL) {
  extern __shared__ __align__(16) unsigned char sm[];
  const GraphDev& g = a.g;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int b = blockIdx.x;
  const int gi = (g.num_graphs == 1) ? 0 : b;
  const int so = g.state_off[gi];
  const int n = g.state_off[gi + 1] - so;
  const int ro = so + gi;
  const int lo = g.lab_off[gi];
  const int nl = g.lab_off[gi + 1] - lo;
  const int len = min(a.lens[b], a.T);
  const int nlp = (int)a.x_stride_t, arow = a.alpha_row;
  const NumPlan P = num_plan(maxN, maxA, maxL, nlp, arow, UNIT);
  uint4* in_pk = reinterpret_cast<uint4*>(sm + P.in_pk);
  uint4* out_pk = reinterpret_cast<uint4*>(sm + P.out_pk);
  float4* in_w = reinterpret_cast<float4*>(sm + P.in_w);
  float4* out_w = reinterpret_cast<float4*>(sm + P.out_w);
  int* lab_ptr = reinterpret_cast<int*>(sm + P.lab_ptr);
  unsigned* lab_pk = reinterpret_cast<unsigned*>(sm + P.lab_pk);
  float* lab_w = reinterpret_cast<float*>(sm + P.lab_w);
  float* abuf0 = reinterpret_cast<float*>(sm + P.a0);
  float* abuf1 = reinterpret_cast<float*>(sm + P.a1);
  float* fin = reinterpret_cast<float*>(sm + P.fin);
  float* gr = reinterpret_cast<float*>(sm + P.gr);
  int* hl = reinterpret_cast<int*>(sm + P.hl);
  int* hx = reinterpret_cast<int*>(sm + P.hx);
  float* hp = reinterpret_cast<float*>(sm + P.hp);
  int* wmax = reinterpret_cast<int*>(sm + P.red);  // [3] floor(max) of the frame, rotating slots
  int* nheavy = wmax + 3;
  float* kt = reinterpret_cast<float*>(wmax + 4);  // [2] posterior scale exponent of the frame
  float* scr = reinterpret_cast<float*>(wmax + 16);  // [2 * NWARPS] scratch
  float* xring = reinterpret_cast<float*>(sm + P.xring);
  float* aring = reinterpret_cast<float*>(sm + P.aring);

  // ---- graph -> shared memory: DEG arcs per state in one 16-byte word ----
  const int arc0 = g.in_ptr[ro];
  const int narcs = g.in_ptr[ro + n] - arc0;
  for (int s = tid; s < n; s += NTH) {
    unsigned pi[DEG], po[DEG];
    float wi[DEG], wo[DEG];
#pragma unroll
    for (int k = 0; k < DEG; ++k) {
      const int ei = g.in_ptr[ro + s] + k, eo = g.out_ptr[ro + s] + k;
      const bool vi = ei < g.in_ptr[ro + s + 1], vo = eo < g.out_ptr[ro + s + 1];
      pi[k] = vi ? ((unsigned)g.in_src[ei] | ((unsigned)g.in_clab[ei] << 16)) : NOARC;
      po[k] = vo ? ((unsigned)g.out_dst[eo] | ((unsigned)g.out_clab[eo] << 16)) : NOARC;
      wi[k] = (vi && !UNIT) ? g.in_w[ei] * kLog2e : 0.f;
      wo[k] = (vo && !UNIT) ? g.out_w[eo] * kLog2e : 0.f;
    }
    in_pk[s] = make_uint4(pi[0], pi[1], pi[2], pi[3]);
    out_pk[s] = make_uint4(po[0], po[1], po[2], po[3]);
    if (!UNIT) {
      in_w[s] = make_float4(wi[0], wi[1], wi[2], wi[3]);
      out_w[s] = make_float4(wo[0], wo[1], wo[2], wo[3]);
    }
    const float f = g.final_w[so + s];
    fin[s] = (f > kDeadWeight) ? f * kLog2e : neg_inf();
    abuf0[s] = (s == 0) ? 0.f : neg_inf();
  }
  for (int j = tid; j <= nl; j += NTH) lab_ptr[j] = g.lab_ptr[lo + gi + j] - arc0;
  for (int e = tid; e < narcs; e += NTH) {
    lab_pk[e] = (unsigned)g.lab_src[arc0 + e] | ((unsigned)g.lab_dst[arc0 + e] << 16);
    if (!UNIT) lab_w[e] = g.lab_w[arc0 + e] * kLog2e;
  }
  if (tid < 3) wmax[tid] = (tid == 0) ? 0 : IMIN;  // max alpha_0 = 0
  if (tid == 0) abuf0[maxN] = abuf1[maxN] = neg_inf();
  const float* xb = a.x + (long long)b * a.x_stride_b;
  float* aw = a.alpha_ws + (long long)b * a.T * arow;
  const int xchunks = nlp / 4, achunks = arow / 4;
  bar_all();
  if (warp == W_ROW) {  // labels with many arcs (the blank), ascending
    int cnt = 0;
    for (int base = 0; base < nl; base += 32) {
      const int j = base + lane;
      const bool hv = j < nl && (lab_ptr[j + 1] - lab_ptr[j]) > HEAVY;
      const unsigned m = __ballot_sync(0xffffffffu, hv);
      const int idx = cnt + __popc(m & ((1u << lane) - 1u));
      if (j < nl) hx[j] = (hv && idx < NHMAX) ? idx : -1;
      if (hv && idx < NHMAX) hl[idx] = j;
      cnt += __popc(m);
    }
    if (lane == 0) *nheavy = min(cnt, NHMAX);
  }

  // =============================== forward ===============================
  float* a_cur = abuf0;
  float* a_nxt = abuf1;
  double c = 0.0;  // io warp: alpha_t = alpha_hat_t + c (log2 units)
  if (SPT > 0 && warp < W_LAB) {
    // arcs into the thread's states as byte offsets: .x source state in the alpha buffer
    // (an absent arc points at the -inf slot), .y label in the logits row
    constexpr int NSP = SPT > 0 ? SPT : 1;
    unsigned so4[NSP][DEG], lo4[NSP][DEG];
    float wi[NSP][DEG];
    float* arp[NSP];
#pragma unroll
    for (int i = 0; i < NSP; ++i) {
      const int s = tid + i * NS;
      const uint4 pk = (s < n) ? in_pk[s] : make_uint4(NOARC, NOARC, NOARC, NOARC);
      const unsigned pkk[DEG] = {pk.x, pk.y, pk.z, pk.w};
      const float4 w = (!UNIT && s < n) ? in_w[s] : make_float4(0.f, 0.f, 0.f, 0.f);
      const float ww[DEG] = {w.x, w.y, w.z, w.w};
#pragma unroll
      for (int k = 0; k < DEG; ++k) {
        so4[i][k] = (pkk[k] == NOARC) ? (unsigned)maxN * 4u : (pkk[k] & 0xffffu) * 4u;
        lo4[i][k] = (pkk[k] == NOARC) ? 0u : (pkk[k] >> 16) * 4u;
        wi[i][k] = ww[k];
      }
      arp[i] = aw + (s < n ? s : 0);
    }
    unsigned ac = sm_u32(a_cur), an = sm_u32(a_nxt);
    const unsigned xr0 = sm_u32(xring), wm_a = sm_u32(wmax);
    int r = 0, slot = 0;
    for (int t = 0; t < len; ++t) {
      bar_fwd();
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
        float v[DEG];
#pragma unroll
        for (int k = 0; k < DEG; ++k) {
          v[k] = fmaf(lds_f(xa + lo4[i][k]), kLog2e, lds_f(ac + so4[i][k]) - mx);
          if (!UNIT) v[k] += wi[i][k];
        }
        const float res = lse4(v);
        if (s < n) {
          *arp[i] = lds_f(ac + s * 4) - mx;
          sts_f(an + s * 4, res);
        }
        arp[i] += arow;
        lm = fmaxf(lm, res);
      }
      const int wm = warp_floor_max(lm);
      if (lane == 0) atomicMax(&wmax[r], wm);
      const unsigned tmp = ac;
      ac = an;
      an = tmp;
    }
    if (len & 1) {
      float* tmp = a_cur;
      a_cur = a_nxt;
      a_nxt = tmp;
    }
  } else if (warp < W_LAB) {
    int r = 0;
    for (int t = 0; t < len; ++t) {
      bar_fwd();
      const float mx = norm_of(wmax[r]);
      r = next3(r);
      const float* xrow = xring + (t % XR) * nlp;
      float* ar = aw + (long long)t * arow;
      float lm = neg_inf();
      for (int s = tid; s < n; s += NS) {
        ar[s] = a_cur[s] - mx;
        const uint4 pk = in_pk[s];
        const unsigned pkk[DEG] = {pk.x, pk.y, pk.z, pk.w};
        float v[DEG];
#pragma unroll
        for (int k = 0; k < DEG; ++k) {
          const unsigned q = (pkk[k] == NOARC) ? 0u : pkk[k];
          const float val = fmaf(xrow[q >> 16], kLog2e, a_cur[q & 0xffffu] - mx);
          v[k] = (pkk[k] == NOARC) ? neg_inf() : val;
        }
        if (!UNIT) {
          const float4 w = in_w[s];
          v[0] += w.x;
          v[1] += w.y;
          v[2] += w.z;
          v[3] += w.w;
        }
        const float res = lse4(v);
        a_nxt[s] = res;
        lm = fmaxf(lm, res);
      }
      const int wm = warp_floor_max(lm);
      if (lane == 0) atomicMax(&wmax[r], wm);
      float* tmp = a_cur;
      a_cur = a_nxt;
      a_nxt = tmp;
    }
  } else {
    if (warp == W_IO) {
      auto issue_x = [&](int t, int slot) {
        if (t >= 0 && t < len) {
          const float* src = xb + (long long)t * nlp;
          float* dst = xring + slot * nlp;
          for (int cc = lane; cc < xchunks; cc += 32) cp_async16(dst + 4 * cc, src + 4 * cc);
        }
        cp_async_commit();
      };
      for (int i = 0; i < XR - 1; ++i) issue_x(i, i);
      cp_async_wait<XR - 2>();
      int r = 0;
      for (int t = 0; t < len; ++t) {
        bar_fwd();
        if (!(a.dbg & 64)) issue_x(t + XR - 1, (t + XR - 1) % XR);
        c += (double)norm_of(wmax[r]);
        r = next3(r);
        if (lane == 0) {
          float* ar = aw + (long long)t * arow;
          const float hi = (float)c;
          ar[arow - 2] = hi;
          ar[arow - 1] = (float)(c - (double)hi);
          wmax[next3(r)] = IMIN;  // the slot the state warps fill during the next frame
        }
        if (!(a.dbg & 64)) cp_async_wait<XR - 2>();
      }
      cp_async_wait<0>();
    }
    if (len & 1) {
      float* tmp = a_cur;
      a_cur = a_nxt;
      a_nxt = tmp;
    }
  }
  bar_all();
  // terminal: cost = c_len + LSE_n(alpha_hat_len[n] + final[n])
  float tm = neg_inf();
  for (int s = tid; s < n; s += NTH) tm = fmaxf(tm, a_cur[s] + fin[s]);
  tm = warp_max(tm);
  if (lane == 0) scr[warp] = tm;
  bar_all();
  float TM = scr[0];
#pragma unroll
  for (int w = 1; w < NWARPS; ++w) TM = fmaxf(TM, scr[w]);
  float ts = 0.f;
  if (TM > neg_inf())
    for (int s = tid; s < n; s += NTH) ts += ex2(a_cur[s] + fin[s] - TM);
  ts = warp_sum(ts);
  if (lane == 0) scr[NWARPS + warp] = ts;
  bar_all();
  float TS = 0.f;
#pragma unroll
  for (int w = 0; w < NWARPS; ++w) TS += scr[NWARPS + w];
  const bool feasible = TM > neg_inf() && TS > 0.f;
  const double cost2 = feasible ? (c + (double)TM + (double)lg2(TS)) : 0.0;  // meaningful in the io warp only
  if (warp == W_IO && lane == 0) a.cost[b] = feasible ? cost2 * kLn2 : -1e20;
  if (!a.write_grad) return;
  float* gb = a.gout + (long long)b * a.g_stride_b;
  if (!feasible) {
    for (int t = 0; t < len; ++t)
      for (int j = tid; j < nl; j += NTH) gb[(long long)t * a.g_stride_t + j] = 0.f;
    return;
  }

  // =============================== backward ==============================
  const float gmul = a.g_sign * (a.g_scale ? a.g_scale[b] : 1.f);
  float* b_cur = a_nxt;  // the buffer not holding alpha_len
  float* b_nxt = a_cur;
  for (int s = tid; s < n; s += NTH) b_cur[s] = fin[s];
  if (tid < 3) wmax[tid] = (tid == 0) ? 0 : IMIN;  // the first frame is not normalised
  const int nh = *nheavy;
  if (SPT > 0 && warp < W_LAB) {
    // ---- state warps, arcs out of the thread's states in registers ----
    constexpr int NSP = SPT > 0 ? SPT : 1;
    unsigned do4[NSP][DEG], lo4[NSP][DEG];
    float wo[NSP][DEG];
#pragma unroll
    for (int i = 0; i < NSP; ++i) {
      const int s = tid + i * NS;
      const uint4 pk = (s < n) ? out_pk[s] : make_uint4(NOARC, NOARC, NOARC, NOARC);
      const unsigned pkk[DEG] = {pk.x, pk.y, pk.z, pk.w};
      const float4 w = (!UNIT && s < n) ? out_w[s] : make_float4(0.f, 0.f, 0.f, 0.f);
      const float ww[DEG] = {w.x, w.y, w.z, w.w};
#pragma unroll
      for (int k = 0; k < DEG; ++k) {
        do4[i][k] = (pkk[k] == NOARC) ? (unsigned)maxN * 4u : (pkk[k] & 0xffffu) * 4u;
        lo4[i][k] = (pkk[k] == NOARC) ? 0u : (pkk[k] >> 16) * 4u;
        wo[i][k] = ww[k];
      }
    }
    unsigned bc = sm_u32(b_cur), bn = sm_u32(b_nxt);
    const unsigned xr0 = sm_u32(xring), wm_a = sm_u32(wmax);
    bar_all();
    int r = 0, slot = 0;
    for (int i = 0; i < len; ++i) {
      bar_all();
      int wmi;
      asm volatile("ld.shared.s32 %0, [%1];" : "=r"(wmi) : "r"(wm_a + r * 4));
      const float mxb = norm_of(wmi);
      r = next3(r);
      const unsigned xa = xr0 + slot * nlp * 4;
      slot = (slot + 1 == XR) ? 0 : slot + 1;
      float lm = neg_inf();
      if (a.dbg & 8) continue;
#pragma unroll
      for (int q = 0; q < NSP; ++q) {
        const int s = tid + q * NS;
        float v[DEG];
#pragma unroll
        for (int k = 0; k < DEG; ++k) {
          v[k] = fmaf(lds_f(xa + lo4[q][k]), kLog2e, lds_f(bc + do4[q][k]) - mxb);
          if (!UNIT) v[k] += wo[q][k];
        }
        const float res = lse4(v);
        if (s < n) sts_f(bn + s * 4, res);
        lm = fmaxf(lm, res);
      }
      const int wm = warp_floor_max(lm);
      if (lane == 0) atomicMax(&wmax[r], wm);
      const unsigned tmp = bc;
      bc = bn;
      bn = tmp;
    }
    bar_all();
  } else if (warp < W_LAB) {
    // ---- state warps: beta_t[s] = LSE_{arcs out of s}(w + x[label] + beta_{t+1}[dst]) ----
    bar_all();
    int r = 0;
    for (int i = 0; i < len; ++i) {
      bar_all();
      const float mxb = norm_of(wmax[r]);
      r = next3(r);
      const float* xrow = xring + (i % XR) * nlp;
      float lm = neg_inf();
      for (int s = tid; s < n; s += NS) {
        const uint4 pk = out_pk[s];
        const unsigned pkk[DEG] = {pk.x, pk.y, pk.z, pk.w};
        float v[DEG];
#pragma unroll
        for (int k = 0; k < DEG; ++k) {
          const unsigned q = (pkk[k] == NOARC) ? 0u : pkk[k];
          const float val = fmaf(xrow[q >> 16], kLog2e, b_cur[q & 0xffffu] - mxb);
          v[k] = (pkk[k] == NOARC) ? neg_inf() : val;
        }
        if (!UNIT) {
          const float4 w = out_w[s];
          v[0] += w.x;
          v[1] += w.y;
          v[2] += w.z;
          v[3] += w.w;
        }
        const float res = lse4(v);
        b_nxt[s] = res;
        lm = fmaxf(lm, res);
      }
      const int wm = warp_floor_max(lm);
      if (lane == 0) atomicMax(&wmax[r], wm);
      float* tmp = b_cur;
      b_cur = b_nxt;
      b_nxt = tmp;
    }
    bar_all();
  } else if (SPT > 0 && warp < W_HV) {
    // ---- label warps, the (at most LARC) arcs of the thread's labels in registers; an absent arc
    // ends in the -inf slot of the beta buffer.  Labels with more arcs take the loop. ----
    constexpr int NSP = SPT > 0 ? SPT : 1;
    const int j0 = tid - NS;
    unsigned sa[NSP][LARC], da[NSP][LARC];
    float wl[NSP][LARC];
    int kind[NSP];  // 0: nothing to do (no such label / heavy), 1: registers, 2: loop
#pragma unroll
    for (int q = 0; q < NSP; ++q) {
      const int j = j0 + q * NLB;
      kind[q] = 0;
#pragma unroll
      for (int k = 0; k < LARC; ++k) {
        sa[q][k] = 0u;
        da[q][k] = (unsigned)maxN * 4u;
        wl[q][k] = 0.f;
      }
      if (j < nl && hx[j] < 0) {
        const int e0 = lab_ptr[j], e1 = lab_ptr[j + 1];
        kind[q] = (e1 - e0 <= LARC) ? 1 : 2;
        if (kind[q] == 1) {
#pragma unroll
          for (int k = 0; k < LARC; ++k) {
            if (e0 + k < e1) {
              const unsigned pk = lab_pk[e0 + k];
              sa[q][k] = (pk & 0xffffu) * 4u;
              da[q][k] = (pk >> 16) * 4u;
              if (!UNIT) wl[q][k] = lab_w[e0 + k];
            }
          }
        }
      }
    }
    unsigned bc = sm_u32(b_cur), bn = sm_u32(b_nxt);
    const unsigned xr0 = sm_u32(xring), ar0 = sm_u32(aring), kt_a = sm_u32(kt), gr_a = sm_u32(gr);
    bar_all();
    int slot = 0;
    for (int i = 0; i < len; ++i) {
      bar_all();
      const float Kt = lds_f(kt_a + (i & 1) * 4);
      const unsigned xa = xr0 + slot * nlp * 4, ata = ar0 + slot * arow * 4;
      slot = (slot + 1 == XR) ? 0 : slot + 1;
      const unsigned grw = gr_a + (i & 1) * maxL * 4;
      if (a.dbg & 2) continue;
#pragma unroll
      for (int q = 0; q < NSP; ++q) {
        const int j = j0 + q * NLB;
        if (j < nl) {
          float acc = 0.f;
          if (kind[q] == 1) {
            const float kx = fmaf(lds_f(xa + j * 4), kLog2e, Kt);
#pragma unroll
            for (int k = 0; k < LARC; ++k)
              acc += post_term(lds_f(ata + sa[q][k]), lds_f(bc + da[q][k]), UNIT ? kx : kx + wl[q][k]);
          } else if (kind[q] == 2) {
            const float* at = aring + (i % XR) * arow;
            const float* bcur = (i & 1) ? b_nxt : b_cur;
            const float kx = fmaf(lds_f(xa + j * 4), kLog2e, Kt);
            for (int e = lab_ptr[j]; e < lab_ptr[j + 1]; ++e) {
              const unsigned pk = lab_pk[e];
              acc += post_term(at[pk & 0xffffu], bcur[pk >> 16], UNIT ? kx : kx + lab_w[e]);
            }
          }
          sts_f(grw + j * 4, acc);  // zero for a heavy label: the row warp fills it in
        }
      }
      const unsigned tmp = bc;
      bc = bn;
      bn = tmp;
    }
    bar_all();
  } else if (warp < W_HV) {
    // ---- label warps: gamma[j] = sum_{arcs labelled j} 2^(x[j] + w + alpha_t[src] + beta_{t+1}[dst] - cost);
    // every term is a probability (<= 1), so no running maximum is needed ----
    const int j0 = tid - NS;
    bar_all();
    for (int i = 0; i < len; ++i) {
      bar_all();
      const float Kt = kt[i & 1];
      const float* xrow = xring + (i % XR) * nlp;
      const float* at = aring + (i % XR) * arow;
      float* grw = gr + (i & 1) * maxL;
      for (int j = j0; j < nl; j += NLB) {
        float acc = 0.f;
        if (hx[j] < 0) {
          const int e0 = lab_ptr[j], e1 = lab_ptr[j + 1];
          const float kx = fmaf(xrow[j], kLog2e, Kt);
          for (int e = e0; e < e1; ++e) {
            const unsigned pk = lab_pk[e];
            acc += post_term(at[pk & 0xffffu], b_cur[pk >> 16], UNIT ? kx : kx + lab_w[e]);
          }
        }
        grw[j] = acc;  // zero for a heavy label: the row warp fills it in
      }
      float* tmp = b_cur;
      b_cur = b_nxt;
      b_nxt = tmp;
    }
    bar_all();
  } else if (SPT > 0 && warp < W_IO) {
    // ---- heavy warps, the thread's (at most HARC) arcs of the first heavy label -- the blank --
    // in registers; further heavy labels, or a longer one, take the loop ----
    constexpr int HARC = 4;
    const int k0 = tid - NS - NLB;
    unsigned sa[HARC], da[HARC];
    float wl[HARC];
    int hj = 0;
    bool fast = false;
#pragma unroll
    for (int q = 0; q < HARC; ++q) {
      sa[q] = 0u;
      da[q] = (unsigned)maxN * 4u;
      wl[q] = 0.f;
    }
    if (nh > 0) {
      hj = hl[0];
      const int e0 = lab_ptr[hj], e1 = lab_ptr[hj + 1];
      fast = (e1 - e0) <= HARC * NHV;
      if (fast) {
#pragma unroll
        for (int q = 0; q < HARC; ++q) {
          const int e = e0 + k0 + q * NHV;
          if (e < e1) {
            const unsigned pk = lab_pk[e];
            sa[q] = (pk & 0xffffu) * 4u;
            da[q] = (pk >> 16) * 4u;
            if (!UNIT) wl[q] = lab_w[e];
          }
        }
      }
    }
    unsigned bc = sm_u32(b_cur), bn = sm_u32(b_nxt);
    const unsigned xr0 = sm_u32(xring), ar0 = sm_u32(aring), kt_a = sm_u32(kt), hp_a = sm_u32(hp);
    bar_all();
    int slot = 0;
    for (int i = 0; i < len; ++i) {
      bar_all();
      const float Kt = lds_f(kt_a + (i & 1) * 4);
      const unsigned xa = xr0 + slot * nlp * 4, ata = ar0 + slot * arow * 4;
      const float* xrow = xring + slot * nlp;
      const float* at = aring + slot * arow;
      const float* bcur = (i & 1) ? b_nxt : b_cur;
      slot = (slot + 1 == XR) ? 0 : slot + 1;
      const unsigned hpw = hp_a + (i & 1) * NHMAX * NHV * 4;
      if (a.dbg & 4) continue;
      if (fast) {
        const float kx = fmaf(lds_f(xa + hj * 4), kLog2e, Kt);
        float acc = 0.f;
#pragma unroll
        for (int q = 0; q < HARC; ++q)
          acc += post_term(lds_f(ata + sa[q]), lds_f(bc + da[q]), UNIT ? kx : kx + wl[q]);
        sts_f(hpw + k0 * 4, acc);
      }
      for (int h = fast ? 1 : 0; h < nh; ++h) {
        const int j = hl[h];
        const int e0 = lab_ptr[j], e1 = lab_ptr[j + 1];
        const float kx = fmaf(xrow[j], kLog2e, Kt);
        float acc = 0.f;
        for (int e = e0 + k0; e < e1; e += NHV) {
          const unsigned pk = lab_pk[e];
          acc += post_term(at[pk & 0xffffu], bcur[pk >> 16], UNIT ? kx : kx + lab_w[e]);
        }
        sts_f(hpw + (h * NHV + k0) * 4, acc);
      }
      const unsigned tmp = bc;
      bc = bn;
      bn = tmp;
    }
    bar_all();
  } else if (warp < W_IO) {
    // ---- heavy warps: one arc of a many-arc label per thread, partial sums to the row warp ----
    const int k0 = tid - NS - NLB;
    bar_all();
    for (int i = 0; i < len; ++i) {
      bar_all();
      const float Kt = kt[i & 1];
      const float* xrow = xring + (i % XR) * nlp;
      const float* at = aring + (i % XR) * arow;
      float* hpw = hp + (i & 1) * NHMAX * NHV;
      for (int h = 0; h < nh; ++h) {
        const int j = hl[h];
        const int e0 = lab_ptr[j], e1 = lab_ptr[j + 1];
        const float kx = fmaf(xrow[j], kLog2e, Kt);
        float acc = 0.f;
        for (int e = e0 + k0; e < e1; e += NHV) {
          const unsigned pk = lab_pk[e];
          acc += post_term(at[pk & 0xffffu], b_cur[pk >> 16], UNIT ? kx : kx + lab_w[e]);
        }
        hpw[h * NHV + k0] = acc;
      }
      float* tmp = b_cur;
      b_cur = b_nxt;
      b_nxt = tmp;
    }
    bar_all();
  } else if (warp == W_IO) {
    auto issue_b = [&](int i, int slot) {
      if (i >= 0 && i < len) {
        const int t = len - 1 - i;
        const float* src = xb + (long long)t * nlp;
        float* dst = xring + slot * nlp;
        for (int cc = lane; cc < xchunks; cc += 32) cp_async16(dst + 4 * cc, src + 4 * cc);
        const float* asrc = aw + (long long)t * arow;
        float* adst = aring + slot * arow;
        for (int cc = lane; cc < achunks; cc += 32) cp_async16(adst + 4 * cc, asrc + 4 * cc);
      }
      cp_async_commit();
    };
    // Kt of frame i = c_t + d_i - cost (fp64), d_i = sum of the beta normalisers before iteration i
    auto publish_kt = [&](int i, double d) {
      if (i < len) {
        const float* at = aring + (i % XR) * arow;
        const double ct = (double)at[arow - 2] + (double)at[arow - 1];
        if (lane == 0) kt[i & 1] = (float)(ct + d - cost2);
      }
    };
    for (int i = 0; i < XR - 1; ++i) issue_b(i, i);
    cp_async_wait<XR - 2>();
    __syncwarp();
    double d = 0.0;  // beta_{t+1} = beta_hat + d
    publish_kt(0, d);
    bar_all();
    int r = 0;
    for (int i = 0; i < len; ++i) {
      bar_all();
      if (!(a.dbg & 32)) issue_b(i + XR - 1, (i + XR - 1) % XR);
      d += (double)norm_of(wmax[r]);
      r = next3(r);
      if (lane == 0) wmax[next3(r)] = IMIN;
      if (!(a.dbg & 32)) cp_async_wait<XR - 2>();
      __syncwarp();
      if (!(a.dbg & 16)) publish_kt(i + 1, d);
    }
    cp_async_wait<0>();
    bar_all();
  } else {
    // ---- row warp: finishes the posterior row of iteration i - 1 while the others run iteration i ----
    auto finish = [&](int i) {
      float* grp = gr + (i & 1) * maxL;
      const float* hpp = hp + (i & 1) * NHMAX * NHV;
      float ls = 0.f;
      for (int j = lane; j < nl; j += 32) ls += grp[j];
      float v0 = 0.f;
      if (nh > 0) {
#pragma unroll
        for (int k = 0; k < NHV / 32; ++k) v0 += hpp[k * 32 + lane];
      }
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) {  // two independent chains, interleaved
        ls += __shfl_xor_sync(0xffffffffu, ls, o);
        v0 += __shfl_xor_sync(0xffffffffu, v0, o);
      }
      float tot = ls + v0;
      if (nh > 0 && lane == 0) grp[hl[0]] = v0;
      for (int h = 1; h < nh; ++h) {
        float v = 0.f;
#pragma unroll
        for (int k = 0; k < NHV / 32; ++k) v += hpp[h * NHV + k * 32 + lane];
        v = warp_sum(v);
        if (lane == 0) grp[hl[h]] = v;
        tot += v;  // identical in every lane
      }
      __syncwarp();
      const float sc = tot > 0.f ? gmul / tot : 0.f;
      float* dst = gb + (long long)(len - 1 - i) * a.g_stride_t;
      for (int j = lane; j < nl; j += 32) dst[j] = grp[j] * sc;
    };
    bar_all();
    for (int i = 0; i < len; ++i) {
      bar_all();
      if (i > 0 && !(a.dbg & 1)) finish(i - 1);
    }
    bar_all();
    if (len > 0) finish(len - 1);
  }
}


// =====================================================================================
// Split numerator: the alpha and the beta recursion run AT THE SAME TIME on separate CTAs
// (grid (B, 2)), each spilling its normalised rows; the posteriors are then formed by a
// time-parallel combine kernel.  Halves the serial chain of the fused kernel above (T frames
// instead of 2T) and takes the posterior work off it altogether.
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
        const float res = lse4(v);
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
__global__ void __launch_bounds__(CB_TH) num_combine_kernel(const PathArgs a, const int maxA, const int maxL) {
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
  float* rows_s = reinterpret_cast<float*>(sm + take(CB_W * (2 * arow + 2 * nlp) * 4));
  const int arc0 = g.in_ptr[ro];
  const int narcs = g.in_ptr[ro + n] - arc0;
  for (int j = tid; j <= nl; j += CB_TH) lab_ptr[j] = g.lab_ptr[lo + gi + j] - arc0;
  for (int e = tid; e < narcs; e += CB_TH) {
    lab_pk[e] = (unsigned)g.lab_src[arc0 + e] | ((unsigned)g.lab_dst[arc0 + e] << 16);
    if (!UNIT) lab_w[e] = g.lab_w[arc0 + e] * kLog2e;
  }
  __syncthreads();
  float* at = rows_s + warp * (2 * arow + 2 * nlp);
  float* bt = at + arow;
  float* xt = bt + arow;
  float* gt = xt + nlp;
  const double costn = a.cost[b];
  const bool feasible = costn > -1e19;
  const float gmul = a.g_sign * (a.g_scale ? a.g_scale[b] : 1.f);
  const int V = a.scatter_V;
  for (int q = 0; q < CB_FPW; ++q) {
    const int t = blockIdx.x * CB_FR + q * CB_W + warp;
    if (t >= len) break;
    float* dst = a.gout + (long long)b * a.g_stride_b + (long long)t * a.g_stride_t;
    if (!feasible) {  // no accepting path: zero posterior
      if (!a.scatter_grad)
        for (int j = lane; j < nl; j += 32) dst[j] = 0.f;
      continue;
    }
    __syncwarp();
    const float4* as = reinterpret_cast<const float4*>(a.alpha_ws + ((long long)b * a.T + t) * arow);
    const float4* bs = reinterpret_cast<const float4*>(a.beta_ws + ((long long)b * a.T + t) * arow);
    const float4* xs = reinterpret_cast<const float4*>(a.x + (long long)b * a.x_stride_b + (long long)t * nlp);
    for (int i = lane; i < arow / 4; i += 32) {
      reinterpret_cast<float4*>(at)[i] = as[i];
      reinterpret_cast<float4*>(bt)[i] = bs[i];
    }
    for (int i = lane; i < nlp / 4; i += 32) reinterpret_cast<float4*>(xt)[i] = xs[i];
    __syncwarp();
    const double ct = (double)at[arow - 2] + (double)at[arow - 1];
    const double dt = (double)bt[arow - 2] + (double)bt[arow - 1];
    const float Kt = (float)(ct + dt - costn / kLn2);
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
      // the labels of an utterance are distinct: all loads of a batch of 8 go out before the first store
      float* out = a.scatter_grad + ((long long)t * a.B + b) * V;
      for (int j0 = 0; j0 < nl; j0 += 256) {
        float* ptr[8];
        float cur[8];
#pragma unroll
        for (int u = 0; u < 8; ++u) {
          const int j = j0 + lane + 32 * u;
          ptr[u] = out + ((j < nl) ? g.lab_list[lo + j] : 0);
          if (j < nl) cur[u] = *ptr[u];
        }
#pragma unroll
        for (int u = 0; u < 8; ++u) {
          const int j = j0 + lane + 32 * u;
          if (j < nl) *ptr[u] = cur[u] - gt[j] * sc;
        }
      }
    } else {
      for (int j = lane; j < nl; j += 32) dst[j] = gt[j] * sc;
    }
  }
}

}  // namespace

bool num_lin_supported(const PathArgs& a) {
  if (a.g.max_states > 65535 || a.g.max_labels > 65535 || a.g.max_arcs < 1) return false;
  if (a.g.max_in_degree > DEG || a.g.max_out_degree > DEG) return false;
  const NumPlan P = num_plan(a.g.max_states, a.g.max_arcs, a.g.max_labels, (int)a.x_stride_t, a.alpha_row,
                             a.g.unit_weights != 0);
  return P.bytes <= 200 * 1024;
}

template <bool UNIT, int SPT>
static cudaError_t launch_one(const PathArgs& a, int bytes, cudaStream_t st) {
  cudaError_t e = cudaFuncSetAttribute(num_log_kernel<UNIT, SPT>, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
  if (e != cudaSuccess) return e;
  num_log_kernel<UNIT, SPT><<<a.B, NTH, bytes, st>>>(a, a.g.max_states, a.g.max_arcs, a.g.max_labels);
  return cudaGetLastError();
}

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
  if (!a.beta_ws || !num_lin_supported(a)) return false;
  const int cb = ((a.g.max_labels + 1) * 4 + 16) + (a.g.max_arcs * 4 + 16) * 2 +
                 CB_W * (2 * a.alpha_row + 2 * (int)a.x_stride_t) * 4 + 64;
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
                    CB_W * (2 * a.alpha_row + 2 * (int)a.x_stride_t) * 4 + 64;
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

cudaError_t launch_num_lin(const PathArgs& a_in, cudaStream_t st) {
  PathArgs a = a_in;
  static const int dbg = [] {
    const char* e = std::getenv("CTCG_NUM_DEBUG");
    return e ? std::atoi(e) : 0;
  }();
  a.dbg = dbg;
  const bool unit = a.g.unit_weights != 0;
  const NumPlan P = num_plan(a.g.max_states, a.g.max_arcs, a.g.max_labels, (int)a.x_stride_t, a.alpha_row, unit);
  ProfScope prof("num_log", st);
  // states / labels per thread: register-resident arcs up to 2 per thread, generic loops beyond
  const int spt = std::max((a.g.max_states + NS - 1) / NS, (a.g.max_labels + NLB - 1) / NLB);
  cudaError_t e;
  if (spt <= 1) e = unit ? launch_one<true, 1>(a, P.bytes, st) : launch_one<false, 1>(a, P.bytes, st);
  else if (spt == 2) e = unit ? launch_one<true, 2>(a, P.bytes, st) : launch_one<false, 2>(a, P.bytes, st);
  else e = unit ? launch_one<true, 0>(a, P.bytes, st) : launch_one<false, 0>(a, P.bytes, st);
  ++g_launches;
  return e;
}

}  // namespace ctcg
