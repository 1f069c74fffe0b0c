// Numerator (per-utterance training graph) forward-backward, sm_100a: one WARP per (utterance,
// direction) runs the alpha or the beta recursion (num_sweep_kernel), then a time-parallel kernel
// forms the posteriors (num_combine_kernel).
//
// Reference semantics: PathLogSumExp.forward/backward (att_speech/fst_utils.py:400-488) on the
// graphs of BaseGraphGen.get_training_matrices (fst_utils.py:647-660): N = 2U+1 states with at
// most 3 arcs into / out of a state.
//
// Number domain.  A training graph is a chain, so alpha_t of the states that carry the posterior
// mass sits hundreds of binary orders below max_s alpha_t[s]: numbers scaled per frame underflow
// (measured at T = 400) and a plain fp32 log value of that magnitude has an ulp of ~1.5e-5, which
// is where the round-1 kernel lost the gradient tolerance at T >= 800.  Every value is therefore a
// block-floating-point pair  v = f * 2^E  with an fp32 mantissa f in {0} U [1, 2) and an exact
// int32 exponent E per state: exponents add exactly, mantissas multiply with a relative error of
// 2^-24 whatever the magnitude, a sum aligns its terms to the largest exponent with a power of two
// built in the integer pipe.  No per-frame normaliser, nothing lagged, and no transcendental on
// the recursion at all: the only MUFU work is one ex2 per label and frame for
// e = 2^frac(x log2 e) (the integer part of x log2 e goes to the exponent).
//
// num_sweep_kernel   one 128-thread CTA per (utterance, direction): the vector of the current frame in
//                    shared memory (int2 per state, double buffered), the arcs of <= 4 states per thread
//                    in registers, a 16-frame cp.async ring of the label columns this utterance touches
//                    -- gathered straight from the (T, B, V) logits (the round-1 gather kernel is gone;
//                    the alpha CTA writes the compact rows the combine reads) -- and the (exponent,
//                    mantissa) form of the next frame's row.  One block barrier per frame, no atomics.
// num_combine_kernel one warp per frame: gamma_t[j] = sum_{arcs e labelled j}
//                    alpha_t[src] e_t[j] w_e beta_{t+1}[dst] / cost, fixed summation order (arcs grouped
//                    by label on the host), row normalised to one; with a factorised denominator the
//                    result is subtracted from the dense gradient rows directly.
#include <algorithm>
#include <cstdlib>
#include <type_traits>

#include "device_utils.cuh"
#include "kernels.cuh"

namespace ctcg {
namespace {

constexpr int NT = 128;         // threads of a sweep CTA
constexpr int XR = 16;          // frames of label columns in flight per sweep
constexpr int DEG = 4;          // arcs packed per state
constexpr int HEAVY = 12;       // labels with more arcs are summed by the whole warp
constexpr unsigned NOARC = 0xffffffffu;
constexpr int IDEAD = -(1 << 28);          // integer part of "no path"
constexpr float MAGIC = 12582912.f;        // 1.5 * 2^23: x + MAGIC has rint(x) in its low mantissa bits
constexpr int MAGIC_BITS = 0x4B400000;
constexpr float XCLAMP = 65536.f;          // |x log2 e| is clamped to 2^16: T * 2^16 stays above IDEAD for T <= 2048

__device__ __forceinline__ int2 lds64(unsigned a) {
  int2 v;
  asm volatile("ld.shared.v2.s32 {%0,%1}, [%2];" : "=r"(v.x), "=r"(v.y) : "r"(a));
  return v;
}
__device__ __forceinline__ void sts64(unsigned a, int2 v) {
  asm volatile("st.shared.v2.s32 [%0], {%1,%2};" ::"r"(a), "r"(v.x), "r"(v.y) : "memory");
}
__device__ __forceinline__ unsigned sm_u32(const void* p) { return static_cast<unsigned>(__cvta_generic_to_shared(p)); }
__device__ __forceinline__ void cp_async4(unsigned dst, const void* src) {
  asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(dst), "l"(src) : "memory");
}

// x (log2 units, clamped) -> (E, f): 2^x = f * 2^E with E = rint(x), f = 2^(x - E) in [0.707, 1.415]
__device__ __forceinline__ int2 split_pow2(float xs) {
  xs = fminf(fmaxf(xs, -XCLAMP), XCLAMP);
  const float t = xs + MAGIC;
  return make_int2(__float_as_int(t) - MAGIC_BITS, __float_as_int(ex2(xs - (t - MAGIC))));
}
// 2^d for d <= 127; anything at or below -127 is exactly 0
__device__ __forceinline__ float pow2c(int d) { return __int_as_float((max(d, -127) + 127) << 23); }
// v > 0 normal -> (floor(log2 v), mantissa in [1, 2))
__device__ __forceinline__ int2 renorm(float v, int base) {
  const int bits = __float_as_int(v);
  return make_int2(base + ((bits >> 23) - 127), (bits & 0x007fffff) | 0x3f800000);
}

struct SweepPlan {
  int pk, w, a0, a1, fin, lab, land, xt, bytes;  // byte offsets inside one warp's slice
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
  p.a0 = take((maxN + 1) * 8);
  p.a1 = take((maxN + 1) * 8);
  p.fin = take(maxN * 4);
  p.lab = take(nlp * 4);
  p.land = take(XR * nlp * 4);
  p.xt = take(2 * nlp * 8);
  p.bytes = o;
  return p;
}

// One state: sum over its DK arcs, terms aligned to the largest exponent.  An absent / dead arc has
// mantissa 0 and an exponent near IDEAD, far below any live one.
template <int DK>
__device__ __forceinline__ int2 sum_state(const int (&J)[DK], const float (&p)[DK]) {
  int Jm = J[0];
#pragma unroll
  for (int k = 1; k < DK; ++k) Jm = max(Jm, J[k]);
  float acc = 0.f;
#pragma unroll
  for (int k = 0; k < DK; ++k) acc = fmaf(p[k], pow2c(J[k] - Jm), acc);
  const int2 out = renorm(acc, Jm);
  return (acc > 0.f) ? out : make_int2(IDEAD, 0);
}

// One CTA of NT threads per sweep.  SPT > 0: the arcs of the SPT states of a thread (state = tid + NT i)
// live in registers as shared-memory byte offsets; SPT == 0: any number of states, arcs re-read from
// shared memory.  DK: arcs per state that exist anywhere in the batch (3 for CTC training graphs,
// fst_utils.py:592-605).  NLJ > 0: at most NT * NLJ distinct labels, their class ids in registers.
template <bool UNIT, int SPT, int DK, int NLJ>
__global__ void __launch_bounds__(NT) num_sweep_kernel(const PathArgs a, const int maxN) {
  extern __shared__ __align__(16) unsigned char sm[];
  __shared__ int red_i[NT / 32];
  __shared__ float red_f[NT / 32];
  const GraphDev& g = a.g;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const bool both = a.write_grad != 0;
  const int q = blockIdx.x;
  const int b = both ? (q >> 1) : q;
  const int DIR = both ? (q & 1) : 0;
  const int gi = (g.num_graphs == 1) ? 0 : b;
  const int so = g.state_off[gi];
  const int n = g.state_off[gi + 1] - so;
  const int ro = so + gi;
  const int lo = g.lab_off[gi];
  const int nl = g.lab_off[gi + 1] - lo;
  const int len = min(a.lens[b], a.T);
  const int nlp = (int)a.x_stride_t, arow = a.alpha_row;
  const SweepPlan P = sweep_plan(maxN, nlp, UNIT);
  uint4* pkv = reinterpret_cast<uint4*>(sm + P.pk);
  float4* wv = reinterpret_cast<float4*>(sm + P.w);
  int2* abuf0 = reinterpret_cast<int2*>(sm + P.a0);
  int2* abuf1 = reinterpret_cast<int2*>(sm + P.a1);
  float* fin = reinterpret_cast<float*>(sm + P.fin);
  int* labs = reinterpret_cast<int*>(sm + P.lab);
  float* land = reinterpret_cast<float*>(sm + P.land);
  int2* xt = reinterpret_cast<int2*>(sm + P.xt);

  const int32_t* ptr = DIR ? g.out_ptr : g.in_ptr;
  const int32_t* oth = DIR ? g.out_dst : g.in_src;
  const int32_t* clab = DIR ? g.out_clab : g.in_clab;
  const float* wgt = DIR ? g.out_w : g.in_w;
  for (int s = tid; s < n; s += NT) {
    unsigned pi[DEG];
    float wi[DEG];
    const int e0 = ptr[ro + s], e1 = ptr[ro + s + 1];
#pragma unroll
    for (int k = 0; k < DEG; ++k) {
      const int e = e0 + k;
      const bool v = e < e1;
      pi[k] = v ? ((unsigned)oth[e] | ((unsigned)clab[e] << 16)) : NOARC;
      wi[k] = (v && !UNIT) ? wgt[e] * kLog2e : 0.f;
    }
    pkv[s] = make_uint4(pi[0], pi[1], pi[2], pi[3]);
    if (!UNIT) wv[s] = make_float4(wi[0], wi[1], wi[2], wi[3]);
    const float f = g.final_w[so + s];
    fin[s] = (f > kDeadWeight) ? f * kLog2e : neg_inf();
    int2 v0 = make_int2(IDEAD, 0);
    if (DIR) {
      if (f > kDeadWeight) v0 = split_pow2(f * kLog2e);
    } else if (s == 0) {
      v0 = make_int2(0, __float_as_int(1.f));
    }
    abuf0[s] = v0;
  }
  if (tid == 0) abuf0[maxN] = abuf1[maxN] = make_int2(IDEAD, 0);  // the slot absent arcs point at
  for (int j = tid; j < nlp; j += NT) labs[j] = (j < nl) ? g.lab_list[lo + j] : 0;
  __syncthreads();

  auto frame_t = [&](int i) { return DIR ? len - 1 - i : i; };
  const bool direct = a.dense_x != nullptr;
  const float* xb = a.x + (long long)b * a.x_stride_b;  // compact rows [t][nlp]: input (gathered by a kernel) or output (direct)
  float* xout = (direct && DIR == 0 && a.write_grad) ? const_cast<float*>(xb) : nullptr;
  const float* dense = direct ? a.dense_x + (long long)b * a.dense_stride_b : nullptr;
  const unsigned land_a = sm_u32(land);
  constexpr int NLR = NLJ > 0 ? NLJ : 1;
  int labr[NLR];
#pragma unroll
  for (int jj = 0; jj < NLR; ++jj) labr[jj] = (tid + NT * jj < nl) ? labs[tid + NT * jj] : -1;
  auto issue_x = [&](int i) {
    if (i < len) {
      const int t = frame_t(i);
      const int slot = i % XR;
      if (direct) {
        const float* src = dense + (long long)t * a.dense_stride_t;
        const unsigned d0 = land_a + (slot * nlp + tid) * 4;
        if (NLJ > 0) {
#pragma unroll
          for (int jj = 0; jj < NLR; ++jj)
            if (labr[jj] >= 0) cp_async4(d0 + jj * NT * 4, src + labr[jj]);
        } else {
          for (int j = tid; j < nl; j += NT) cp_async4(land_a + (slot * nlp + j) * 4, src + labs[j]);
        }
      } else {
        const float* src = xb + (long long)t * nlp;
        for (int cc = tid; cc < nlp / 4; cc += NT) cp_async16(land + slot * nlp + 4 * cc, src + 4 * cc);
      }
    }
    cp_async_commit();
  };
  // landed row of frame i -> (exponent, mantissa) of e = 2^(x log2 e); the alpha sweep also writes the compact row
  auto transform = [&](int i) {
    if (i >= len) return;
    const float* src = land + (i % XR) * nlp;
    int2* dst = xt + (i & 1) * nlp;
    float* xo = xout ? xout + (long long)frame_t(i) * nlp : nullptr;
    if (NLJ > 0) {
#pragma unroll
      for (int jj = 0; jj < NLR; ++jj) {
        const int j = tid + NT * jj;
        if (j < nlp) {
          const float x = (labr[jj] >= 0) ? src[j] : 0.f;
          dst[j] = split_pow2(x * kLog2e);
          if (xo) xo[j] = x;
        }
      }
    } else {
      for (int j = tid; j < nlp; j += NT) {
        const float x = (j < nl) ? src[j] : 0.f;
        dst[j] = split_pow2(x * kLog2e);
        if (xo) xo[j] = x;
      }
    }
  };
  int2* rows = reinterpret_cast<int2*>((DIR ? a.beta_ws : a.alpha_ws) + (long long)b * a.T * arow);
  const bool spill = a.write_grad != 0;

  for (int i = 0; i < XR - 1; ++i) issue_x(i);
  cp_async_wait<XR - 2>();
  __syncthreads();
  transform(0);
  if (spill && len > 0) {
    int2* r0 = rows + (long long)frame_t(0) * (arow / 2);
    for (int s = tid; s < n; s += NT) r0[s] = abuf0[s];
  }

  unsigned ac = sm_u32(abuf0), an = sm_u32(abuf1);
  const unsigned xt_a = sm_u32(xt);
  constexpr int NSP = SPT > 0 ? SPT : 1;
  unsigned asrc[NSP][DK], alab[NSP][DK];  // byte offsets of the source state / the label
  float wreg[NSP][UNIT ? 1 : DK];
  if (SPT > 0) {
#pragma unroll
    for (int i = 0; i < NSP; ++i) {
      const int s = tid + NT * i;
      const uint4 pk = (s < n) ? pkv[s] : make_uint4(NOARC, NOARC, NOARC, NOARC);
      const unsigned pkk[DEG] = {pk.x, pk.y, pk.z, pk.w};
      const float4 w = (!UNIT && s < n) ? wv[s] : make_float4(0.f, 0.f, 0.f, 0.f);
      const float ww[DEG] = {w.x, w.y, w.z, w.w};
#pragma unroll
      for (int k = 0; k < DK; ++k) {
        asrc[i][k] = (pkk[k] == NOARC) ? (unsigned)maxN * 8u : (pkk[k] & 0xffffu) * 8u;
        alab[i][k] = (pkk[k] == NOARC) ? 0u : (pkk[k] >> 16) * 8u;
        if (!UNIT) wreg[i][k] = ww[k];
      }
    }
  }
  // ONE block barrier per frame.  It publishes (a) the vector and the split label row written during
  // the previous frame, (b) the landed label row of the next frame (every thread has waited for its own
  // copies), and it separates this frame's writes from the previous frame's reads of the same buffers.
  for (int i = 0; i < len; ++i) {
    issue_x(i + XR - 1);
    cp_async_wait<XR - 2>();  // this thread's share of the row of frame i + 1 has landed
    __syncthreads();
    transform(i + 1);
    const unsigned xa = xt_a + (i & 1) * nlp * 8;
    int2* rn = (spill && i + 1 < len) ? rows + (long long)frame_t(i + 1) * (arow / 2) : nullptr;
    if (SPT > 0) {
      // three phases, so that every load is in flight before the first sum
      int J[NSP][DK];
      float pm[NSP][DK];
#pragma unroll
      for (int u = 0; u < NSP; ++u) {
#pragma unroll
        for (int k = 0; k < DK; ++k) {
          const int2 av = lds64(ac + asrc[u][k]);
          const int2 xv = lds64(xa + alab[u][k]);
          J[u][k] = av.x + xv.x;
          pm[u][k] = __int_as_float(av.y) * __int_as_float(xv.y);
          if (!UNIT) {
            const int2 ws = split_pow2(wreg[u][k]);
            J[u][k] += ws.x;
            pm[u][k] *= __int_as_float(ws.y);
          }
        }
      }
      int2 res[NSP];
#pragma unroll
      for (int u = 0; u < NSP; ++u) res[u] = sum_state<DK>(J[u], pm[u]);
#pragma unroll
      for (int u = 0; u < NSP; ++u) {
        const int s = tid + NT * u;
        if (s < n) {
          sts64(an + s * 8, res[u]);
          if (rn) rn[s] = res[u];
        }
      }
    } else {
      constexpr int CH = 2;
      for (int s0 = tid; s0 < n; s0 += NT * CH) {
        int J[CH][DEG];
        float pm[CH][DEG];
#pragma unroll
        for (int u = 0; u < CH; ++u) {
          const int s = s0 + NT * u;
          const uint4 pk = (s < n) ? pkv[s] : make_uint4(NOARC, NOARC, NOARC, NOARC);
          const unsigned pkk[DEG] = {pk.x, pk.y, pk.z, pk.w};
          float ww[DEG] = {0.f, 0.f, 0.f, 0.f};
          if (!UNIT && s < n) {
            const float4 w = wv[s];
            ww[0] = w.x, ww[1] = w.y, ww[2] = w.z, ww[3] = w.w;
          }
#pragma unroll
          for (int k = 0; k < DEG; ++k) {
            const bool v = pkk[k] != NOARC;
            const int2 av = lds64(ac + (v ? (pkk[k] & 0xffffu) : (unsigned)maxN) * 8u);
            const int2 xv = lds64(xa + (v ? (pkk[k] >> 16) : 0u) * 8u);
            J[u][k] = av.x + xv.x;
            pm[u][k] = __int_as_float(av.y) * __int_as_float(xv.y);
            if (!UNIT) {
              const int2 ws = split_pow2(ww[k]);
              J[u][k] += ws.x;
              pm[u][k] *= __int_as_float(ws.y);
            }
          }
        }
#pragma unroll
        for (int u = 0; u < CH; ++u) {
          const int s = s0 + NT * u;
          const int2 res = sum_state<DEG>(J[u], pm[u]);
          if (s < n) {
            sts64(an + s * 8, res);
            if (rn) rn[s] = res;
          }
        }
      }
    }
    const unsigned tmp = ac;
    ac = an;
    an = tmp;
  }
  cp_async_wait<0>();
  if (DIR) return;
  __syncthreads();
  // terminal: cost = log sum_n alpha_len[n] final[n]
  const int2* a_end = (len & 1) ? abuf1 : abuf0;
  int Im = IDEAD;
  for (int s = tid; s < n; s += NT)
    if (fin[s] > neg_inf() && a_end[s].x > IDEAD / 2) Im = max(Im, a_end[s].x);
  Im = __reduce_max_sync(0xffffffffu, Im);
  if (lane == 0) red_i[warp] = Im;
  __syncthreads();
#pragma unroll
  for (int w = 0; w < NT / 32; ++w) Im = max(Im, red_i[w]);
  float ts = 0.f;
  if (Im > IDEAD / 2)
    for (int s = tid; s < n; s += NT)
      if (fin[s] > neg_inf()) ts += __int_as_float(a_end[s].y) * pow2c(a_end[s].x - Im) * ex2(fin[s]);
  ts = warp_sum(ts);
  if (lane == 0) red_f[warp] = ts;
  __syncthreads();
  if (tid == 0) {
    float tot = 0.f;
#pragma unroll
    for (int w = 0; w < NT / 32; ++w) tot += red_f[w];
    const bool feasible = Im > IDEAD / 2 && tot > 0.f;
    a.cost[b] = feasible ? ((double)Im + log2((double)tot)) * kLn2 : -1e20;
  }
}

constexpr int CB_W = 8;             // warps per combine CTA (fewer when the rows are long), one frame at a time each
constexpr int CB_FPW = 4;           // frames per warp: the graph tables are staged once per warps * CB_FPW frames
constexpr int CB_TH = CB_W * 32;

// alpha[src] * beta[dst] * k * 2^(exponents): kE / km hold e_t[j] (and w) with the cost's exponent taken off
__device__ __forceinline__ float post_term(int2 av, int2 bv, int kE, float km) {
  return (__int_as_float(av.y) * __int_as_float(bv.y)) * (km * pow2c(min(av.x + bv.x + kE, 64)));
}

// With a.scatter_grad set the normalised posteriors are subtracted from the dense gradient rows the
// denominator wrote (grad[t][b][lab_list[j]] -= scale_b * gamma_t[j], the scatter_sub step fused in:
// every (t, b, label) is touched by exactly one lane, no atomics); otherwise the compact rows
// gout[b][t][j] are written for the row kernels of plain CTC.
template <bool UNIT>
__global__ void __launch_bounds__(CB_TH, 2) num_combine_kernel(const PathArgs a, const int maxA, const int maxL) {
  extern __shared__ __align__(16) unsigned char sm[];
  const GraphDev& g = a.g;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  // utterances in descending order: with the scatter fused in, the gradient rows written last by the
  // denominator (the highest b) are the ones most likely to be L2 resident still
  const int b = a.B - 1 - (int)blockIdx.y;
  const int len = min(a.lens[b], a.T);
  const int nwarps = blockDim.x >> 5, nthreads = blockDim.x;
  const int cb_fr = nwarps * CB_FPW;
  if (blockIdx.x * cb_fr >= len) return;
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
  float* rows_s = reinterpret_cast<float*>(sm + take(nwarps * (2 * stg + nlp) * 4));
  const int arc0 = g.in_ptr[ro];
  const int narcs = g.in_ptr[ro + n] - arc0;
  for (int j = tid; j <= nl; j += nthreads) lab_ptr[j] = g.lab_ptr[lo + gi + j] - arc0;
  for (int e = tid; e < narcs; e += nthreads) {
    lab_pk[e] = (unsigned)g.lab_src[arc0 + e] | ((unsigned)g.lab_dst[arc0 + e] << 16);
    if (!UNIT) lab_w[e] = g.lab_w[arc0 + e] * kLog2e;
  }
  __syncthreads();
  float* wbase = rows_s + warp * (2 * stg + nlp);
  float* gt = wbase + 2 * stg;
  const double costn = a.cost[b];
  const bool feasible = costn > -1e19;
  // the cost's exponent keeps the terms in range; its mantissa cancels in the row normalisation
  const int cE = (int)rint(costn / kLn2);
  const float gmul = a.g_sign * (a.g_scale ? a.g_scale[b] : 1.f);
  const int V = a.scatter_V;
  auto stage = [&](int q) {  // always commits a group (possibly empty), so that the wait counts stay uniform
    const int t = blockIdx.x * cb_fr + q * nwarps + warp;
    if (q < CB_FPW && t < len && feasible) {
      float* d = wbase + (q & 1) * stg;
      const float* as = a.alpha_ws + ((long long)b * a.T + t) * arow;
      const float* bs = a.beta_ws + ((long long)b * a.T + t) * arow;
      const float* xs = a.x + (long long)b * a.x_stride_b + (long long)t * nlp;
      for (int i = lane; i < (2 * n + 3) / 4; i += 32) {
        cp_async16(d + 4 * i, as + 4 * i);
        cp_async16(d + arow + 4 * i, bs + 4 * i);
      }
      for (int i = lane; i < nlp / 4; i += 32) cp_async16(d + 2 * arow + 4 * i, xs + 4 * i);
    }
    cp_async_commit();
  };
  stage(0);
  for (int q = 0; q < CB_FPW; ++q) {
    const int t = blockIdx.x * cb_fr + q * nwarps + warp;
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
    const int2* at = reinterpret_cast<const int2*>(wbase + (q & 1) * stg);
    const int2* bt = reinterpret_cast<const int2*>(wbase + (q & 1) * stg + arow);
    const float* xt = wbase + (q & 1) * stg + 2 * arow;
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
        const int2 xs = split_pow2(xt[j] * kLog2e);
        const int kE = xs.x - cE;
        const float km = __int_as_float(xs.y);
        for (int e = e0; e < e1; ++e) {
          const unsigned pk = lab_pk[e];
          const int2 ws = UNIT ? make_int2(0, __float_as_int(1.f)) : split_pow2(lab_w[e]);
          acc += post_term(at[pk & 0xffffu], bt[pk >> 16], kE + ws.x, UNIT ? km : km * __int_as_float(ws.y));
        }
      }
      unsigned hm = __ballot_sync(0xffffffffu, heavy);
      while (hm) {  // labels with many arcs (the blank): the whole warp strides over the arcs
        const int src_lane = __ffs(hm) - 1;
        hm &= hm - 1;
        const int jh = base + src_lane;
        const int h0 = __shfl_sync(0xffffffffu, e0, src_lane), h1 = __shfl_sync(0xffffffffu, e1, src_lane);
        const int2 xs = split_pow2(xt[jh] * kLog2e);
        const int kE = xs.x - cE;
        const float km = __int_as_float(xs.y);
        float part = 0.f;
        for (int e = h0 + lane; e < h1; e += 32) {
          const unsigned pk = lab_pk[e];
          const int2 ws = UNIT ? make_int2(0, __float_as_int(1.f)) : split_pow2(lab_w[e]);
          part += post_term(at[pk & 0xffffu], bt[pk >> 16], kE + ws.x, UNIT ? km : km * __int_as_float(ws.y));
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

constexpr int kSmemBudget = 200 * 1024;
int combine_smem_bytes(const PathArgs& a, bool unit, int warps) {
  return ((a.g.max_labels + 1) * 4 + 16) + (a.g.max_arcs * 4 + 16) * (unit ? 1 : 2) +
         warps * (2 * (2 * a.alpha_row + (int)a.x_stride_t) + (int)a.x_stride_t) * 4 + 64;
}
// most warps (<= CB_W) whose staging buffers fit; 0: not even one
int combine_warps(const PathArgs& a, bool unit) {
  int w = CB_W;
  while (w > 0 && combine_smem_bytes(a, unit, w) > kSmemBudget) --w;
  return w;
}
int sweep_smem_bytes(const PathArgs& a, bool unit) { return sweep_plan(a.g.max_states, (int)a.x_stride_t, unit).bytes; }
}  // namespace

template <bool UNIT, int SPT, int DK, int NLJ>
static cudaError_t launch_sweep_nl(const PathArgs& a, int nsweeps, cudaStream_t st) {
  const int bytes = sweep_smem_bytes(a, UNIT);
  cudaError_t e = cudaFuncSetAttribute(num_sweep_kernel<UNIT, SPT, DK, NLJ>, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
  if (e != cudaSuccess) return e;
  num_sweep_kernel<UNIT, SPT, DK, NLJ><<<nsweeps, NT, bytes, st>>>(a, a.g.max_states);
  return cudaGetLastError();
}
template <bool UNIT, int SPT, int DK>
static cudaError_t launch_sweep_dk(const PathArgs& a, int nsweeps, cudaStream_t st) {
  if (SPT > 0 && a.x_stride_t <= NT * 2) return launch_sweep_nl<UNIT, SPT, DK, 2>(a, nsweeps, st);
  return launch_sweep_nl<UNIT, SPT, DK, 0>(a, nsweeps, st);
}
template <bool UNIT>
static cudaError_t launch_sweep(const PathArgs& a, int nsweeps, cudaStream_t st) {
  const int spt = (a.g.max_states + NT - 1) / NT;
  const bool d3 = a.g.max_in_degree <= 3 && a.g.max_out_degree <= 3;
  if (d3 && spt <= 1) return launch_sweep_dk<UNIT, 1, 3>(a, nsweeps, st);
  if (d3 && spt <= 2) return launch_sweep_dk<UNIT, 2, 3>(a, nsweeps, st);
  if (d3 && spt <= 4) return launch_sweep_dk<UNIT, 4, 3>(a, nsweeps, st);
  if (spt <= 2) return launch_sweep_dk<UNIT, 2, DEG>(a, nsweeps, st);
  return launch_sweep_dk<UNIT, 0, DEG>(a, nsweeps, st);
}

bool num_split_supported(const PathArgs& a) {
  if (!a.beta_ws) return false;
  if (a.g.max_states > 65535 || a.g.max_labels > 65535 || a.g.max_arcs < 1) return false;
  if (a.g.max_in_degree > DEG || a.g.max_out_degree > DEG) return false;
  if (a.alpha_row < 2 * a.g.max_states) return false;
  const bool unit = a.g.unit_weights != 0;
  return sweep_smem_bytes(a, unit) <= kSmemBudget && combine_warps(a, unit) >= 1;
}

cudaError_t launch_num_sweeps(const PathArgs& a, cudaStream_t st) {
  const bool unit = a.g.unit_weights != 0;
  ProfScope prof("num_sweep", st);
  const int nsweeps = a.write_grad ? 2 * a.B : a.B;
  const cudaError_t e = unit ? launch_sweep<true>(a, nsweeps, st) : launch_sweep<false>(a, nsweeps, st);
  ++g_launches;
  return e;
}

cudaError_t launch_num_combine(const PathArgs& a, cudaStream_t st) {
  if (!a.write_grad) return cudaSuccess;
  const bool unit = a.g.unit_weights != 0;
  cudaError_t e;
  ProfScope prof("num_combine", st);
  const int warps = combine_warps(a, unit);
  const int bytes = combine_smem_bytes(a, unit, warps);
  const int cb_fr = warps * CB_FPW;
  const dim3 grid((a.T + cb_fr - 1) / cb_fr, a.B);
  if (unit) {
    e = cudaFuncSetAttribute(num_combine_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
    if (e != cudaSuccess) return e;
    num_combine_kernel<true><<<grid, warps * 32, bytes, st>>>(a, a.g.max_arcs, a.g.max_labels);
  } else {
    e = cudaFuncSetAttribute(num_combine_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
    if (e != cudaSuccess) return e;
    num_combine_kernel<false><<<grid, warps * 32, bytes, st>>>(a, a.g.max_arcs, a.g.max_labels);
  }
  ++g_launches;
  return cudaGetLastError();
}

}  // namespace ctcg
