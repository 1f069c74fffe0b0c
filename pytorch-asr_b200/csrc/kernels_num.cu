// Numerator (per-utterance training graph) forward-backward, sm_100a: one CTA per (utterance,
// direction) runs the alpha-type or the beta-type recursion (num_sweep_kernel), then a time-parallel
// kernel forms the posteriors (num_combine_kernel).
//
// Reference semantics: PathLogSumExp.forward/backward (att_speech/fst_utils.py:400-488) on the
// graphs of BaseGraphGen.get_training_matrices (fst_utils.py:647-660): N = 2U+1 states with at
// most 3 arcs into / out of a state.
//
// One recursion for both directions.  In a training graph (decoding graph composed with the
// transcript chain) every arc into a state carries the same label, lab(s) -- the graph compiler
// checks it (GraphDev::in_label_uniform); other graphs take the generic CSR kernel.  With
// e_t[j] = exp(x_t[j]):
//   forward   alpha_{t+1}[s] = e_t[lab(s)]     * sum_{arcs k into s}   w_k alpha_t[src_k]
//   backward  C_{t-1}[s]     = e_{t-1}[lab(s)] * sum_{arcs k out of s} w_k C_t[dst_k],
//             C_t[d] = e_t[lab(d)] beta_{t+1}[d],   C_{len-1} = e_{len-1}[lab] * final
// i.e. v' = e[lab] (.) (M v) with M the in-arc or the out-arc matrix: one label look-up per state
// instead of one per arc.  The backward sweep spills beta_{t+1} = M C_{t+1} (the sum before the label
// factor), the forward sweep alpha_{t+1}; the posterior of a label is then the occupancy of its states,
//   gamma_t[j] = sum_{states d with lab(d) = j} alpha_{t+1}[d] beta_{t+1}[d] / Z,
// one product per state with no arcs, weights or logits left in it.
//
// Number domain.  A training graph is a chain, so alpha_t of the states that carry the posterior
// mass sits hundreds of binary orders below max_s alpha_t[s]: numbers scaled per frame underflow
// (measured at T = 400) and a plain fp32 log value of that magnitude has an ulp of ~1.5e-5, which
// is where the round-1 kernel lost the gradient tolerance at T >= 800.  Every value is therefore a
// block-floating-point pair  v = f * 2^E  with an fp32 mantissa f in {0} U [1, 2) and an exact
// int32 exponent E per state: exponents add exactly, mantissas multiply with a relative error of
// 2^-24 whatever the magnitude, a sum aligns its terms to the largest exponent with a power of two
// built in the integer pipe.  No per-frame normaliser, nothing lagged, and no transcendental on
// the recursion: the only MUFU work is one ex2 per label and frame for e = 2^frac(x log2 e) (the
// integer part of x log2 e goes to the exponent), done once per batch by gather_kernel.
//
// num_sweep_kernel   256 threads.  7 compute warps own 1..4 states per thread (arcs as shared-memory
//                    byte offsets in registers); the vector lives in shared memory (int2 per state,
//                    double buffered).  One producer warp streams the rows of (exponent, mantissa)
//                    pairs of the label columns this utterance touches -- written once per batch by
//                    gather_kernel (kernels_rows.cu), which spreads the scattered 4-byte reads of the
//                    (T, B, V) logits over the whole GPU -- into a 16-row ring with 16-byte cp.async
//                    copies.  One block barrier per frame, no atomics, no MUFU.
// num_combine_kernel one warp per frame: the sum above in a fixed order (states grouped by label on the
//                    host), row normalised to one; with a factorised denominator the compact rows go
//                    to the denominator's combine kernel, which subtracts them while it writes.
#include <algorithm>
#include <cstdlib>
#include <type_traits>

#include "device_utils.cuh"
#include "kernels.cuh"

namespace ctcg {
namespace {

constexpr int NCW = 7;           // compute warps of a sweep CTA
constexpr int NPW = 1;           // producer warp
constexpr int NCT = NCW * 32;    // compute threads: state s = tid + NCT * u
constexpr int NPT = NPW * 32;    // producer threads
constexpr int NT = NCT + NPT;
constexpr int MAXSPT = 4;        // states per compute thread
constexpr int ER = 16;           // (exponent, mantissa) rows of the label columns in flight
constexpr int DEG = 4;           // arcs per state the sweeps take
constexpr int HEAVY = 12;        // combine: labels with more arcs are summed by the whole warp
constexpr int IDEAD = -(1 << 28);          // exponent of "no path"
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

// x (log2 units, clamped) -> (E, f): 2^x = f * 2^E with E = rint(x), f = 2^(x - E) in [0.707, 1.415]
__device__ __forceinline__ int2 split_pow2(float xs) {
  xs = fminf(fmaxf(xs, -XCLAMP), XCLAMP);
  const float t = xs + MAGIC;
  return make_int2(__float_as_int(t) - MAGIC_BITS, __float_as_int(ex2(xs - (t - MAGIC))));
}
// 2^d for d <= 127; anything at or below -127 is exactly 0
__device__ __forceinline__ float pow2c(int d) { return __int_as_float((max(d, -127) + 127) << 23); }
// v > 0 normal -> (base + floor(log2 v), mantissa in [1, 2)); v == 0 -> dead
__device__ __forceinline__ int2 renorm(float v, int base) {
  const int bits = __float_as_int(v);
  const int2 out = make_int2(base + ((bits >> 23) - 127), (bits & 0x007fffff) | 0x3f800000);
  return (v > 0.f) ? out : make_int2(IDEAD, 0);
}

struct SweepPlan {
  int a0, a1, er, bytes;
};
__host__ __device__ inline SweepPlan sweep_plan(int maxN, int nlp) {
  SweepPlan p;
  int o = 0;
  auto take = [&](int bytes) {
    const int at = o;
    o += (bytes + 15) & ~15;
    return at;
  };
  p.a0 = take((maxN + 1) * 8);
  p.a1 = take((maxN + 1) * 8);
  p.er = take(ER * nlp * 8);
  p.bytes = o;
  return p;
}

// One state: sum over its DK arcs, terms aligned to the largest exponent, times the label factor.
// An absent / dead arc has mantissa 0 and an exponent near IDEAD, far below any live one.
template <int DK>
__device__ __forceinline__ int2 step_state(const int (&J)[DK], const float (&p)[DK], const int2 ev, int2* pre) {
  int Jm = J[0];
#pragma unroll
  for (int k = 1; k < DK; ++k) Jm = max(Jm, J[k]);
  float acc = 0.f;
#pragma unroll
  for (int k = 0; k < DK; ++k) acc = fmaf(p[k], pow2c(J[k] - Jm), acc);
  *pre = renorm(acc, Jm);  // (M v)[s]: beta_{t+1}[s] in the backward sweep
  return renorm(acc * __int_as_float(ev.y), Jm + ev.x);
}

// SPT: states per compute thread (state s = tid + NCT u).  DK: arcs per state that exist anywhere in
// the batch (3 for CTC training graphs, fst_utils.py:592-605).
template <bool UNIT, int SPT, int DK>
__global__ void __launch_bounds__(NT) num_sweep_kernel(const PathArgs a, const int maxN) {
  extern __shared__ __align__(16) unsigned char sm[];
  __shared__ int red_i[NCW];
  __shared__ float red_f[NCW];
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
  const int len = min(a.lens[b], a.T);
  const int nlp = (int)a.x_stride_t, arow2 = a.alpha_row / 2;
  const SweepPlan P = sweep_plan(maxN, nlp);
  const unsigned a0_a = sm_u32(sm + P.a0), a1_a = sm_u32(sm + P.a1);
  const unsigned er_a = sm_u32(sm + P.er);
  const bool spill = a.write_grad != 0;
  int2* rows = reinterpret_cast<int2*>((DIR ? a.beta_ws : a.alpha_ws) + (long long)b * (a.T + 1) * a.alpha_row);  // T + 1 rows: alpha_len has a slot too
  const bool producer = tid >= NCT;
  const int i0 = DIR;  // the backward sweep spends frame len-1 in its start vector

  // ---- compute threads: arcs, label and start value of their states ----
  unsigned asrc[SPT][DK], lab8[SPT];
  int wE[SPT][UNIT ? 1 : DK];
  float wm[SPT][UNIT ? 1 : DK];
  if (!producer) {
    const int32_t* ptr = DIR ? g.out_ptr : g.in_ptr;
    const int32_t* oth = DIR ? g.out_dst : g.in_src;
    const float* wgt = DIR ? g.out_w : g.in_w;
#pragma unroll
    for (int u = 0; u < SPT; ++u) {
      const int s = tid + NCT * u;
      const bool valid = s < n;
      const int e0 = valid ? ptr[ro + s] : 0, e1 = valid ? ptr[ro + s + 1] : 0;
#pragma unroll
      for (int k = 0; k < DK; ++k) {
        const bool v = e0 + k < e1;
        asrc[u][k] = v ? (unsigned)oth[e0 + k] * 8u : (unsigned)maxN * 8u;
        if (!UNIT) {
          const int2 ws = split_pow2(v ? wgt[e0 + k] * kLog2e : 0.f);
          wE[u][k] = ws.x;
          wm[u][k] = __int_as_float(ws.y);
        }
      }
      int cl = 0;  // the label on the arcs INTO the state, for both directions; none: the value is never read
      if (valid && g.in_ptr[ro + s] < g.in_ptr[ro + s + 1]) cl = g.in_clab[g.in_ptr[ro + s]];
      lab8[u] = (unsigned)cl * 8u;
    }
    if (tid == 0) {
      sts64(a0_a + maxN * 8, make_int2(IDEAD, 0));  // the slot absent arcs point at
      sts64(a1_a + maxN * 8, make_int2(IDEAD, 0));
    }
  }

  // ---- producer warp: rows of (exponent, mantissa) pairs written by gather_kernel -> ring, ER - 1 frames ahead ----
  auto frame_t = [&](int i) { return DIR ? len - 1 - i : i; };
  const int ptid = tid - NCT;
  const int2* ebase = reinterpret_cast<const int2*>(a.x) + (long long)b * a.x_stride_b;  // [t][nlp]
  const int chunks = nlp / 2;  // 16-byte chunks per row (nlp is a multiple of 4)
  // this lane's chunks of a row: ptid, ptid + NPT, ...; the source pointer walks the rows in sweep order
  const int n_mine = producer ? max(0, (chunks - ptid + NPT - 1) / NPT) : 0;
  const long long pstep = (DIR ? -8LL : 8LL) * nlp;
  const char* psrc = reinterpret_cast<const char*>(ebase + (long long)frame_t(0) * nlp) + 16 * ptid;
  auto issue = [&](int r) {  // copies of row r (frame frame_t(r)); always commits a group
    if (r < len) {
      unsigned d = er_a + (r & (ER - 1)) * nlp * 8 + 16 * ptid;
      const char* sp = psrc;
      for (int k = 0; k < n_mine; ++k) {
        asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(d), "l"(sp) : "memory");
        d += 16 * NPT;
        sp += 16 * NPT;
      }
      psrc += pstep;
    }
    cp_async_commit();
  };
  if (producer) {
    for (int r = 0; r < ER - 1 + i0; ++r) issue(r);
    cp_async_wait<0>();
  }
  __syncthreads();

  // ---- start vector: alpha_0 = [1, 0, ...];  C_{len-1} = e_{len-1}[lab] * final ----
  if (!producer) {
#pragma unroll
    for (int u = 0; u < SPT; ++u) {
      const int s = tid + NCT * u;
      if (s < n) {
        int2 v0 = make_int2(IDEAD, 0);
        if (DIR && len > 0) {
          const float f = g.final_w[so + s];
          if (f > kDeadWeight) {
            const int2 fs = split_pow2(f * kLog2e);
            const int2 ev = lds64(er_a + lab8[u]);  // row 0 of the backward order = frame len-1 (ring slot 0)
            v0 = renorm(__int_as_float(fs.y) * __int_as_float(ev.y), fs.x + ev.x);
          }
        } else if (!DIR && s == 0) {
          v0 = make_int2(0, __float_as_int(1.f));
        }
        sts64(a0_a + s * 8, v0);
        if (spill) {  // forward: row 0 = alpha_0; backward: row len = beta_len = final
          int2 r0 = v0;
          if (DIR) {
            const float f = g.final_w[so + s];
            r0 = (f > kDeadWeight) ? split_pow2(f * kLog2e) : make_int2(IDEAD, 0);
          }
          rows[(DIR ? (long long)len : 0LL) * arow2 + s] = r0;
        }
      }
    }
  }
  unsigned ac = a0_a, an = a1_a;
  // what a step spills: forward alpha_{i+1} -> row i+1; backward beta_{t+1} = M C_{t+1}, t = len-1-i -> row len-i
  // (the combine reads the state occupancies alpha_{t+1} beta_{t+1} from rows t+1 of both)
  long long orow = (DIR ? (long long)(len - i0) : (long long)(i0 + 1)) * arow2;
  const long long ostep = DIR ? -(long long)arow2 : (long long)arow2;

  // ONE block barrier per frame.  It publishes the vector written during the previous step and the
  // label row the producer's copies completed during the previous step, and it separates this step's
  // writes (vector buffer, ring slot of row i - 1) from the previous step's reads of them.
  for (int i = i0; i < len; ++i) {
    __syncthreads();
    if (producer) {
      issue(i + ER - 1);
      cp_async_wait<ER - 2>();  // row i + 1 has landed
    } else {
      const unsigned ea = er_a + (i & (ER - 1)) * nlp * 8;  // ER is a power of two
      int J[SPT][DK];
      float pm[SPT][DK];
      int2 ev[SPT];
#pragma unroll
      for (int u = 0; u < SPT; ++u) {  // every load in flight before the first sum
        ev[u] = lds64(ea + lab8[u]);
#pragma unroll
        for (int k = 0; k < DK; ++k) {
          const int2 av = lds64(ac + asrc[u][k]);
          J[u][k] = UNIT ? av.x : av.x + wE[u][k];
          pm[u][k] = UNIT ? __int_as_float(av.y) : __int_as_float(av.y) * wm[u][k];
        }
      }
#pragma unroll
      for (int u = 0; u < SPT; ++u) {
        const int s = tid + NCT * u;
        int2 pre;
        const int2 res = step_state<DK>(J[u], pm[u], ev[u], &pre);
        if (s < n) {
          sts64(an + s * 8, res);
          if (spill) rows[orow + s] = DIR ? pre : res;
        }
      }
    }
    orow += ostep;
    const unsigned tmp = ac;
    ac = an;
    an = tmp;
  }
  if (producer) cp_async_wait<0>();
  if (DIR) return;
  __syncthreads();
  // terminal: cost = log sum_n alpha_len[n] final[n]
  int Im = IDEAD;
  int2 ae[SPT];
  float fw[SPT];
#pragma unroll
  for (int u = 0; u < SPT; ++u) {
    const int s = tid + NCT * u;
    ae[u] = make_int2(IDEAD, 0);
    fw[u] = 0.f;
    if (!producer && s < n) {
      const float f = g.final_w[so + s];
      if (f > kDeadWeight) {
        ae[u] = lds64(ac + s * 8);
        fw[u] = ex2(f * kLog2e);
        if (ae[u].x > IDEAD / 2) Im = max(Im, ae[u].x);
      }
    }
  }
  Im = __reduce_max_sync(0xffffffffu, Im);
  if (lane == 0 && !producer) red_i[warp] = Im;
  __syncthreads();
#pragma unroll
  for (int w = 0; w < NCW; ++w) Im = max(Im, red_i[w]);
  float ts = 0.f;
  if (Im > IDEAD / 2) {
#pragma unroll
    for (int u = 0; u < SPT; ++u) ts += __int_as_float(ae[u].y) * pow2c(ae[u].x - Im) * fw[u];
  }
  ts = warp_sum(ts);
  if (lane == 0 && !producer) red_f[warp] = ts;
  __syncthreads();
  if (tid == 0) {
    float tot = 0.f;
#pragma unroll
    for (int w = 0; w < NCW; ++w) tot += red_f[w];
    const bool feasible = Im > IDEAD / 2 && tot > 0.f;
    a.cost[b] = feasible ? ((double)Im + log2((double)tot)) * kLn2 : -1e20;
  }
}

constexpr int CB_W = 8;             // warps per combine CTA (fewer when the rows are long), one frame at a time each
constexpr int CB_FPW = 4;           // frames per warp: the graph tables are staged once per warps * CB_FPW frames
constexpr int CB_TH = CB_W * 32;

// alpha_{t+1}[d] * beta_{t+1}[d] * 2^(exponents - cost exponent): the occupancy of state d after frame t
__device__ __forceinline__ float post_term(int2 av, int2 bv, int kE) {
  return (__int_as_float(av.y) * __int_as_float(bv.y)) * pow2c(min(av.x + bv.x + kE, 64));
}

// gamma_t[j] = sum_{states d with lab(d) = j} alpha_{t+1}[d] beta_{t+1}[d] / Z: one product per STATE (the arcs,
// their weights and the logits are already inside alpha_{t+1} and beta_{t+1}), summed in a fixed order (states
// grouped by label on the host) and normalised so that the row sums to one.  Writes the compact posterior rows
// gout[b][t][j] (times g_sign * g_scale[b]): the row kernels of plain CTC and the denominator's combine kernel
// (CTC-G) fold them into the dense gradient.
// SMALL: 128-thread CTAs capped at 80 registers (10 k registers per CTA), the footprint that fits next to a
// denominator sweep CTA (416 threads x 126 registers = 52 k of the SM's 64 k): with a factorised denominator the
// kernel then runs DURING the sweeps instead of waiting for their SMs.
template <bool SMALL>
__global__ void __launch_bounds__(SMALL ? 128 : CB_TH, SMALL ? 6 : 2)
    num_combine_kernel(const PathArgs a, const int maxN, const int maxL) {
  extern __shared__ __align__(16) unsigned char sm[];
  const GraphDev& g = a.g;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int b = blockIdx.y;
  const int len = min(a.lens[b], a.T);
  const int nwarps = blockDim.x >> 5, nthreads = blockDim.x;
  const int cb_fr = nwarps * CB_FPW;
  if (blockIdx.x * cb_fr >= len) return;
  const int gi = (g.num_graphs == 1) ? 0 : b;
  const int so = g.state_off[gi];
  const int n = g.state_off[gi + 1] - so;
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
  int* lab_st = reinterpret_cast<int*>(sm + take(maxN * 4));
  // per warp: two staging buffers {alpha row, beta row} (the next frame's rows are copied in with cp.async
  // while this frame's posteriors are summed) and one posterior row
  const int stg = 2 * arow;
  float* rows_s = reinterpret_cast<float*>(sm + take(nwarps * (2 * stg + nlp) * 4));
  for (int j = tid; j <= nl; j += nthreads) lab_ptr[j] = g.slab_ptr[lo + gi + j] - so;
  for (int e = tid; e < n; e += nthreads) lab_st[e] = g.slab_state[so + e];
  __syncthreads();
  float* wbase = rows_s + warp * (2 * stg + nlp);
  float* gt = wbase + 2 * stg;
  const double costn = a.cost[b];
  const bool feasible = costn > -1e19;
  // the cost's exponent keeps the terms in range; its mantissa cancels in the row normalisation
  const int kE = -(int)rint(costn / kLn2);
  const float gmul = a.g_sign * (a.g_scale ? a.g_scale[b] : 1.f);
  auto stage = [&](int q) {  // always commits a group (possibly empty), so that the wait counts stay uniform
    const int t = blockIdx.x * cb_fr + q * nwarps + warp;
    if (q < CB_FPW && t < len && feasible) {
      float* d = wbase + (q & 1) * stg;
      const float* as = a.alpha_ws + ((long long)b * (a.T + 1) + t + 1) * arow;   // alpha_{t+1}
      const float* bs = a.beta_ws + ((long long)b * (a.T + 1) + t + 1) * arow;    // beta_{t+1}
      for (int i = lane; i < (2 * n + 3) / 4; i += 32) {
        cp_async16(d + 4 * i, as + 4 * i);
        cp_async16(d + arow + 4 * i, bs + 4 * i);
      }
    }
    cp_async_commit();
  };
  stage(0);
  for (int q = 0; q < CB_FPW; ++q) {
    const int t = blockIdx.x * cb_fr + q * nwarps + warp;
    if (t >= len) break;
    // row-major [t][j], or segment-major [t / seg][j][t % seg] for the denominator's combine kernel
    const int seg = a.g_seg;
    float* dst = a.gout + (long long)b * a.g_stride_b +
                 (seg > 0 ? (long long)(t / seg) * nlp * seg + (t % seg) : (long long)t * a.g_stride_t);
    const int dstep = seg > 0 ? seg : 1;
    if (!feasible) {  // no accepting path: zero posterior
      for (int j = lane; j < nl; j += 32) dst[j * dstep] = 0.f;
      continue;
    }
    __syncwarp();       // the previous frame's reads of the buffer about to be refilled are done
    stage(q + 1);
    cp_async_wait<1>();
    __syncwarp();
    const int2* at = reinterpret_cast<const int2*>(wbase + (q & 1) * stg);
    const int2* bt = reinterpret_cast<const int2*>(wbase + (q & 1) * stg + arow);
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
        for (int e = e0; e < e1; ++e) {
          const int d = lab_st[e];
          acc += post_term(at[d], bt[d], kE);
        }
      }
      unsigned hm = __ballot_sync(0xffffffffu, heavy);
      while (hm) {  // labels with many states (the blank): the whole warp strides over them
        const int src_lane = __ffs(hm) - 1;
        hm &= hm - 1;
        const int h0 = __shfl_sync(0xffffffffu, e0, src_lane), h1 = __shfl_sync(0xffffffffu, e1, src_lane);
        float part = 0.f;
        for (int e = h0 + lane; e < h1; e += 32) {
          const int d = lab_st[e];
          part += post_term(at[d], bt[d], kE);
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
    for (int j = lane; j < nl; j += 32) dst[j * dstep] = gt[j] * sc;
  }
}

constexpr int kSmemBudget = 200 * 1024;
int combine_smem_bytes(const PathArgs& a, int warps) {
  return ((a.g.max_labels + 1) * 4 + 16) + (a.g.max_states * 4 + 16) +
         warps * (2 * (2 * a.alpha_row) + (int)a.x_stride_t) * 4 + 64;
}
// most warps (<= CB_W) whose staging buffers fit; 0: not even one
int combine_warps(const PathArgs& a, int most = CB_W) {
  int w = most;
  while (w > 0 && combine_smem_bytes(a, w) > kSmemBudget) --w;
  return w;
}
int sweep_smem_bytes(const PathArgs& a) { return sweep_plan(a.g.max_states, (int)a.x_stride_t).bytes; }

}  // namespace

template <bool UNIT, int SPT, int DK>
static cudaError_t launch_sweep_dk(const PathArgs& a, int nsweeps, cudaStream_t st) {
  const int bytes = sweep_smem_bytes(a);
  static int attr_cache[64] = {};
  cudaError_t e = smem_attr_once(num_sweep_kernel<UNIT, SPT, DK>, bytes, attr_cache);
  if (e != cudaSuccess) return e;
  num_sweep_kernel<UNIT, SPT, DK><<<nsweeps, NT, bytes, st>>>(a, a.g.max_states);
  return cudaGetLastError();
}
template <bool UNIT, int SPT>
static cudaError_t launch_sweep_spt(const PathArgs& a, int nsweeps, cudaStream_t st) {
  if (a.g.max_in_degree <= 3 && a.g.max_out_degree <= 3) return launch_sweep_dk<UNIT, SPT, 3>(a, nsweeps, st);
  return launch_sweep_dk<UNIT, SPT, DEG>(a, nsweeps, st);
}
template <bool UNIT>
static cudaError_t launch_sweep(const PathArgs& a, int nsweeps, cudaStream_t st) {
  const int spt = (a.g.max_states + NCT - 1) / NCT;
  if (spt <= 1) return launch_sweep_spt<UNIT, 1>(a, nsweeps, st);
  if (spt <= 2) return launch_sweep_spt<UNIT, 2>(a, nsweeps, st);
  return launch_sweep_spt<UNIT, MAXSPT>(a, nsweeps, st);
}

bool num_split_supported(const PathArgs& a) {
  if (!a.beta_ws) return false;
  if (a.g.max_states > 65535 || a.g.max_labels > 65535 || a.g.max_arcs < 1) return false;
  if (a.g.max_in_degree > DEG || a.g.max_out_degree > DEG) return false;
  if (a.alpha_row < 2 * a.g.max_states) return false;
  if (!a.g.in_label_uniform || a.g.max_states > NCT * MAXSPT) return false;  // the generic CSR kernel takes these
  return sweep_smem_bytes(a) <= kSmemBudget && combine_warps(a) >= 1;
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
  ProfScope prof("num_combine", st);
  // segment-major output = the denominator's combine will read it = the denominator sweeps are running now
  const bool small = a.g_seg > 0;
  const int warps = combine_warps(a, small ? 4 : CB_W);
  const int bytes = combine_smem_bytes(a, warps);
  const int cb_fr = warps * CB_FPW;
  const dim3 grid((a.T + cb_fr - 1) / cb_fr, a.B);
  cudaError_t e;
  if (small) {
    static int attr_cache[64] = {};
    e = smem_attr_once(num_combine_kernel<true>, bytes, attr_cache);
    if (e != cudaSuccess) return e;
    num_combine_kernel<true><<<grid, warps * 32, bytes, st>>>(a, a.g.max_states, a.g.max_labels);
  } else {
    static int attr_cache[64] = {};
    e = smem_attr_once(num_combine_kernel<false>, bytes, attr_cache);
    if (e != cudaSuccess) return e;
    num_combine_kernel<false><<<grid, warps * 32, bytes, st>>>(a, a.g.max_states, a.g.max_labels);
  }
  ++g_launches;
  return cudaGetLastError();
}

}  // namespace ctcg
