// Fused CTC-G / CTC-GB denominator for the bi-char decoding graph, one CTA per
// utterance, linear (scaled) domain, sm_100a.
//
// Reference semantics: PathLogSumExp.forward/backward (att_speech/fst_utils.py:400-488)
// on the decoding graph of build_ctc_bigram_decoding_fst (fst_utils.py:729-835),
// called from FSTDecoder.get_fst_loss (advanced_decoder.py:497-500).
//
// The decoding graph has S*S states (c,l) and ~S^3 arcs, but every arc entering
// (c2,l2) carries the label of (c2,l2) and its predecessor set depends on c2
// only (fst_utils.py:783-816).  With a[c][l] = exp(alpha) and e = exp(x - max):
//     P[c2]      = [c2 != 0] * sum_c1 a[c1][c2] + a[c2][0]            (49 column sums)
//     a'[c2][l2] = e[lab(c2,l2)] * (P[c2] + loop(c2,l2) * a[c2][l2])  (elementwise)
// and symmetrically for beta with row sums Q[c] = sum_l e*b.  The posterior of
// label lab(c,l) at frame t is a_{t+1}[c][l] * b_{t+1}[c][l] / Z: no scatter.
//
// HBM traffic per valid frame: x read in the forward sweep (4V), x read again in
// the backward sweep (4V), gradient written once (4V), plus the P spill (S+3
// floats) and one alpha checkpoint every KSEG frames (8V/KSEG): ~13V against the
// algorithmic 8V.  alpha is never spilled densely: the backward sweep recomputes
// it elementwise from P_t (no reductions) KSEG frames ahead of the beta recursion.
//
// Roles inside the CTA (lockstep, ONE __syncthreads per frame):
//   main warps   the serial recursion; registers hold a (or b), 8 states/thread
//   4 producers  TMA-landed logits row -> e = exp2((x-m)*log2e) in place
//   4 recompute  backward sweep only: a_{t+1} = 2^k e (P_t + loop a_t) elementwise
//   scale warp   total mass -> lagged power-of-two rescale, final cost
//   io warp      issues cp.async.bulk loads (UBLKCP), bulk-stores gradient rows,
//                zero-fills the rows beyond the utterance's end
#include <cstdio>

#include "device_utils.cuh"
#include "kernels.cuh"

namespace ctcg {
namespace {

constexpr int RPW = 4;        // majors (rows fwd / columns bwd) per main warp
constexpr int NPROD = 4;      // exp producer warps
constexpr int NRECO = 4;      // alpha-recompute warps (backward sweep)
constexpr int PTHREADS = NPROD * 32;
constexpr int PARTW = 64;     // width of a partial-sum row (2 slots per lane)
constexpr int TARGET = 60;    // scaled mass is kept near 2^TARGET
constexpr int BAND = 20;
constexpr int QRING = 32;     // frames-in-flight tables (power of two, > 2*K + D + 5)
constexpr int DF = 8;         // TMA lookahead of the forward sweep (frames)

// ---- PTX wrappers -----------------------------------------------------------
__device__ __forceinline__ unsigned smem_u32(const void* p) {
  return static_cast<unsigned>(__cvta_generic_to_shared(p));
}
__device__ __forceinline__ void mbar_init(uint64_t* bar, unsigned count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, unsigned bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes)
               : "memory");
}
__device__ __forceinline__ bool mbar_try(uint64_t* bar, unsigned parity) {
  unsigned ok;
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n"
      "selp.u32 %0, 1, 0, p;\n"
      "}\n"
      : "=r"(ok)
      : "r"(smem_u32(bar)), "r"(parity)
      : "memory");
  return ok != 0;
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, unsigned parity) {
  while (!mbar_try(bar, parity)) {
  }
}
__device__ __forceinline__ void mbar_inval(uint64_t* bar) {
  asm volatile("mbarrier.inval.shared::cta.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}
// 1-D bulk copy global -> shared (TMA engine, SASS UBLKCP); 16-byte aligned, size % 16 == 0.
__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, unsigned bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                   smem_u32(dst)),
               "l"(src), "r"(bytes), "r"(smem_u32(bar))
               : "memory");
}
__device__ __forceinline__ void bulk_s2g(void* dst, const void* src, unsigned bytes) {
  asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(dst), "r"(smem_u32(src)),
               "r"(bytes)
               : "memory");
}
__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void bulk_wait_read() {
  asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(N) : "memory");
}
template <int N>
__device__ __forceinline__ void bulk_wait_all() {
  asm volatile("cp.async.bulk.wait_group %0;" ::"n"(N) : "memory");
}
__device__ __forceinline__ void fence_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

__device__ __forceinline__ float pow2i(int k) {  // 2^k for -126 <= k <= 127
  return __int_as_float((k + 127) << 23);
}
__device__ __forceinline__ int exponent_of(float v) {  // floor(log2 v) for normal v > 0
  return ((__float_as_int(v) >> 23) & 0xff) - 127;
}


// ---- shared-memory plan -------------------------------------------------------
struct Smem {
  int rows, part, prow, maxpart, tab_e, tab_a, freel, phase, ksm, csm, scal, mbar;
  int bytes;
};

__host__ __device__ inline Smem smem_plan(int NR, int rowf, int nmain, int prowf) {
  Smem s;
  int o = 0;
  s.rows = o;     o += NR * rowf * 4;
  s.part = o;     o += 2 * (nmain + 1) * PARTW * 4;
  s.prow = o;     o += QRING * prowf * 4;
  s.maxpart = o;  o += 2 * NPROD * 4;
  s.tab_e = o;    o += QRING * 4;
  s.tab_a = o;    o += QRING * 4;
  s.freel = o;    o += (NR + 2) * 4;
  s.phase = o;    o += NR * 4;
  s.ksm = o;      o += 2 * 4;
  s.csm = o;      o += 4 * 4;
  o = (o + 15) & ~15;
  s.scal = o;     o += 64;          // double msum; float inv_mass; int klen; int dead; int nfree
  s.mbar = o;     o += NR * 8;
  s.bytes = o;
  return s;
}

struct Scal {
  double msum;
  float inv_mass;
  int klen;
  int dead;
  int nfree;
  int zrow;
};

// Geometry of one (t, b) logits row inside its 16-byte aligned window.
struct RowGeo {
  long long w0;   // first float of the window (multiple of 4)
  int r;          // row starts at window float r (0..3)
  int nbulk;      // floats fetched by the bulk copy (multiple of 4)
  int nfl;        // floats of the padded window, pad4(r + V)
};
__device__ __forceinline__ RowGeo row_geo(int t, int b, int B, int V, long long total) {
  RowGeo g;
  const long long i0 = ((long long)t * B + b) * V;
  g.r = (int)(i0 & 3);
  g.w0 = i0 - g.r;
  g.nfl = pad4(g.r + V);
  const long long room = (total - g.w0) & ~3LL;
  g.nbulk = (int)(room < (long long)g.nfl ? room : (long long)g.nfl);
  return g;
}

// Production order of the backward sweep: segments of K frames from the last to
// the first, frames ascending inside a segment (the recompute runs forward in time).
__device__ __forceinline__ int frame_of_q(int q, int len, int K) {
  const int J = (len + K - 1) / K;
  const int L0 = len - (J - 1) * K;
  if (q < L0) return (J - 1) * K + q;
  const int qq = q - L0;
  return (J - 2 - qq / K) * K + qq % K;
}
__device__ __forceinline__ int q_of_frame(int t, int len, int K) {
  const int J = (len + K - 1) / K;
  const int L0 = len - (J - 1) * K;
  const int j = t / K;
  return (j == J - 1) ? (t - j * K) : (L0 + (J - 2 - j) * K + (t - j * K));
}

// In-warp reduction of the partial-sum rows: lanes 8k..8k+7 add the NROWS
// partials of column col[k]; every lane returns the RPW totals.
__device__ __forceinline__ void reduce_partials(const float* pb, int nrows, const int (&col)[RPW], int lane,
                                                float (&out)[RPW]) {
  const int grp = lane >> 3, j8 = lane & 7;
  int c = col[0];
#pragma unroll
  for (int k = 1; k < RPW; ++k) c = (grp == k) ? col[k] : c;
  float s = 0.f;
  for (int rr = j8; rr < nrows; rr += 8) s += pb[rr * PARTW + c];
  s += __shfl_xor_sync(0xffffffffu, s, 4);
  s += __shfl_xor_sync(0xffffffffu, s, 2);
  s += __shfl_xor_sync(0xffffffffu, s, 1);
#pragma unroll
  for (int k = 0; k < RPW; ++k) out[k] = __shfl_sync(0xffffffffu, s, 8 * k);
}

// ---- per-CTA context ----------------------------------------------------------
struct Ctx {
  BicharArgs A;
  int b, len, tid, lane, warp;
  int S, V, T, B, NR, rowf, nmain, prowf, K, DBk, partrows, LEAD;
  float *rows, *part, *prow, *maxpart, *csm, *psp, *ckp;
  int *tab_e, *tab_a, *freel, *phase, *ksm;
  Scal* sc;
  uint64_t* mbar;
  float gsc;
  bool loops, ctx;
};

__device__ __forceinline__ void cta_sync() { asm volatile("bar.sync 0;" ::: "memory"); }

__device__ __forceinline__ int row_r(const Ctx& c, int t) {
  return (int)((((long long)t * c.B + c.b) * c.V) & 3);
}

// Every role calls this between the sweeps (same number of barriers on every path).
// Returns false when there is no backward sweep to run.
__device__ __forceinline__ bool between_sweeps(const Ctx& c) {
  if (!c.A.grad) return false;
  if (c.sc->dead) {  // no accepting path: cost -1e20 (the reference's sentinel), zero gradient
    for (int t = 0; t < c.len; ++t) {
      float* dst = c.A.grad + ((long long)t * c.B + c.b) * c.V;
      for (int j = c.tid; j < c.V; j += blockDim.x) dst[j] = 0.f;
    }
    return false;
  }
  if (c.tid == 0) {
    for (int i = 0; i < c.NR; ++i) {
      mbar_inval(&c.mbar[i]);
      mbar_init(&c.mbar[i], 1);
      c.freel[i] = i;
      c.phase[i] = 0;
    }
    c.sc->nfree = c.NR;
    c.ksm[0] = c.ksm[1] = 0;
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  cta_sync();
  return true;
}

__device__ __forceinline__ void issue_row(const Ctx& c, int t, int row, const float* extra_src, float* extra_dst) {
  const RowGeo g = row_geo(t, c.b, c.B, c.V, c.A.total);
  float* dst = c.rows + row * c.rowf;
  fence_async_smem();
  mbar_expect_tx(&c.mbar[row], g.nbulk * 4 + (extra_src ? c.prowf * 4 : 0));
  bulk_g2s(dst, c.A.x + g.w0, g.nbulk * 4, &c.mbar[row]);
  if (extra_src) bulk_g2s(extra_dst, extra_src, c.prowf * 4, &c.mbar[row]);
  for (int idx = g.nbulk; idx < g.r + c.V; ++idx) dst[idx] = c.A.x[g.w0 + idx];  // <= 3 floats at the tensor's end
}

// e = exp2((x - m) * log2e) in place over the row's aligned window.
__device__ __forceinline__ void exp_row(float* rowp, int nfl, int p, float mL) {
  float4* rp = reinterpret_cast<float4*>(rowp);
  for (int q4 = p; q4 * 4 < nfl; q4 += PTHREADS) {
    float4 v = rp[q4];
    v.x = ex2(fmaf(v.x, kLog2e, -mL));
    v.y = ex2(fmaf(v.y, kLog2e, -mL));
    v.z = ex2(fmaf(v.z, kLog2e, -mL));
    v.w = ex2(fmaf(v.w, kLog2e, -mL));
    rp[q4] = v;
  }
}

// ---- main warps: the serial recursion -----------------------------------------
__device__ __forceinline__ void main_role(const Ctx& c) {
  const int lane = c.lane, warp = c.warp, S = c.S, K = c.K, len = c.len, nmain = c.nmain;
  const int partrows = c.partrows, rowf = c.rowf, prowf = c.prowf;
  // majors M (a strip per warp; warp 0 owns major 0 alone), minors n = lane, lane + 32
  int Mv[RPW];
  bool kval[RPW];
#pragma unroll
  for (int k = 0; k < RPW; ++k) {
    Mv[k] = (warp == 0) ? 0 : 1 + (warp - 1) * RPW + k;
    kval[k] = (warp == 0) ? (k == 0) : (Mv[k] < S);
    if (!kval[k]) Mv[k] = 0;
  }
  const int n0 = lane, n1 = lane + 32;
  const bool nv0 = n0 < S, nv1 = n1 < S;
  const float2 pm2 = make_float2(nv0 ? 1.f : 0.f, nv1 ? 1.f : 0.f);
  const bool loops = c.loops, ctx = c.ctx;

  {  // ------------------------------ forward: majors = contexts c, minors = letters l
    float2 a2[RPW], m2[RPW];
    int eo0[RPW], eo1[RPW];
#pragma unroll
    for (int k = 0; k < RPW; ++k) {
      a2[k] = make_float2(0.f, 0.f);
      const int cc = Mv[k];
      eo0[k] = (kval[k] && nv0 && (n0 != 0 || ctx)) ? cc * S + n0 : 0;  // (c,0) reads the global blank
      eo1[k] = (kval[k] && nv1) ? cc * S + n1 : 0;
      m2[k].x = (loops && kval[k] && nv0 && n0 != 0 && n0 != cc) ? 1.f : 0.f;
      m2[k].y = (loops && kval[k] && nv1 && n1 != cc) ? 1.f : 0.f;
    }
    if (warp == 0 && lane == 0) {
      a2[0].x = pow2i(TARGET);
      c.part[nmain * PARTW + 0] = a2[0].x;  // extra row of buffer 0 carries a[c][0]
    }
    cta_sync();
    for (int g = 0; g <= len + 2; ++g) {
      const int t = g - 2;
      if (t >= 0 && t < len) {
        const float* rowp = c.rows + (t % c.NR) * rowf + row_r(c, t);
        float P[RPW];
        reduce_partials(c.part + (t & 1) * partrows * PARTW, partrows, Mv, lane, P);
        if (lane < RPW) {
          float pv = P[0];
          int mc = Mv[0];
          bool ok = kval[0];
#pragma unroll
          for (int k = 1; k < RPW; ++k) {
            pv = (lane == k) ? P[k] : pv;
            mc = (lane == k) ? Mv[k] : mc;
            ok = (lane == k) ? kval[k] : ok;
          }
          if (ok) c.psp[t * prowf + mc] = pv;
        }
        const int kk = c.ksm[t & 1];
#pragma unroll
        for (int k = 0; k < RPW; ++k) {
          const float2 e2 = make_float2(rowp[eo0[k]], rowp[eo1[k]]);
          const float2 P2 = fmul2(splat2(kval[k] ? P[k] : 0.f), pm2);
          a2[k] = fmul2(e2, ffma2(m2[k], a2[k], P2));
        }
        if (kk != 0) {
          const float2 s2 = splat2(pow2i(kk));
#pragma unroll
          for (int k = 0; k < RPW; ++k) a2[k] = fmul2(a2[k], s2);
        }
        float2 cs = a2[0];
#pragma unroll
        for (int k = 1; k < RPW; ++k) cs = fadd2(cs, a2[k]);
        float* pn = c.part + ((t + 1) & 1) * partrows * PARTW;
        pn[warp * PARTW + n0] = (lane == 0) ? 0.f : cs.x;  // column 0 is not a context sum
        pn[warp * PARTW + n1] = cs.y;
        if (lane == 0) {
#pragma unroll
          for (int k = 0; k < RPW; ++k)
            if (kval[k]) pn[nmain * PARTW + Mv[k]] = a2[k].x;
        }
        if ((t + 1) % K == 0 && t + 1 < len) {
          float* cp = c.ckp + (long long)((t + 1) / K) * c.V;
#pragma unroll
          for (int k = 0; k < RPW; ++k) {
            if (kval[k] && nv0) cp[Mv[k] * S + n0] = a2[k].x;
            if (kval[k] && nv1) cp[Mv[k] * S + n1] = a2[k].y;
          }
        }
      }
      cta_sync();
    }
  }
  if (!between_sweeps(c)) return;

  // -------------------------------- backward: majors = letters l, minors = contexts c
  float2 b2[RPW], f2[RPW], m2[RPW];
  int eo0[RPW], eo1[RPW], vo0[RPW], vo1[RPW];
#pragma unroll
  for (int k = 0; k < RPW; ++k) {
    const int l = Mv[k];
    const bool v0 = kval[k] && nv0, v1 = kval[k] && nv1;
    b2[k] = make_float2(v0 ? pow2i(TARGET) : 0.f, v1 ? pow2i(TARGET) : 0.f);
    f2[k] = make_float2(0.f, 0.f);
    vo0[k] = v0 ? n0 * S + l : 0;
    vo1[k] = v1 ? n1 * S + l : 0;
    eo0[k] = (v0 && (l != 0 || ctx)) ? n0 * S + l : 0;
    eo1[k] = (v1 && (l != 0 || ctx)) ? n1 * S + l : 0;
    m2[k].x = (loops && v0 && l != 0 && l != n0) ? 1.f : 0.f;
    m2[k].y = (loops && v1 && l != 0 && l != n1) ? 1.f : 0.f;
  }
  const int GEND = len + c.LEAD + 2;
  for (int g = 0; g < GEND; ++g) {
    const int i = g - c.LEAD;
    if (i >= 0 && i < len) {
      const int t = len - 1 - i;
      const int slot = q_of_frame(t, len, K) & (QRING - 1);
      const int r = row_r(c, t);
      const float* er = c.rows + (c.tab_e[slot] & 0xff) * rowf + r;
      float* ar = c.rows + c.tab_a[slot] * rowf + r;
      if (i > 0) {  // (a) b_{t+1} = 2^h (Q + loop * f_{t+1}) from the partials of the previous step
        const float* pb = c.part + (i & 1) * partrows * PARTW;
        const float hs = pow2i(c.ksm[i & 1]);
        if (warp == 0) {
          float q0 = 0.f, q1 = 0.f;
          for (int rr = 0; rr < nmain; ++rr) {
            q0 += pb[rr * PARTW + n0];
            q1 += pb[rr * PARTW + n1];
          }
          b2[0] = make_float2(nv0 ? q0 * hs : 0.f, nv1 ? q1 * hs : 0.f);
        } else {
          float Q[RPW];
          reduce_partials(pb, nmain, Mv, lane, Q);
#pragma unroll
          for (int k = 0; k < RPW; ++k) {
            const float2 Q2 = fmul2(splat2(kval[k] ? Q[k] * hs : 0.f), pm2);
            b2[k] = ffma2(fmul2(m2[k], splat2(hs)), f2[k], Q2);
          }
        }
      }
      // (b) posteriors of frame t and f_t = e_t * b_{t+1}
      const float c1 = c.csm[(i & 1) * 2], c2 = c.csm[(i & 1) * 2 + 1];
      float2 cs;
      if (warp == 0) {
        const float* pr = c.prow + slot * prowf;
        const float kt = pow2i((int)pr[S + 1]);
        const float2 e2 = make_float2(er[eo0[0]], er[eo1[0]]);
        const float2 al = fmul2(e2, make_float2(nv0 ? pr[n0] * kt : 0.f, nv1 ? pr[n1] * kt : 0.f));
        f2[0] = fmul2(e2, b2[0]);
        const float2 gm = fmul2(fmul2(al, splat2(c1)), fmul2(b2[0], splat2(c2)));
        cs = f2[0];
        if (ctx) {
          if (nv0) ar[vo0[0]] = gm.x;
          if (nv1) ar[vo1[0]] = gm.y;
        } else {  // one global blank: its posterior is the sum over the S blank states
          const float tot = warp_sum(gm.x + gm.y);
          if (nv0) ar[vo0[0]] = (lane == 0) ? tot : 0.f;
          if (nv1) ar[vo1[0]] = 0.f;
        }
      } else {
        cs = make_float2(0.f, 0.f);
#pragma unroll
        for (int k = 0; k < RPW; ++k) {
          const float2 e2 = make_float2(er[eo0[k]], er[eo1[k]]);
          const float2 al = make_float2(ar[vo0[k]], ar[vo1[k]]);
          f2[k] = fmul2(e2, b2[k]);
          const float2 gm = fmul2(fmul2(al, splat2(c1)), fmul2(b2[k], splat2(c2)));
          if (kval[k] && nv0) ar[vo0[k]] = gm.x;
          if (kval[k] && nv1) ar[vo1[k]] = gm.y;
          cs = fadd2(cs, f2[k]);
        }
      }
      float* pn = c.part + ((i + 1) & 1) * partrows * PARTW;
      pn[warp * PARTW + n0] = cs.x;
      pn[warp * PARTW + n1] = cs.y;
    }
    cta_sync();
  }
}

// ---- producer warps: row maximum and exp in place -------------------------------
__device__ __forceinline__ void prod_role(const Ctx& c) {
  const int pw = c.warp - c.nmain;
  const int p = pw * 32 + c.lane;
  const int len = c.len, S = c.S, V = c.V, NR = c.NR;
  cta_sync();
  for (int g = 0; g <= len + 2; ++g) {
    if (g < len) {  // pass A: maximum of frame g (valid elements only)
      const int row = g % NR;
      mbar_wait(&c.mbar[row], (g / NR) & 1);
      const RowGeo ge = row_geo(g, c.b, c.B, V, c.A.total);
      const float4* rp = reinterpret_cast<const float4*>(c.rows + row * c.rowf);
      const int lo = ge.r, hi = ge.r + V;
      float mx = neg_inf();
      for (int q4 = p; q4 * 4 < ge.nfl; q4 += PTHREADS) {
        const float4 v = rp[q4];
        const int i0 = q4 * 4;
        if (i0 >= lo && i0 + 3 < hi) {
          mx = fmaxf(fmaxf(mx, v.x), fmaxf(v.y, fmaxf(v.z, v.w)));
        } else {
          if (i0 >= lo && i0 < hi) mx = fmaxf(mx, v.x);
          if (i0 + 1 >= lo && i0 + 1 < hi) mx = fmaxf(mx, v.y);
          if (i0 + 2 >= lo && i0 + 2 < hi) mx = fmaxf(mx, v.z);
          if (i0 + 3 >= lo && i0 + 3 < hi) mx = fmaxf(mx, v.w);
        }
      }
      mx = warp_max(mx);
      if (c.lane == 0) c.maxpart[(g & 1) * NPROD + pw] = mx;
    }
    const int f = g - 1;
    if (f >= 0 && f < len) {  // pass B: e = exp(x - max) in place
      const float* mp = c.maxpart + (f & 1) * NPROD;
      float m = fmaxf(fmaxf(mp[0], mp[1]), fmaxf(mp[2], mp[3]));
      if (!(m > neg_inf())) m = 0.f;
      const RowGeo ge = row_geo(f, c.b, c.B, V, c.A.total);
      exp_row(c.rows + (f % NR) * c.rowf, ge.nfl, p, m * kLog2e);
      if (p == 0) {
        c.sc->msum += (double)m;
        c.psp[f * c.prowf + S] = m;
      }
    }
    cta_sync();
  }
  if (!between_sweeps(c)) return;
  const int GEND = len + c.LEAD + 2;
  for (int g = 0; g < GEND; ++g) {
    const int q = g - c.DBk;
    if (q >= 0 && q < len) {
      const int te = c.tab_e[q & (QRING - 1)];
      mbar_wait(&c.mbar[te & 0xff], te >> 8);
      const int tau = frame_of_q(q, len, c.K);
      const RowGeo ge = row_geo(tau, c.b, c.B, V, c.A.total);
      const float mL = c.prow[(q & (QRING - 1)) * c.prowf + S] * kLog2e;
      exp_row(c.rows + (te & 0xff) * c.rowf, ge.nfl, p, mL);
    }
    cta_sync();
  }
}

// ---- recompute warps: a_{t+1} = 2^k e_t (P_t + loop a_t), elementwise ------------
template <int RU>
__device__ __forceinline__ void reco_role(const Ctx& c) {
  const int rw = c.warp - c.nmain - NPROD;
  const int lane = c.lane, len = c.len, S = c.S, V = c.V, K = c.K, T = c.T;
  const int n0 = lane, n1 = lane + 32;
  const bool nv0 = n0 < S, nv1 = n1 < S;
  cta_sync();
  {  // forward sweep: these warps only zero the gradient rows beyond the utterance's end
    const int nz = (c.A.grad && T > len) ? T - len : 0;
    const int zper = (nz + len + 2) / (len + 3);
    for (int g = 0; g <= len + 2; ++g) {
      for (int z = g * zper + rw; z < (g + 1) * zper && z < nz; z += NRECO) {
        float* dst = c.A.grad + ((long long)(len + z) * c.B + c.b) * V;
        for (int j = lane; j < V; j += 32) dst[j] = 0.f;
      }
      cta_sync();
    }
  }
  if (!between_sweeps(c)) return;
  float2 ra[RU], rnx[RU];
#pragma unroll
  for (int u = 0; u < RU; ++u) ra[u] = rnx[u] = make_float2(0.f, 0.f);
  auto load_ckpt = [&](int tau) {
    const float* cp = c.ckp + (long long)(tau / K) * V;
#pragma unroll
    for (int u = 0; u < RU; ++u) {
      const int cc = rw + 4 * u;
      if (cc < S) {
        if (nv0) rnx[u].x = cp[cc * S + n0];
        if (nv1) rnx[u].y = cp[cc * S + n1];
      }
    }
  };
  if (len > 0) {
    const int tau0 = frame_of_q(0, len, K);
    if (tau0 > 0) load_ckpt(tau0);
  }
  const int GEND = len + c.LEAD + 2;
  for (int g = 0; g < GEND; ++g) {
    const int q = g - c.DBk - 1;
    if (q >= 0 && q < len) {
      const int tau = frame_of_q(q, len, K);
      const int r = row_r(c, tau);
      const float* er = c.rows + (c.tab_e[q & (QRING - 1)] & 0xff) * c.rowf + r;
      float* ar = c.rows + c.tab_a[q & (QRING - 1)] * c.rowf + r;
      const float* pr = c.prow + (q & (QRING - 1)) * c.prowf;
      const float scl = pow2i((int)pr[S + 1]);
      if (tau % K == 0) {
        if (tau == 0) {
#pragma unroll
          for (int u = 0; u < RU; ++u) ra[u] = make_float2(0.f, 0.f);
          if (rw == 0 && lane == 0) ra[0].x = pow2i(TARGET);
        } else {
#pragma unroll
          for (int u = 0; u < RU; ++u) ra[u] = rnx[u];
        }
      }
#pragma unroll
      for (int u = 0; u < RU; ++u) {
        const int cc = rw + 4 * u;
        if (cc < S) {
          const float Pc = pr[cc] * scl;
          const float2 e2 = make_float2(nv0 ? er[cc * S + n0] : 0.f, nv1 ? er[cc * S + n1] : 0.f);
          const float2 s2 =
              make_float2((c.loops && n0 != 0 && n0 != cc) ? scl : 0.f, (c.loops && n1 != cc) ? scl : 0.f);
          ra[u] = fmul2(e2, ffma2(s2, ra[u], splat2(Pc)));
          if (nv0) ar[cc * S + n0] = ra[u].x;
          if (nv1) ar[cc * S + n1] = ra[u].y;
        }
      }
      if (q + 1 < len) {  // prefetch the checkpoint that starts the next production segment
        const int tn = frame_of_q(q + 1, len, K);
        if (tn % K == 0 && tn > 0) load_ckpt(tn);
      }
    }
    cta_sync();
  }
}

// ---- scale warp: lagged power-of-two rescaling, final cost, posterior constants ----
__device__ __forceinline__ void scale_role(const Ctx& c) {
  const int lane = c.lane, len = c.len, S = c.S, partrows = c.partrows, nmain = c.nmain, prowf = c.prowf;
  const float inf = __int_as_float(0x7f800000);
  int kcur = 0, Ksum = TARGET;
  cta_sync();
  for (int g = 0; g <= len + 2; ++g) {
    const int t = g - 2;
    if (t >= 0 && t <= len) {
      const float* pb = c.part + (t & 1) * partrows * PARTW;
      float s = 0.f;
      for (int rr = 0; rr < partrows; ++rr) s += pb[rr * PARTW + lane] + pb[rr * PARTW + lane + 32];
      s = warp_sum(s);
      const bool ok = (s > 0.f) && (s < inf);
      if (t < len) {
        const int Knext = Ksum + kcur;  // K_{t+1}: alpha_{t+1} = a_{t+1} 2^-K_{t+1} exp(sum of maxima)
        if (lane == 0) {
          c.psp[t * prowf + S + 1] = (float)kcur;
          c.psp[t * prowf + S + 2] = (float)Knext;
        }
        int knext = 0;
        if (ok) {
          const int ex = exponent_of(s) + kcur;
          if (ex > TARGET + BAND || ex < TARGET - BAND) knext = max(-120, min(120, TARGET - ex));
        } else if (lane == 0) {
          c.sc->dead = 1;
        }
        if (lane == 0) c.ksm[(t + 1) & 1] = knext;
        Ksum = Knext;
        kcur = knext;
      } else if (lane == 0) {
        const bool dead = !ok || c.sc->dead;
        if (dead) c.sc->dead = 1;
        c.A.cost[c.b] = dead ? -1e20 : (log((double)s) - (double)Ksum * kLn2 + c.sc->msum);
        c.sc->inv_mass = dead ? 0.f : 1.f / s;
        c.sc->klen = Ksum;
      }
    }
    cta_sync();
  }
  if (!between_sweeps(c)) return;
  const float inv_mass = c.sc->inv_mass;
  const int klen = c.sc->klen;
  const int E1 = max(-120, min(120, -TARGET - exponent_of(inv_mass)));
  const float c1 = inv_mass * pow2i(E1) * c.gsc;
  int hcur = 0, Hsum = TARGET;
  const int GEND = len + c.LEAD + 2;
  for (int g = 0; g < GEND; ++g) {
    const int i = g - c.LEAD;
    const int ii = i + 1;  // decisions for main step ii (frame len-1-ii)
    if (ii >= 0 && ii < len) {
      int hnext = 0;
      if (i >= 1) {
        const float* pb = c.part + (i & 1) * partrows * PARTW;
        float s = 0.f;
        for (int rr = 0; rr < nmain; ++rr) s += pb[rr * PARTW + lane] + pb[rr * PARTW + lane + 32];
        s = warp_sum(s);
        if ((s > 0.f) && (s < inf)) {
          const int ex = exponent_of(s) + hcur;
          if (ex > TARGET + BAND || ex < TARGET - BAND) hnext = max(-120, min(120, TARGET - ex));
        }
      }
      Hsum += hnext;
      hcur = hnext;
      const int slot = q_of_frame(len - 1 - ii, len, c.K) & (QRING - 1);
      const int te = c.tab_e[slot];
      mbar_wait(&c.mbar[te & 0xff], te >> 8);
      const int EA = klen - (int)c.prow[slot * prowf + S + 2];
      const int E2 = max(-126, min(126, EA - Hsum - E1));
      if (lane == 0) {
        c.ksm[ii & 1] = hnext;
        c.csm[(ii & 1) * 2] = c1;
        c.csm[(ii & 1) * 2 + 1] = pow2i(E2);
      }
    }
    cta_sync();
  }
}

// ---- io warp: bulk loads, bulk stores of finished gradient rows, row pool ----------
__device__ __forceinline__ void io_role(const Ctx& c) {
  const int lane = c.lane, len = c.len, V = c.V, NR = c.NR, K = c.K;
  if (lane == 0)
    for (int f = 0; f < DF && f < len; ++f) issue_row(c, f, f % NR, nullptr, nullptr);
  cta_sync();
  for (int g = 0; g <= len + 2; ++g) {
    const int f = g + DF;
    if (lane == 0 && f < len) issue_row(c, f, f % NR, nullptr, nullptr);
    cta_sync();
  }
  if (!between_sweeps(c)) return;
  auto pop_row = [&]() -> int {
    const int nf = c.sc->nfree;
    if (nf <= 0) {
      printf("ctcg bichar: row pool exhausted (b=%d)\n", c.b);
      __trap();
    }
    c.sc->nfree = nf - 1;
    return c.freel[nf - 1];
  };
  auto push_row = [&](int row) {
    const int nf = c.sc->nfree;
    c.freel[nf] = row;
    c.sc->nfree = nf + 1;
  };
  const int GEND = len + c.LEAD + 2;
  for (int g = 0; g < GEND; ++g) {
    const int i = g - c.LEAD;  // main step of this global step; its frame is len-1-i
    // 1. the logits row consumed by main step i-1 is free again
    if (lane == 0 && i - 1 >= 0 && i - 1 < len) push_row(c.tab_e[q_of_frame(len - i, len, K) & (QRING - 1)] & 0xff);
    // 2. store the gradient row main step i-1 finished
    if (i - 1 >= 0 && i - 1 < len) {
      const int t = len - i;
      const RowGeo ge = row_geo(t, c.b, c.B, V, c.A.total);
      const float* src = c.rows + c.tab_a[q_of_frame(t, len, K) & (QRING - 1)] * c.rowf;
      float* dst = c.A.grad + ge.w0;
      const int i0 = ge.r ? 4 : 0, i1 = (ge.r + V) & ~3;
      if (lane == 0) {
        fence_async_smem();
        if (i1 > i0) bulk_s2g(dst + i0, src + i0, (i1 - i0) * 4);
      }
      // head / tail floats share a 16-byte chunk with a neighbouring row: plain stores
      if (lane >= 8 && lane < 12) {
        const int idx = ge.r + (lane - 8);
        if (idx < i0 && idx < ge.r + V) dst[idx] = src[idx];
      } else if (lane >= 16 && lane < 20) {
        const int idx = max(i1, i0) + (lane - 16);
        if (idx < ge.r + V) dst[idx] = src[idx];
      }
    }
    if (lane == 0) {
      bulk_commit();
      bulk_wait_read<1>();  // 3. the row stored one step earlier has been read out of shared memory
      if (i - 2 >= 0 && i - 2 < len) push_row(c.tab_a[q_of_frame(len + 1 - i, len, K) & (QRING - 1)]);
      if (g < len) {  // 4. fetch logits row + P row of production index g
        const int row = pop_row();
        const int ph = c.phase[row];
        c.phase[row] = ph ^ 1;
        c.tab_e[g & (QRING - 1)] = row | (ph << 8);
        const int tau = frame_of_q(g, len, K);
        issue_row(c, tau, row, c.psp + tau * c.prowf, c.prow + (g & (QRING - 1)) * c.prowf);
      }
      const int qa = g - c.DBk;  // 5. row that receives the recomputed alpha of production index g - DB
      if (qa >= 0 && qa < len) c.tab_a[qa & (QRING - 1)] = pop_row();
    }
    cta_sync();
  }
  if (lane == 0) bulk_wait_all<0>();
}

// ---- the kernel ---------------------------------------------------------------
template <int RU, int MAXT>
__global__ void __launch_bounds__(MAXT, 1) bichar_den_kernel(const BicharArgs A) {
  extern __shared__ __align__(128) unsigned char smraw[];
  Ctx c;
  c.A = A;
  c.tid = threadIdx.x;
  c.lane = c.tid & 31;
  c.warp = c.tid >> 5;
  c.b = blockIdx.x;
  c.S = A.S; c.V = A.V; c.T = A.T; c.B = A.B; c.NR = A.NR; c.rowf = A.rowf; c.nmain = A.nmain;
  c.prowf = A.prowf; c.K = A.K; c.DBk = A.DB;
  c.partrows = A.nmain + 1;
  c.LEAD = A.DB + 1 + A.K;
  c.len = min(A.lens[c.b], A.T);
  const Smem L = smem_plan(A.NR, A.rowf, A.nmain, A.prowf);
  c.rows = reinterpret_cast<float*>(smraw + L.rows);
  c.part = reinterpret_cast<float*>(smraw + L.part);
  c.prow = reinterpret_cast<float*>(smraw + L.prow);
  c.maxpart = reinterpret_cast<float*>(smraw + L.maxpart);
  c.tab_e = reinterpret_cast<int*>(smraw + L.tab_e);
  c.tab_a = reinterpret_cast<int*>(smraw + L.tab_a);
  c.freel = reinterpret_cast<int*>(smraw + L.freel);
  c.phase = reinterpret_cast<int*>(smraw + L.phase);
  c.ksm = reinterpret_cast<int*>(smraw + L.ksm);
  c.csm = reinterpret_cast<float*>(smraw + L.csm);
  c.sc = reinterpret_cast<Scal*>(smraw + L.scal);
  c.mbar = reinterpret_cast<uint64_t*>(smraw + L.mbar);
  c.gsc = A.g_scale ? A.g_scale[c.b] : 1.f;
  c.psp = A.pspill + (long long)c.b * A.T * A.prowf;
  c.ckp = A.ckpt + (long long)c.b * A.nck * A.V;
  c.loops = A.loops != 0;
  c.ctx = A.ctx != 0;

  for (int i = c.tid; i < 2 * c.partrows * PARTW; i += blockDim.x) c.part[i] = 0.f;
  if (c.tid == 0) {
    for (int i = 0; i < A.NR; ++i) mbar_init(&c.mbar[i], 1);
    c.ksm[0] = c.ksm[1] = 0;
    c.sc->msum = 0.0;
    c.sc->dead = 0;
    c.sc->inv_mass = 0.f;
    c.sc->klen = TARGET;
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  cta_sync();
  const int w = c.warp, nm = A.nmain;
  if (w < nm) main_role(c);
  else if (w < nm + NPROD) prod_role(c);
  else if (w < nm + NPROD + NRECO) reco_role<RU>(c);
  else if (w == nm + NPROD + NRECO) scale_role(c);
  else io_role(c);
}

}  // namespace

// ---- host side ------------------------------------------------------------------
BicharPlan bichar_plan(int T, int B, int V, int S) {
  BicharPlan p{};
  p.ok = 0;
  if (S < 4 || S > 64 || V != S * S || T < 1 || B < 1) return p;
  p.nmain = 1 + (S - 1 + RPW - 1) / RPW;
  p.threads = (p.nmain + NPROD + NRECO + 2) * 32;
  p.rowf = pad4(V + 3);
  p.prowf = pad4(S + 3);
  p.ru = (S + 3) / 4;
  if (p.ru > 13 || p.threads > 736) return p;  // S <= 49 with the kernels instantiated below
  const int budget = 227 * 1024;
  int NR = 0;
  for (int n = 40; n >= 12; --n) {
    if (smem_plan(n, p.rowf, p.nmain, p.prowf).bytes <= budget) {
      NR = n;
      break;
    }
  }
  if (NR < 12) return p;
  // rows needed by the backward sweep: 2K + DB + 5
  p.DB = (NR >= 25) ? 4 : 3;
  p.K = (NR - p.DB - 5) / 2;
  if (p.K > 8) p.K = 8;
  if (p.K < 2 || NR < DF + 3) return p;
  p.NR = NR;
  p.nck = T / p.K + 1;
  p.smem_bytes = smem_plan(NR, p.rowf, p.nmain, p.prowf).bytes;
  p.pspill_bytes = (size_t)B * T * p.prowf * 4;
  p.ckpt_bytes = (size_t)B * p.nck * V * 4;
  p.ok = 1;
  return p;
}

cudaError_t launch_bichar_den(const BicharArgs& a, const BicharPlan& p, cudaStream_t st) {
  ProfScope prof("bichar_den", st);
  cudaError_t e = cudaSuccess;
  if (p.ru <= 9) {
    e = cudaFuncSetAttribute(bichar_den_kernel<9, 736>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                             (int)p.smem_bytes);
    if (e != cudaSuccess) return e;
    bichar_den_kernel<9, 736><<<a.B, p.threads, p.smem_bytes, st>>>(a);
  } else {
    e = cudaFuncSetAttribute(bichar_den_kernel<13, 736>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                             (int)p.smem_bytes);
    if (e != cudaSuccess) return e;
    bichar_den_kernel<13, 736><<<a.B, p.threads, p.smem_bytes, st>>>(a);
  }
  ++g_launches;
  return cudaGetLastError();
}

}  // namespace ctcg
