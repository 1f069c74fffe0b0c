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
#include <cstdlib>
#include <vector>

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
constexpr int TRS = 8;        // clock stamps per warp and step in a CTCG_BICHAR_TRACE capture

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
struct ProdCursor {
  int tau, u, size, K;
  __device__ __forceinline__ void init(int len, int K_) {
    K = K_;
    const int J = (len + K - 1) / K;
    size = len - (J - 1) * K;
    tau = (J - 1) * K;
    u = 0;
  }
  __device__ __forceinline__ void next() {
    ++u;
    ++tau;
    if (u == size) {
      tau -= size + K;
      u = 0;
      size = K;
    }
  }
};
// Consumption order t = len-1, len-2, ...: q is the production index of frame t.
struct ConsCursor {
  int q, u, base, size, K;
  __device__ __forceinline__ void init(int len, int K_) {
    K = K_;
    const int J = (len + K - 1) / K;
    size = len - (J - 1) * K;
    base = 0;
    u = size - 1;
    q = u;
  }
  __device__ __forceinline__ void prev() {
    if (u > 0) {
      --u;
      --q;
    } else {
      base += size;
      size = K;
      u = K - 1;
      q = base + K - 1;
    }
  }
};

// Keep a value in a register: the compiler may not rematerialise it inside the hot loops.
__device__ __forceinline__ void pin(int& v) { asm volatile("" : "+r"(v)); }
__device__ __forceinline__ void pin(unsigned& v) { asm volatile("" : "+r"(v)); }
__device__ __forceinline__ void pin(float& v) { asm volatile("" : "+f"(v)); }
__device__ __forceinline__ void pin(float2& v) {
  pin(v.x);
  pin(v.y);
}

// ---- per-CTA context ----------------------------------------------------------
struct Ctx {
  BicharArgs A;
  int b, len, tid, lane, warp;
  int S, V, T, B, NR, rowf, nmain, prowf, K, DBk, partrows, LEAD;
  int B3, b3, V3;  // (B, b, V) mod 4: row alignment arithmetic
  float *rows, *part, *prow, *maxpart, *csm, *psp, *ckp;
  int *tab_e, *tab_a, *freel, *phase, *ksm;
  Scal* sc;
  uint64_t* mbar;
  float gsc;
  bool loops, ctx;
  long long* trace;  // optional per-step timeline of CTA 0 (CTCG_BICHAR_TRACE), else null
  int trace_steps;
};

// Compile-time sizes when the kernel is instantiated for one S (SCT > 0), run-time otherwise.
template <int SCT>
struct Dims {
  int S, V, nmain, partrows, rowf, prowf;
  __device__ __forceinline__ explicit Dims(const Ctx& c) {
    if (SCT > 0) {
      S = SCT;
      V = SCT * SCT;
      nmain = 1 + (SCT - 1 + RPW - 1) / RPW;
      rowf = (SCT * SCT + 3 + 3) & ~3;
      prowf = (SCT + 3 + 3) & ~3;
    } else {
      S = c.S;
      V = c.V;
      nmain = c.nmain;
      rowf = c.rowf;
      prowf = c.prowf;
    }
    partrows = nmain + 1;
  }
};

__device__ __forceinline__ void cta_sync() { asm volatile("bar.sync 0;" ::: "memory"); }

// Barrier of the per-frame loops.  With tracing on, lane 0 of every warp of CTA 0 stamps
// the clock when its work of the step ends (before the barrier) and when the barrier releases.
__device__ __forceinline__ void step_sync(const Ctx& c, int& n) {
#ifdef CTCG_BICHAR_TRACE_BUILD
  if (c.trace) {
    const bool rec = c.b == 0 && c.lane == 0 && n < c.trace_steps;
    long long* tr = c.trace + ((long long)c.warp * c.trace_steps + (rec ? n : 0)) * TRS;
    if (rec) tr[0] = clock64();
    cta_sync();  // one barrier instruction for the whole warp: no divergence around it
    if (rec) tr[1] = clock64();
    ++n;
    return;
  }
#endif
  cta_sync();
}

// Intermediate stamp `slot` (2..TRS-1) of the current step, taken once `dep` is available.
// Compiled in only with -DCTCG_BICHAR_TRACE_BUILD (the stamps cost ~3 % of the instructions).
__device__ __forceinline__ void tstamp(const Ctx& c, int n, int slot, float dep) {
#ifdef CTCG_BICHAR_TRACE_BUILD
  if (c.trace && c.b == 0 && c.lane == 0 && n < c.trace_steps) {
    long long t;
    asm volatile("{ .reg .f32 d; mov.f32 d, %1; mov.u64 %0, %%clock64; }" : "=l"(t) : "f"(dep));
    c.trace[((long long)c.warp * c.trace_steps + n) * TRS + slot] = t;
  }
#endif
}

// Offset (floats) of row (t, b) inside its 16-byte aligned window: ((t*B + b) * V) mod 4.
__device__ __forceinline__ int row_r(const Ctx& c, int t) { return ((((t & 3) * c.B3 + c.b3) & 3) * c.V3) & 3; }
// The same in bytes for consecutive frames: r4(t + 1) = (r4(t) + d4) & 12 with d4 = 4 * ((B * V) mod 4).
__device__ __forceinline__ unsigned row_d4(const Ctx& c) { return (unsigned)((c.B3 * c.V3) & 3) * 4u; }

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

// In-warp reduction of the partial-sum rows: lanes 8k..8k+7 add the partials of the
// group's column (at most 24 rows); returns the group's total in every lane of the group.
__device__ __forceinline__ float group_sum(const float* pcol, int nrows, int j8) {
  float s = (j8 < nrows) ? pcol[j8 * PARTW] : 0.f;
  if (j8 + 8 < nrows) s += pcol[(j8 + 8) * PARTW];
  if (j8 + 16 < nrows) s += pcol[(j8 + 16) * PARTW];
  s += __shfl_xor_sync(0xffffffffu, s, 4);
  s += __shfl_xor_sync(0xffffffffu, s, 2);
  s += __shfl_xor_sync(0xffffffffu, s, 1);
  return s;
}

// Shared-memory accesses through 32-bit shared-space addresses (one IADD per access,
// immediates folded by ptxas); volatile keeps them ordered against the barriers.
__device__ __forceinline__ float lds(unsigned a) {
  float v;
  asm volatile("ld.shared.f32 %0, [%1];" : "=f"(v) : "r"(a));
  return v;
}
__device__ __forceinline__ float4 lds4(unsigned a) {
  float4 v;
  asm volatile("ld.shared.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(a));
  return v;
}
__device__ __forceinline__ void sts(unsigned a, float v) { asm volatile("st.shared.f32 [%0], %1;" ::"r"(a), "f"(v)); }

// e = exp(x - m) for a pair: exp2(x * log2e - m * log2e)
__device__ __forceinline__ float2 exp_pair(float x0, float x1, float nmL) {
  const float2 y = ffma2(make_float2(x0, x1), splat2(kLog2e), splat2(nmL));
  return make_float2(ex2(y.x), ex2(y.y));
}

// Same with a shared-space address of the lane's first partial (row j8 of the group's column).
__device__ __forceinline__ float group_sum_a(unsigned a, int nrows, int j8) {
  float s = (j8 < nrows) ? lds(a) : 0.f;
  if (j8 + 8 < nrows) s += lds(a + 8 * PARTW * 4);
  if (j8 + 16 < nrows) s += lds(a + 16 * PARTW * 4);
  s += __shfl_xor_sync(0xffffffffu, s, 4);
  s += __shfl_xor_sync(0xffffffffu, s, 2);
  s += __shfl_xor_sync(0xffffffffu, s, 1);
  return s;
}

// ---- main warps: the serial recursion -----------------------------------------
template <int SCT>
__device__ __forceinline__ void main_role(const Ctx& c) {
  const Dims<SCT> D(c);
  int nstep = 0;
  const int S = D.S, nmain = D.nmain, partrows = D.partrows, rowf = D.rowf, prowf = D.prowf;
  const int lane = c.lane, warp = c.warp, K = c.K, len = c.len, NR = c.NR;
  // majors M0..M0+nk-1 (a strip per warp; warp 0 owns major 0 alone), minors n = lane, lane + 32
  const int M0 = (warp == 0) ? 0 : 1 + (warp - 1) * RPW;
  const int nk = (warp == 0) ? 1 : min(RPW, S - M0);
  const int n0 = lane, n1 = lane + 32;
  const bool nv0 = n0 < S, nv1 = n1 < S;
  const bool loops = c.loops, ctx = c.ctx;
  const int grp = lane >> 3, j8 = lane & 7;
  const int gcol = (grp < nk) ? M0 + grp : 0;  // column this lane group reduces
  const unsigned rows_a = smem_u32(c.rows), part_a = smem_u32(c.part);
  const unsigned rowb = rowf * 4, partb = partrows * PARTW * 4;

  {  // ------------------------------ forward: majors = contexts c, minors = letters l
    float2 a2[RPW], m2[RPW];
    unsigned a0[RPW], a1[RPW];  // byte offsets of the element's label inside the logits row
#pragma unroll
    for (int k = 0; k < RPW; ++k) {
      a2[k] = make_float2(0.f, 0.f);
      const bool kv = k < nk;
      const int cc = kv ? M0 + k : 0;
      a0[k] = (nv0 && (n0 != 0 || ctx)) ? (cc * S + n0) * 4 : 0;  // (c,0) reads the global blank
      a1[k] = (cc * S + n1) * 4;  // an invalid slot reads past the row (inside shared memory); never used
      m2[k].x = (loops && kv && nv0 && n0 != 0 && n0 != cc) ? 1.f : 0.f;
      m2[k].y = (loops && kv && nv1 && n1 != cc) ? 1.f : 0.f;
      pin(m2[k]);
      pin(a0[k]);
      pin(a1[k]);
    }
    float col0 = (lane == 0) ? 0.f : 1.f;  // column 0 is not a context sum
    int pspo = ((j8 == 0) && grp < nk) ? M0 + grp : -1;
    float gmask = (grp < nk) ? 1.f : 0.f;
    unsigned gaddr = part_a + (gcol + j8 * PARTW) * 4;           // this lane's first partial of its group column
    unsigned paddr = part_a + (warp * PARTW + n0) * 4;           // this warp's partial row, slot 0
    unsigned xaddr = part_a + (nmain * PARTW + M0) * 4;          // extra row entries of this warp
    pin(col0);
    pin(pspo);
    pin(gmask);
    pin(gaddr);
    pin(paddr);
    pin(xaddr);
    if (warp == 0 && lane == 0) {
      a2[0].x = pow2i(TARGET);
      c.part[nmain * PARTW + 0] = a2[0].x;  // extra row of buffer 0 carries a[c][0]
    }
    int ck = K;            // frames until the next checkpoint
    unsigned rowoff = 0;   // byte offset of the ring row being consumed
    unsigned r4 = row_r(c, 0) * 4;  // byte offset of the row inside its aligned window
    const unsigned d4 = row_d4(c);
    float* pspw = c.psp + (pspo >= 0 ? pspo : 0);  // this lane's slot of the P spill row of frame t
    const unsigned ksm_a = smem_u32(c.ksm), maxp_a = smem_u32(c.maxpart);
    cta_sync();
    for (int g = 0; g <= len + 1; ++g) {
      const int t = g - 1;
      if (t >= 0 && t < len && !(c.A.dbg & 4)) {
        const unsigned base = rows_a + rowoff + r4;
        r4 = (r4 + d4) & 12u;
        rowoff += rowb;
        if (rowoff == NR * rowb) rowoff = 0;
        const float4 mp = lds4(maxp_a + (t & 1) * 16);
        int kk;
        asm volatile("ld.shared.s32 %0, [%1];" : "=r"(kk) : "r"(ksm_a + (t & 1) * 4));
        float m = fmaxf(fmaxf(mp.x, mp.y), fmaxf(mp.z, mp.w));
        if (!(m > neg_inf())) m = 0.f;
        // a rescale by 2^kk rides in the exponent of e: exp2(x*log2e - m*log2e + kk)
        const float nmL = fmaf(m, -kLog2e, (float)kk);
        tstamp(c, nstep, 2, nmL);
        float xr0[RPW], xr1[RPW];
#pragma unroll
        for (int k = 0; k < RPW; ++k) {
          xr0[k] = lds(base + a0[k]);
          xr1[k] = lds(base + a1[k]);
        }
        tstamp(c, nstep, 3, xr0[0] + xr1[RPW - 1]);
        const float s = group_sum_a(gaddr + (t & 1) * partb, partrows, j8) * gmask;
        tstamp(c, nstep, 4, s);
        float P[RPW];
#pragma unroll
        for (int k = 0; k < RPW; ++k) P[k] = __shfl_sync(0xffffffffu, s, 8 * k);
        if (pspo >= 0) *pspw = s;
        pspw += prowf;
#pragma unroll
        for (int k = 0; k < RPW; ++k) {
          const float2 e2 = exp_pair(xr0[k], xr1[k], nmL);
          a2[k] = fmul2(e2, ffma2(m2[k], a2[k], splat2(P[k])));
        }
        float2 cs = a2[0];
#pragma unroll
        for (int k = 1; k < RPW; ++k) cs = fadd2(cs, a2[k]);
        tstamp(c, nstep, 5, cs.x + cs.y);
        const unsigned pno = ((t + 1) & 1) * partb;
        sts(paddr + pno, cs.x * col0);
        sts(paddr + pno + 128, cs.y);
        if (lane == 0) {
          const unsigned ex = xaddr + pno;
          sts(ex, a2[0].x);
          if (nk > 1) sts(ex + 4, a2[1].x);
          if (nk > 2) sts(ex + 8, a2[2].x);
          if (nk > 3) sts(ex + 12, a2[3].x);
        }
        if (--ck == 0) {
          ck = K;
          if (t + 1 < len) {
            float* cp = c.ckp + (long long)((t + 1) / K) * D.V + M0 * S;
#pragma unroll
            for (int k = 0; k < RPW; ++k) {
              if (k < nk && nv0) cp[k * S + n0] = a2[k].x;
              if (k < nk && nv1) cp[k * S + n1] = a2[k].y;
            }
          }
        }
      }
      step_sync(c, nstep);
    }
  }
  if (!between_sweeps(c)) return;

  // -------------------------------- backward: majors = letters l, minors = contexts c
  // Element (c, l) of a row sits at byte (c*S + l)*4: with c = lane (+32) and l = M0 + k the four
  // majors of a lane are consecutive floats, so one base address per slot serves all k as an
  // immediate offset.  Warp 0 (l = 0) reads the one global blank at offset 0 unless blanks are
  // contextual.  FULL: every warp but warp 0 owns exactly RPW letters (S = 49, 37, ...).
  constexpr bool FULL = (SCT > 0) && ((SCT - 1) % RPW == 0);
  constexpr bool WIDE = (SCT >= 32);  // slot 0 (c = lane) is always a valid context
  float2 b2[RPW], f2[RPW], m2[RPW];
#pragma unroll
  for (int k = 0; k < RPW; ++k) {
    const bool kv = k < nk;
    const int l = kv ? M0 + k : 0;
    const bool v0 = kv && nv0, v1 = kv && nv1;
    b2[k] = make_float2(v0 ? pow2i(TARGET) : 0.f, v1 ? pow2i(TARGET) : 0.f);
    f2[k] = make_float2(0.f, 0.f);
    m2[k].x = (loops && v0 && l != 0 && l != n0) ? 1.f : 0.f;
    m2[k].y = (loops && v1 && l != 0 && l != n1) ? 1.f : 0.f;
    pin(m2[k]);
  }
  // invalid slots read element 0 of the row (finite after the exp pass); their products are masked
  unsigned vb0 = (nv0 ? n0 * S + M0 : 0) * 4, vb1 = (nv1 ? n1 * S + M0 : 0) * 4;
  unsigned eb0 = (warp == 0 && !ctx) ? 0u : vb0, eb1 = (warp == 0 && !ctx) ? 0u : vb1;
  const bool st0 = WIDE ? true : nv0, st1 = nv1;
  float2 pm = make_float2(nv0 ? 1.f : 0.f, nv1 ? 1.f : 0.f);
  float gmask = (grp < nk) ? 1.f : 0.f;
  unsigned gaddr = part_a + (gcol + j8 * PARTW) * 4;
  unsigned paddr = part_a + (warp * PARTW + n0) * 4;
  pin(vb0);
  pin(vb1);
  pin(eb0);
  pin(eb1);
  pin(pm);
  pin(gmask);
  pin(gaddr);
  pin(paddr);
  ConsCursor cur;
  cur.init(max(len, 1), K);
  const int GEND = len + c.LEAD + 2;
  const unsigned d4 = row_d4(c);
  unsigned r4 = row_r(c, max(len - 1, 0)) * 4;
  const unsigned tabe_a = smem_u32(c.tab_e), taba_a = smem_u32(c.tab_a), ksm_a = smem_u32(c.ksm),
                 csm_a = smem_u32(c.csm);
  for (int g = 0; g < GEND; ++g) {
    const int i = g - c.LEAD;
    if (i >= 0 && i < len && !(c.A.dbg & 2)) {
      const int slot = cur.q & (QRING - 1);
      cur.prev();
      int te, ta, hk;
      asm volatile("ld.shared.s32 %0, [%1];" : "=r"(te) : "r"(tabe_a + slot * 4));
      asm volatile("ld.shared.s32 %0, [%1];" : "=r"(ta) : "r"(taba_a + slot * 4));
      asm volatile("ld.shared.s32 %0, [%1];" : "=r"(hk) : "r"(ksm_a + (i & 1) * 4));
      const unsigned er = rows_a + (te & 0xff) * rowb + r4;
      const unsigned ar = rows_a + ta * rowb + r4;
      r4 = (r4 - d4) & 12u;  // frame t - 1 comes next
      const unsigned e0 = er + eb0, e1 = er + eb1, q0a = ar + vb0, q1a = ar + vb1;
      float2 e2[RPW], al[RPW];
#pragma unroll
      for (int k = 0; k < RPW; ++k) {  // all loads first: they alias the stores below
        // an absent major (k >= nk) re-reads major 0 of the warp: in bounds, result unused
        const unsigned ko = (FULL || k < nk) ? 4u * k : 0u;
        e2[k] = make_float2(lds(e0 + ko), lds(e1 + ko));
        al[k] = make_float2(lds(q0a + ko), lds(q1a + ko));
      }
      if (i > 0) {  // (a) b_{t+1} = 2^h (Q + loop * f_{t+1}) from the partials of the previous step
        const float hs = pow2i(hk);
        if (warp == 0) {
          const float* pb = c.part + (i & 1) * partrows * PARTW;
          float q0 = 0.f, q1 = 0.f;
#pragma unroll 4
          for (int rr = 0; rr < nmain; ++rr) {
            q0 += pb[rr * PARTW + n0];
            q1 += pb[rr * PARTW + n1];
          }
          b2[0] = fmul2(make_float2(q0 * hs, q1 * hs), pm);
        } else {
          const float s = group_sum_a(gaddr + (i & 1) * partb, nmain, j8) * (hs * gmask);
          const float2 h2 = splat2(hs);
#pragma unroll
          for (int k = 0; k < RPW; ++k) {
            const float Qk = __shfl_sync(0xffffffffu, s, 8 * k);
            b2[k] = ffma2(fmul2(m2[k], h2), f2[k], splat2(Qk));
          }
        }
      }
      // (b) posteriors of frame t and f_t = e_t * b_{t+1}
      float2 c12;
      asm volatile("ld.shared.v2.f32 {%0,%1}, [%2];" : "=f"(c12.x), "=f"(c12.y) : "r"(csm_a + (i & 1) * 8));
      float2 cs;
      if (warp == 0) {
        const float* pr = c.prow + slot * prowf;
        const float kt = pow2i((int)pr[S + 1]);
        const float2 a_ = fmul2(fmul2(e2[0], make_float2(pr[nv0 ? n0 : 0] * kt, pr[nv1 ? n1 : 0] * kt)), pm);
        f2[0] = fmul2(e2[0], b2[0]);
        const float2 gm = fmul2(fmul2(a_, splat2(c12.x)), fmul2(b2[0], splat2(c12.y)));
        cs = f2[0];
        if (ctx) {
          if (nv0) sts(q0a, gm.x);
          if (nv1) sts(q1a, gm.y);
        } else {  // one global blank: its posterior is the sum over the S blank states
          const float tot = warp_sum(gm.x + gm.y);
          if (nv0) sts(q0a, (lane == 0) ? tot : 0.f);
          if (nv1) sts(q1a, 0.f);
        }
      } else {
        cs = make_float2(0.f, 0.f);
        const float2 c2 = splat2(c12.y);  // c1 is already in the alpha row (auxiliary warps)
#pragma unroll
        for (int k = 0; k < RPW; ++k) {
          f2[k] = fmul2(e2[k], b2[k]);
          al[k] = fmul2(al[k], fmul2(b2[k], c2));
          cs = fadd2(cs, f2[k]);
        }
#pragma unroll
        for (int k = 0; k < RPW; ++k) {
          if (FULL || k < nk) {
            if (st0) sts(q0a + 4 * k, al[k].x);
            if (st1) sts(q1a + 4 * k, al[k].y);
          }
        }
      }
      const unsigned pno = ((i + 1) & 1) * partb;
      sts(paddr + pno, cs.x);
      sts(paddr + pno + 128, cs.y);
    }
    step_sync(c, nstep);
  }
}

// ---- auxiliary warps ------------------------------------------------------------
// forward sweep : warps 0..3 take the maximum of the frame that the main warps consume in
//                 the next step; warps 4..7 zero the gradient rows beyond the utterance's end.
// backward sweep: all 8 turn a landed logits row into e = exp(x - max) (in place) and the
//                 recomputed a_{t+1} = 2^k e_t (P_t + loop a_t) (own row), elementwise.
template <int SCT, int RU>
__device__ __forceinline__ void aux_role(const Ctx& c) {
  constexpr int NAUX = NPROD + NRECO;
  const Dims<SCT> D(c);
  int nstep = 0;
  const int aw = c.warp - D.nmain;
  const int lane = c.lane, len = c.len, S = D.S, V = D.V, K = c.K, NR = c.NR, rowf = D.rowf, prowf = D.prowf;
  const unsigned rows_a = smem_u32(c.rows), rowb = rowf * 4;
  cta_sync();
  if (aw < NPROD) {
    constexpr int NQ = (640 + PTHREADS - 1) / PTHREADS;  // rows hold at most 640 float4 (V <= 2500)
    const int p = aw * 32 + lane;
    int rowA = 0, parA = 0;  // ring row / mbarrier parity of the frame
    for (int g = 0; g <= len + 1; ++g) {
      if (g < len && !(c.A.dbg & 8)) {
        mbar_wait(&c.mbar[rowA], parA);
        const int r = row_r(c, g);
        const unsigned rp = rows_a + rowA * rowb;
        const int hi = r + V, nq = (hi + 3) >> 2;
        float4 v[NQ];
#pragma unroll
        for (int j = 0; j < NQ; ++j)
          if (p + j * PTHREADS < nq) v[j] = lds4(rp + (p + j * PTHREADS) * 16);
        float mx = neg_inf();
#pragma unroll
        for (int j = 0; j < NQ; ++j) {
          const int i0 = (p + j * PTHREADS) * 4;
          if (p + j * PTHREADS < nq) {
            if (i0 >= r && i0 + 3 < hi) {
              mx = fmaxf(fmaxf(mx, v[j].x), fmaxf(v[j].y, fmaxf(v[j].z, v[j].w)));
            } else {  // first / last float4 of the window: mask the neighbours' floats
              if (i0 >= r && i0 < hi) mx = fmaxf(mx, v[j].x);
              if (i0 + 1 >= r && i0 + 1 < hi) mx = fmaxf(mx, v[j].y);
              if (i0 + 2 >= r && i0 + 2 < hi) mx = fmaxf(mx, v[j].z);
              if (i0 + 3 >= r && i0 + 3 < hi) mx = fmaxf(mx, v[j].w);
            }
          }
        }
        mx = warp_max(mx);
        if (lane == 0) c.maxpart[(g & 1) * 4 + aw] = mx;
        if (++rowA == NR) {
          rowA = 0;
          parA ^= 1;
        }
      }
      step_sync(c, nstep);
    }
  } else {
    const int zw = aw - NPROD;
    const int nz = (c.A.grad && c.T > len) ? c.T - len : 0;
    const int zper = (nz + len + 1) / (len + 2);
    for (int g = 0; g <= len + 1; ++g) {
      for (int z = g * zper + zw; z < (g + 1) * zper && z < nz; z += NRECO) {
        float* dst = c.A.grad + ((long long)(len + z) * c.B + c.b) * V;
        for (int j = lane; j < V; j += 32) dst[j] = 0.f;
      }
      step_sync(c, nstep);
    }
  }
  if (!between_sweeps(c)) return;

  const int n0 = lane, n1 = lane + 32;
  const bool nv0 = n0 < S, nv1 = n1 < S;
  // posterior constant c1 = gscale / mass * 2^E1 (the scale warp's formula): folded into the alpha
  // rows here so that the main warps spend one multiply less per element
  const float inv_mass_ = c.sc->inv_mass;
  const float2 c1v = splat2(inv_mass_ * pow2i(max(-120, min(120, -TARGET - exponent_of(inv_mass_)))) * c.gsc);
  float2 ra[RU], rnx[RU];
#pragma unroll
  for (int u = 0; u < RU; ++u) ra[u] = rnx[u] = make_float2(0.f, 0.f);
  unsigned o0 = (aw * S + (nv0 ? n0 : 0)) * 4;  // element offsets of row aw; row aw + 8u adds 8u*S
  unsigned o1 = (aw * S + (nv1 ? n1 : (nv0 ? n0 : 0))) * 4;
  pin(o0);
  pin(o1);
  // explicit self-loop of state (c, l): l != 0 and l != c (the diagonal loops through P)
  float2 lm[RU];
#pragma unroll
  for (int u = 0; u < RU; ++u) {
    const int cc = aw + NAUX * u;
    lm[u] = make_float2((c.loops && nv0 && n0 != 0 && n0 != cc) ? 1.f : 0.f, (c.loops && nv1 && n1 != cc) ? 1.f : 0.f);
    pin(lm[u]);
  }
  auto load_ckpt = [&](int tau) {
    const float* cp = c.ckp + (long long)(tau / K) * V;
#pragma unroll
    for (int u = 0; u < RU; ++u) {
      const int cc = aw + NAUX * u;
      if (cc < S) {
        if (nv0) rnx[u].x = cp[cc * S + n0];
        if (nv1) rnx[u].y = cp[cc * S + n1];
      }
    }
  };
  ProdCursor pc;
  pc.init(max(len, 1), K);
  if (len > 0 && pc.tau > 0) load_ckpt(pc.tau);
  constexpr bool WIDE = (SCT >= 32);  // slot 0 (l = lane) is always a valid letter
  // row aw + NAUX*u exists for every auxiliary warp when NAUX*(u+1) <= S: known at compile time
  auto row_ok = [&](int u) { return (SCT > 0 && NAUX * (u + 1) <= SCT) ? true : (aw + NAUX * u < S); };
  const int GEND = len + c.LEAD + 2;
  for (int g = 0; g < GEND; ++g) {
    const int q = g - c.DBk;
    if (q >= 0 && q < len && !(c.A.dbg & 1)) {
      const int tau = pc.tau;
      const bool seg0 = pc.u == 0;
      pc.next();
      const int te = c.tab_e[q & (QRING - 1)];
      const unsigned r4 = row_r(c, tau) * 4;
      const unsigned er = rows_a + (te & 0xff) * rowb + r4;
      const unsigned ar = rows_a + c.tab_a[q & (QRING - 1)] * rowb + r4;
      mbar_wait(&c.mbar[te & 0xff], te >> 8);
      const float* pr = c.prow + (q & (QRING - 1)) * prowf;
      const int kq = (int)pr[S + 1];
      const float mq = pr[S];
      const float nmL = -mq * kLog2e;
      if (seg0) {
        if (tau == 0) {
#pragma unroll
          for (int u = 0; u < RU; ++u) ra[u] = make_float2(0.f, 0.f);
          if (aw == 0 && lane == 0) ra[0].x = pow2i(TARGET);
        } else {
#pragma unroll
          for (int u = 0; u < RU; ++u) ra[u] = rnx[u];
        }
      }
      float Pc[RU];
      float2 x2[RU];
#pragma unroll
      for (int u = 0; u < RU; ++u) {  // all shared-memory loads before the first store
        if (row_ok(u)) {
          Pc[u] = pr[aw + NAUX * u];
          x2[u] = make_float2(lds(er + o0 + NAUX * u * S * 4), lds(er + o1 + NAUX * u * S * 4));
        }
      }
      if (kq == 0) {
#pragma unroll
        for (int u = 0; u < RU; ++u) {
          if (row_ok(u)) {
            x2[u] = exp_pair(x2[u].x, x2[u].y, nmL);
            ra[u] = fmul2(x2[u], ffma2(lm[u], ra[u], splat2(Pc[u])));
          }
        }
      } else {  // rare: the forward sweep rescaled at this frame, through the exponent like it did
        const float nmLk = fmaf(mq, -kLog2e, (float)kq);
#pragma unroll
        for (int u = 0; u < RU; ++u) {
          if (row_ok(u)) {
            const float2 es = exp_pair(x2[u].x, x2[u].y, nmLk);
            x2[u] = exp_pair(x2[u].x, x2[u].y, nmL);
            ra[u] = fmul2(es, ffma2(lm[u], ra[u], splat2(Pc[u])));
          }
        }
      }
#pragma unroll
      for (int u = 0; u < RU; ++u) {
        if (row_ok(u)) {
          const float2 as = fmul2(ra[u], c1v);
          if (WIDE || nv0) {
            sts(er + o0 + NAUX * u * S * 4, x2[u].x);
            sts(ar + o0 + NAUX * u * S * 4, as.x);
          }
          if (nv1) {
            sts(er + o1 + NAUX * u * S * 4, x2[u].y);
            sts(ar + o1 + NAUX * u * S * 4, as.y);
          }
        }
      }
      if (q + 1 < len && pc.u == 0 && pc.tau > 0) load_ckpt(pc.tau);  // checkpoint of the next production segment
    }
    step_sync(c, nstep);
  }
}

// ---- scale warp: lagged power-of-two rescaling, final cost, posterior constants ----
template <int SCT>
__device__ __forceinline__ void scale_role(const Ctx& c) {
  const Dims<SCT> D(c);
  int nstep = 0;
  const int lane = c.lane, len = c.len, S = D.S, partrows = D.partrows, nmain = D.nmain, prowf = D.prowf;
  const float inf = __int_as_float(0x7f800000);
  int kcur = 0, Ksum = TARGET;
  cta_sync();
  const bool cv0 = lane < S, cv1 = lane + 32 < S;  // partial columns >= S hold unused (possibly non-finite) values
  for (int g = 0; g <= len + 1; ++g) {
    const int t = g - 1;
    if (t >= 0 && t <= len && !((c.A.dbg & 32) && t < len)) {
      const float* pb = c.part + (t & 1) * partrows * PARTW;
      float s = 0.f;
#pragma unroll 7
      for (int rr = 0; rr < partrows; ++rr) s += (cv0 ? pb[rr * PARTW + lane] : 0.f) + (cv1 ? pb[rr * PARTW + lane + 32] : 0.f);
      s = warp_sum(s);
      const bool ok = (s > 0.f) && (s < inf);
      if (t < len) {
        const int Knext = Ksum + kcur;  // K_{t+1}: alpha_{t+1} = a_{t+1} 2^-K_{t+1} exp(sum of maxima)
        if (lane == 0) {
          const float* mp = c.maxpart + (t & 1) * 4;  // frame maximum found by the producer warps
          float m = fmaxf(fmaxf(mp[0], mp[1]), fmaxf(mp[2], mp[3]));
          if (!(m > neg_inf())) m = 0.f;
          c.sc->msum += (double)m;
          c.psp[t * prowf + S] = m;
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
    step_sync(c, nstep);
  }
  if (!between_sweeps(c)) return;
  const float inv_mass = c.sc->inv_mass;
  const int klen = c.sc->klen;
  const int E1 = max(-120, min(120, -TARGET - exponent_of(inv_mass)));
  const float c1 = inv_mass * pow2i(E1) * c.gsc;
  int hcur = 0, Hsum = TARGET;
  ConsCursor cur;  // frame of main step ii = i + 1
  cur.init(max(len, 1), c.K);
  const int GEND = len + c.LEAD + 2;
  for (int g = 0; g < GEND; ++g) {
    const int i = g - c.LEAD;
    const int ii = i + 1;  // decisions for main step ii (frame len-1-ii)
    if (ii >= 0 && ii < len) {
      int hnext = 0;
      if (i >= 1) {
        const float* pb = c.part + (i & 1) * partrows * PARTW;
        float s = 0.f;
#pragma unroll 7
        for (int rr = 0; rr < nmain; ++rr) s += (cv0 ? pb[rr * PARTW + lane] : 0.f) + (cv1 ? pb[rr * PARTW + lane + 32] : 0.f);
        s = warp_sum(s);
        if ((s > 0.f) && (s < inf)) {
          const int ex = exponent_of(s) + hcur;
          if (ex > TARGET + BAND || ex < TARGET - BAND) hnext = max(-120, min(120, TARGET - ex));
        }
      }
      Hsum += hnext;
      hcur = hnext;
      const int slot = cur.q & (QRING - 1);
      cur.prev();
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
    step_sync(c, nstep);
  }
}

// ---- io warp: bulk loads, bulk stores of finished gradient rows, row pool ----------
template <int SCT>
__device__ __forceinline__ void io_role(const Ctx& c) {
  const Dims<SCT> D(c);
  int nstep = 0;
  const int lane = c.lane, len = c.len, V = D.V, NR = c.NR, K = c.K, rowf = D.rowf, prowf = D.prowf;
  int rowi = 0;
  if (lane == 0)
    for (int f = 0; f < DF && f < len; ++f) {
      issue_row(c, f, rowi, nullptr, nullptr);
      if (++rowi == NR) rowi = 0;
    }
  cta_sync();
  for (int g = 0; g <= len + 1; ++g) {
    const int f = g + DF;
    if (lane == 0 && f < len) {
      issue_row(c, f, rowi, nullptr, nullptr);
      if (++rowi == NR) rowi = 0;
    }
    step_sync(c, nstep);
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
  ConsCursor cur;  // frame consumed by main step i of this global step
  cur.init(max(len, 1), K);
  ProdCursor pc;   // frame of production index g
  pc.init(max(len, 1), K);
  int e1 = -1, a1 = -1, a2 = -1, t1 = 0;  // rows of main steps i-1 (logits, alpha) and i-2 (alpha); frame of i-1
  const int GEND = len + c.LEAD + 2;
  for (int g = 0; g < GEND; ++g) {
    const int i = g - c.LEAD;  // main step of this global step; its frame is len-1-i
    // 1. the logits row consumed by main step i-1 is free again
    if (lane == 0 && e1 >= 0) push_row(e1);
    // 2. store the gradient row main step i-1 finished
    if (a1 >= 0 && !(c.A.dbg & 16)) {
      const RowGeo ge = row_geo(t1, c.b, c.B, V, c.A.total);
      const float* src = c.rows + a1 * rowf;
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
      if (a2 >= 0) push_row(a2);
      if (g < len) {  // 4. fetch logits row + P row of production index g
        const int row = pop_row();
        const int ph = c.phase[row];
        c.phase[row] = ph ^ 1;
        c.tab_e[g & (QRING - 1)] = row | (ph << 8);
        issue_row(c, pc.tau, row, c.psp + pc.tau * prowf, c.prow + (g & (QRING - 1)) * prowf);
        pc.next();
      }
      const int qa = g - c.DBk + 1;  // 5. row that receives the recomputed alpha in the next step
      if (qa >= 0 && qa < len) c.tab_a[qa & (QRING - 1)] = pop_row();
    }
    // shift the row history by one main step
    a2 = a1;
    if (i >= 0 && i < len) {
      const int slot = cur.q & (QRING - 1);
      cur.prev();
      e1 = c.tab_e[slot] & 0xff;
      a1 = c.tab_a[slot];
      t1 = len - 1 - i;
    } else {
      e1 = -1;
      a1 = -1;
    }
    step_sync(c, nstep);
  }
  if (lane == 0) bulk_wait_all<0>();
}

// ---- the kernel ---------------------------------------------------------------
template <int SCT, int RU, int MAXT>
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
  c.LEAD = A.DB + A.K;
  c.B3 = A.B & 3; c.b3 = c.b & 3; c.V3 = A.V & 3;
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
  c.trace = A.trace;
  c.trace_steps = A.trace_steps;

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
  if (w < nm) main_role<SCT>(c);
  else if (w < nm + NPROD + NRECO) aux_role<SCT, RU>(c);
  else if (w == nm + NPROD + NRECO) scale_role<SCT>(c);
  else io_role<SCT>(c);
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
  p.ru = (S + 7) / 8;
  if (p.ru > 7 || p.threads > 736) return p;  // S <= 49 with the kernels instantiated below
  const int budget = 227 * 1024;
  int NR = 0;
  for (int n = 40; n >= 12; --n) {
    if (smem_plan(n, p.rowf, p.nmain, p.prowf).bytes <= budget) {
      NR = n;
      break;
    }
  }
  if (NR < 12) return p;
  // rows in flight in the backward sweep: 2K + DB + 4 ("row pool")
  p.DB = 2;
  p.K = (NR - p.DB - 4) / 2;
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

cudaError_t launch_bichar_den(const BicharArgs& a_in, const BicharPlan& p, cudaStream_t st) {
  BicharArgs a = a_in;
  // CTCG_BICHAR_TRACE=<file>: per-step clock stamps of every warp of CTA 0 (profiles/trace_roles.py reads it).
  const char* trace_path = getenv("CTCG_BICHAR_TRACE");
  const int warps = p.threads / 32;
  a.trace = nullptr;
  a.trace_steps = 0;
  {
    static const int dbg = [] {
      const char* e = getenv("CTCG_BICHAR_DEBUG");
      return e ? atoi(e) : 0;
    }();
    a.dbg = dbg;
  }
  if (trace_path && trace_path[0]) {
    a.trace_steps = 2 * a.T + 64;
    const size_t bytes = (size_t)warps * a.trace_steps * TRS * sizeof(long long);
    if (cudaMalloc(&a.trace, bytes) != cudaSuccess) return cudaGetLastError();
    cudaMemsetAsync(a.trace, 0, bytes, st);
  }
  ProfScope prof("bichar_den", st);
  cudaError_t e = cudaSuccess;
#define CTCG_BICHAR(SCT, RU)                                                                              \
  do {                                                                                                    \
    e = cudaFuncSetAttribute(bichar_den_kernel<SCT, RU, 736>, cudaFuncAttributeMaxDynamicSharedMemorySize, \
                             (int)p.smem_bytes);                                                          \
    if (e != cudaSuccess) return e;                                                                       \
    bichar_den_kernel<SCT, RU, 736><<<a.B, p.threads, p.smem_bytes, st>>>(a);                             \
  } while (0)
  if (a.S == 49) CTCG_BICHAR(49, 7);       // the WSJ character set (egs/wsj/vocabulary.txt)
  else if (a.S == 36) CTCG_BICHAR(36, 5);   // BASELINE.json's nominal V = 1296
  else if (p.ru <= 5) CTCG_BICHAR(0, 5);
  else CTCG_BICHAR(0, 7);
#undef CTCG_BICHAR
  ++g_launches;
  e = cudaGetLastError();
  if (a.trace) {
    const size_t n = (size_t)warps * a.trace_steps * TRS;
    std::vector<long long> host(n);
    cudaStreamSynchronize(st);
    cudaMemcpy(host.data(), a.trace, n * sizeof(long long), cudaMemcpyDeviceToHost);
    cudaFree(a.trace);
    if (FILE* f = fopen(trace_path, "wb")) {
      const int hdr[4] = {warps, a.trace_steps, p.nmain, TRS};
      fwrite(hdr, sizeof(int), 4, f);
      fwrite(host.data(), sizeof(long long), n, f);
      fclose(f);
    }
  }
  return e;
}

}  // namespace ctcg
