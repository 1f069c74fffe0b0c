// Bi-char CTC-G / CTC-GB denominator as two concurrent sweeps plus a streaming combine, sm_100a.
//
// Reference semantics: PathLogSumExp.forward/backward (att_speech/fst_utils.py:400-488) on the
// decoding graph of build_ctc_bigram_decoding_fst (fst_utils.py:759-817), called from
// FSTDecoder.get_fst_loss (advanced_decoder.py:497-500).  oracle/bichar_model.py states the
// same computation in numpy and is checked against the reference algorithm on the CPU.
//
// State (c, l) = (context, last letter).  Every arc into (c, l) carries class cls(c, l) and its
// source set depends on c only, so with e_t = exp(x_t - max x_t), m(c,l) = explicit self-loop:
//   forward  sweep  A'[c][l] = e[c][l] (P[c] + m A[c][l])   l >= 1;   P'[l] = sum_c A'[c][l] + e0[l] P[l]
//   backward sweep  F'[c][l] = e[c][l] (Q[l] + m F[c][l])   l >= 1;   Q'[c] = sum_l F'[c][l] + e0[c] Q[c]
//   posterior       gamma_t[cls(c,l)] = e_t (P_t[c] + m A_t) (Q_{t+1}[l] + m F_{t+1}) / Z,
//                   gamma_t[cls(c,0)] += e0_t[c] P_t[c] Q_{t+1}[c] / Z
// (the blank states (c, 0) are folded into the e0 term: a_{t+1}[c][0] = e0[c] P_t[c]).
//
// bisweep_kernel   grid (B, 2): CTA (b, 0) runs the forward recursion of utterance b, CTA (b, 1)
//                  the backward recursion, at the same time.  Each spills only the S-vector
//                  entering every frame (P_t / Q_{t+1}), the frame maximum and the power-of-two
//                  rescale exponents, plus the S*S per-state values every K frames when the
//                  graph has explicit self-loops.  Roles: an io lane streams logits rows through
//                  a ring of cp.async.bulk (TMA, UBLKCP) copies; producer warps turn a landed row
//                  into e (row held in registers, one MUFU.EX2 per class) in a bank-conflict-free
//                  layout, several frames ahead; the main warps own 8 states per thread and run
//                  the recursion with ONE named barrier per frame among themselves (the per-frame
//                  chain is 2 LDS.128, 4 FFMA2 + 4 FMUL2, an 8-lane shuffle tree, 1 STS); a scale
//                  warp spills the vectors and decides the lagged rescale.
// bicombine_kernel grid (T/K, B): embarrassingly parallel; re-creates the per-state values of a
//                  K-frame segment elementwise from the spills (forward from the A checkpoint,
//                  backward from the F checkpoint) and writes the posterior rows: reads the
//                  logits once, writes the gradient once.
//
// HBM traffic per valid frame (fp32 words): sweeps 2V (+ 2(S+3) + 2V/K), combine 2V (+ 2V/K).
#include <cstdio>
#include <cstdlib>

#include "device_utils.cuh"
#include "kernels.cuh"

namespace ctcg {
namespace {

constexpr int NPROD = 4;     // producer warps
constexpr int NRAW = 12;     // TMA landing rows in flight
constexpr int NE = 6;        // e rows ahead of the recursion
constexpr int TARGET = 60;   // the spilled vectors are kept near 2^TARGET
constexpr int BAND = 20;
constexpr int RBW = 64;      // floats per vector buffer (S <= 64)
constexpr int CTHREADS = 256;
constexpr int KSEG = 8;      // frames per checkpoint segment (power of two)

// ---- PTX wrappers -----------------------------------------------------------
__device__ __forceinline__ unsigned smem_u32(const void* p) {
  return static_cast<unsigned>(__cvta_generic_to_shared(p));
}
__device__ __forceinline__ void mbar_init(unsigned bar, unsigned count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(unsigned bar, unsigned bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(unsigned bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ bool mbar_try(unsigned bar, unsigned parity) {
  unsigned ok;
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n"
      "selp.u32 %0, 1, 0, p;\n"
      "}\n"
      : "=r"(ok)
      : "r"(bar), "r"(parity)
      : "memory");
  return ok != 0;
}
__device__ __forceinline__ void mbar_wait(unsigned bar, unsigned parity) {
  while (!mbar_try(bar, parity)) {
  }
}
// 1-D bulk copy global -> shared (TMA engine, SASS UBLKCP); 16-byte aligned, size % 16 == 0.
__device__ __forceinline__ void bulk_g2s(unsigned dst, const void* src, unsigned bytes, unsigned bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst),
               "l"(src), "r"(bytes), "r"(bar)
               : "memory");
}
__device__ __forceinline__ void fence_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ void named_sync(int id, int count) {
  asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(count) : "memory");
}
__device__ __forceinline__ float lds(unsigned a) {
  float v;
  asm volatile("ld.shared.f32 %0, [%1];" : "=f"(v) : "r"(a));
  return v;
}
__device__ __forceinline__ int ldsi(unsigned a) {
  int v;
  asm volatile("ld.shared.s32 %0, [%1];" : "=r"(v) : "r"(a));
  return v;
}
__device__ __forceinline__ float4 lds4(unsigned a) {
  float4 v;
  asm volatile("ld.shared.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(a));
  return v;
}
__device__ __forceinline__ void sts(unsigned a, float v) { asm volatile("st.shared.f32 [%0], %1;" ::"r"(a), "f"(v) : "memory"); }
__device__ __forceinline__ void stsi(unsigned a, int v) { asm volatile("st.shared.s32 [%0], %1;" ::"r"(a), "r"(v) : "memory"); }

__device__ __forceinline__ float pow2i(int k) {  // 2^k for -126 <= k <= 127
  return __int_as_float((k + 127) << 23);
}
__device__ __forceinline__ int exponent_of(float v) {  // floor(log2 v) for normal v > 0
  return ((__float_as_int(v) >> 23) & 0xff) - 127;
}

// ---- shared-memory plan of a sweep CTA ------------------------------------------
struct SwSmem {
  int raw, eb, rbuf, mbuf, kflag, mbar, bytes;
};
__host__ __device__ inline SwSmem sweep_smem(int rowf, int estride) {
  SwSmem s;
  int o = 0;
  s.raw = o;    o += NRAW * rowf * 4;
  s.eb = o;     o += NE * estride * 4;
  s.rbuf = o;   o += 2 * RBW * 4;
  s.mbuf = o;   o += 8 * 4;
  s.kflag = o;  o += 4 * 4;
  o = (o + 15) & ~15;
  s.mbar = o;   o += (2 * NRAW + 2 * NE) * 8;
  s.bytes = o;
  return s;
}

// Geometry of one (t, b) logits row inside its 16-byte aligned window.
struct RowGeo {
  long long w0;   // first float of the window (multiple of 4)
  int r;          // row starts at window float r (0..3)
  int nbulk;      // floats fetched by the bulk copy (multiple of 4)
};
__device__ __forceinline__ RowGeo row_geo(int t, int b, int B, int V, long long total) {
  RowGeo g;
  const long long i0 = ((long long)t * B + b) * V;
  g.r = (int)(i0 & 3);
  g.w0 = i0 - g.r;
  const int nfl = pad4(g.r + V);
  const long long room = (total - g.w0) & ~3LL;
  g.nbulk = (int)(room < (long long)nfl ? room : (long long)nfl);
  return g;
}

template <int SCT>
struct SDims {
  int S, V, PF, ZO;
  __device__ __forceinline__ explicit SDims(const BisweepArgs& a) {
    S = SCT > 0 ? SCT : a.S;
    V = S * S;
    PF = S | 1;             // odd pitch: the 8 scalar e loads of a main thread are conflict-free
    ZO = pad4(S * PF);      // a 16-byte chunk of zeros behind the e rows
  }
};

// ---- one sweep (DIR 0: forward in time, P vectors; DIR 1: backward in time, Q vectors) ------
template <int SCT, int DIR>
__device__ __forceinline__ void sweep_body(const BisweepArgs& A, unsigned char* smraw) {
  const SDims<SCT> D(A);
  const int S = D.S, V = D.V, PF = D.PF, ZO = D.ZO;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int b = blockIdx.x, B = A.B;
  const int len = min(A.lens[b], A.T);
  const int nmain = A.nmain;
  const int rowf = A.rowf, estride = A.estride, prowf = A.prowf;
  const bool loops = A.loops != 0, ctx = A.ctx != 0;
  const SwSmem L = sweep_smem(rowf, estride);
  const unsigned sm0 = smem_u32(smraw);
  const unsigned raw_a = sm0 + L.raw, eb_a = sm0 + L.eb, rb_a = sm0 + L.rbuf, mb_a = sm0 + L.mbuf,
                 kf_a = sm0 + L.kflag, bar_a = sm0 + L.mbar;
  auto raw_full = [&](int s) { return bar_a + s * 8; };
  auto raw_empty = [&](int s) { return bar_a + (NRAW + s) * 8; };
  auto e_full = [&](int s) { return bar_a + (2 * NRAW + s) * 8; };
  auto e_empty = [&](int s) { return bar_a + (2 * NRAW + NE + s) * 8; };
  const int NB = (nmain + 1) * 32;  // threads of the per-frame barrier: main warps + scale warp
  float* spill = A.spill[DIR] + (long long)b * A.T * prowf;
  float* ck = A.ck[DIR] + (long long)b * A.nck * V;

  // ---- init ----
  {
    float* ebf = reinterpret_cast<float*>(smraw + L.eb);
    for (int i = tid; i < NE * estride; i += blockDim.x) ebf[i] = 0.f;  // pads, l = 0 slots, zero chunks
    float* rb = reinterpret_cast<float*>(smraw + L.rbuf);
    for (int i = tid; i < 2 * RBW; i += blockDim.x) {
      const int g = i & (RBW - 1);
      float v = 0.f;
      if (i < RBW) v = (DIR == 0) ? (g == 0 ? pow2i(TARGET) : 0.f) : (g < S ? pow2i(TARGET) : 0.f);
      rb[i] = v;
    }
    if (tid < 4) reinterpret_cast<int*>(smraw + L.kflag)[tid] = 0;
    if (tid == 0) {
      for (int s = 0; s < NRAW; ++s) {
        mbar_init(raw_full(s), 1);
        mbar_init(raw_empty(s), 1);
      }
      for (int s = 0; s < NE; ++s) {
        mbar_init(e_full(s), 1);
        mbar_init(e_empty(s), nmain + 1);
      }
      asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
  }
  __syncthreads();
  auto frame_t = [&](int f) { return DIR == 0 ? f : len - 1 - f; };

  if (warp < nmain) {
    // ============================ main warps: the recursion ============================
    const int g = warp * 4 + (lane >> 3), k8 = lane & 7;
    const bool gv = g < S;
    unsigned eo[8];   // byte offsets of the 8 states' classes inside an e stage
    float2 m2[4];     // explicit self-loop masks
    float mk[8];
#pragma unroll
    for (int s = 0; s < 8; ++s) {
      const int o = 4 * (k8 + 8 * (s >> 2)) + (s & 3);
      bool valid;
      int c, l;
      if (DIR == 0) { c = o; l = g; valid = gv && o < S && g >= 1; }
      else          { c = g; l = o; valid = gv && o < S && o >= 1; }
      eo[s] = valid ? (unsigned)(c * PF + l) * 4u : (unsigned)(ZO + (s & 3)) * 4u;
      mk[s] = (valid && loops && c != l) ? 1.f : 0.f;
    }
#pragma unroll
    for (int h = 0; h < 4; ++h) m2[h] = make_float2(mk[2 * h], mk[2 * h + 1]);
    const unsigned e0o = (unsigned)((ctx && gv) ? g * PF : 0) * 4u;  // blank class of context g (class 0 when global)
    const unsigned rc0 = (unsigned)(4 * k8) * 4u, rc1 = (unsigned)(4 * (k8 + 8)) * 4u;
    float2 val[4];
#pragma unroll
    for (int h = 0; h < 4; ++h) val[h] = make_float2(0.f, 0.f);
    float Rown = (DIR == 0) ? (g == 0 ? pow2i(TARGET) : 0.f) : (gv ? pow2i(TARGET) : 0.f);
    float ev[8], e0 = 0.f;
#pragma unroll
    for (int s = 0; s < 8; ++s) ev[s] = 0.f;
    if (len > 0) {
      mbar_wait(e_full(0), 0);
#pragma unroll
      for (int s = 0; s < 8; ++s) ev[s] = lds(eb_a + eo[s]);
      e0 = lds(eb_a + e0o);
    }
    int es = 0, esn = 0;           // e stage of this frame
    unsigned epar = 0;             // parity of the stage the NEXT frame waits on
    long long tw = 0, tb = 0, tt0 = clock64();
    const bool tim = (A.dbg & 256) && b == 0 && tid == 0;
    for (int f = 0; f < len; ++f) {
      const int t = frame_t(f);
      // prefetch the next frame's e values (registers evn); its stage is ready several frames ahead
      float evn[8], e0n = 0.f;
      esn = es + 1;
      unsigned eparn = epar;
      if (esn == NE) { esn = 0; eparn ^= 1; }
      if (f + 1 < len) {
        long long c0 = tim ? clock64() : 0;
        mbar_wait(e_full(esn), eparn);
        if (tim) tw += clock64() - c0;
        const unsigned eb = eb_a + esn * estride * 4;
#pragma unroll
        for (int s = 0; s < 8; ++s) evn[s] = lds(eb + eo[s]);
        e0n = lds(eb + e0o);
      } else {
#pragma unroll
        for (int s = 0; s < 8; ++s) evn[s] = 0.f;
      }
      // checkpoint of the per-state values entering this frame (explicit self-loops only)
      if (loops && (DIR == 0 ? (t > 0 && (t & (KSEG - 1)) == 0) : ((t & (KSEG - 1)) == KSEG - 1))) {
        float* cp = ck + (long long)(t / KSEG) * V;
#pragma unroll
        for (int s = 0; s < 8; ++s) {
          const int o = 4 * (k8 + 8 * (s >> 2)) + (s & 3);
          const bool valid = gv && o < S && (DIR == 0 ? g >= 1 : o >= 1);
          const int cls = DIR == 0 ? o * S + g : g * S + o;
          const float v = (s & 1) ? val[s >> 1].y : val[s >> 1].x;
          if (valid) cp[cls] = v;
        }
      }
      const unsigned rb = rb_a + (f & 1) * RBW * 4;
      if (A.dbg & 1) {  // timing experiment: barriers and hand-shakes only
        if (lane == 0) mbar_arrive(e_empty(es));
        named_sync(1, NB);
        es = esn;
        epar = eparn;
        continue;
      }
      const float4 r0 = lds4(rb + rc0), r1 = lds4(rb + rc1);
      const int kk = ldsi(kf_a + (f & 3) * 4);
      val[0] = fmul2(make_float2(ev[0], ev[1]), ffma2(m2[0], val[0], make_float2(r0.x, r0.y)));
      val[1] = fmul2(make_float2(ev[2], ev[3]), ffma2(m2[1], val[1], make_float2(r0.z, r0.w)));
      val[2] = fmul2(make_float2(ev[4], ev[5]), ffma2(m2[2], val[2], make_float2(r1.x, r1.y)));
      val[3] = fmul2(make_float2(ev[6], ev[7]), ffma2(m2[3], val[3], make_float2(r1.z, r1.w)));
      const float2 s2 = fadd2(fadd2(val[0], val[1]), fadd2(val[2], val[3]));
      float tot = s2.x + s2.y;
      tot += __shfl_xor_sync(0xffffffffu, tot, 1);
      tot += __shfl_xor_sync(0xffffffffu, tot, 2);
      tot += __shfl_xor_sync(0xffffffffu, tot, 4);
      float Rn = fmaf(e0, Rown, tot);
      if (kk != 0) {  // rare: power-of-two rescale decided two frames ago
        const float sc = pow2i(kk);
        const float2 sc2 = splat2(sc);
#pragma unroll
        for (int h = 0; h < 4; ++h) val[h] = fmul2(val[h], sc2);
        Rn *= sc;
      }
      Rown = Rn;
      if (k8 == 0 && gv) sts(rb_a + ((f + 1) & 1) * RBW * 4 + g * 4, Rn);
      if (lane == 0) mbar_arrive(e_empty(es));
      {
        long long c0 = tim ? clock64() : 0;
        named_sync(1, NB);
        if (tim) tb += clock64() - c0;
      }
#pragma unroll
      for (int s = 0; s < 8; ++s) ev[s] = evn[s];
      e0 = e0n;
      es = esn;
      epar = eparn;
    }
    if (tim) printf("sweep dir %d main: total %lld  wait_e_full %lld  barrier %lld  (len %d)\n", DIR, clock64() - tt0, tw, tb, len);
  } else if (warp == nmain) {
    // ============================ scale / spill warp ============================
    int kcur = 0, knext = 0, Kcum = TARGET;
    double msum = 0.0;
    int es = 0;
    unsigned epar = 0;
    const float inf = __int_as_float(0x7f800000);
    bool dead = false;
    for (int f = 0; f < len; ++f) {
      const int t = frame_t(f);
      const unsigned rb = rb_a + (f & 1) * RBW * 4;
      if (A.dbg & 8) {
        mbar_wait(e_full(es), epar);
        if (lane == 0) mbar_arrive(e_empty(es));
        if (++es == NE) { es = 0; epar ^= 1; }
        named_sync(1, NB);
        continue;
      }
      const float v0 = lds(rb + lane * 4), v1 = lds(rb + (lane + 32) * 4);
      float* sp = spill + (long long)t * prowf;
      if (lane < S) sp[lane] = v0;
      if (lane + 32 < S) sp[lane + 32] = v1;
      const float mx = warp_max(fmaxf(v0, v1));
      int knext2 = 0;
      if (mx > 0.f && mx < inf) {
        const int ex = exponent_of(mx) + kcur + knext;
        if (ex > TARGET + BAND || ex < TARGET - BAND) knext2 = max(-120, min(120, TARGET - ex));
      } else {
        dead = true;
      }
      mbar_wait(e_full(es), epar);
      const float m = lds(mb_a + es * 4);
      if (lane == 0) {
        msum += (double)m;
        sp[S] = m;
        sp[S + 1] = (float)kcur;
        sp[S + 2] = (float)Kcum;
        stsi(kf_a + ((f + 2) & 3) * 4, knext2);
      }
      __syncwarp();
      if (lane == 0) mbar_arrive(e_empty(es));
      Kcum += kcur;
      kcur = knext;
      knext = knext2;
      if (++es == NE) { es = 0; epar ^= 1; }
      named_sync(1, NB);
    }
    if (DIR == 0) {  // total mass = sum_c P_len[c]; cost = log(mass) - K ln2 + sum of the frame maxima
      const unsigned rb = rb_a + (len & 1) * RBW * 4;
      const float v0 = lds(rb + lane * 4), v1 = lds(rb + (lane + 32) * 4);
      const float s = warp_sum(v0 + v1);
      if (lane == 0) {
        const bool ok = !dead && (s > 0.f) && (s < inf);
        A.cost[b] = ok ? (log((double)s) - (double)Kcum * kLn2 + msum) : -1e20;  // the reference's sentinel
        BiFin fin;
        fin.inv_mass = ok ? 1.f / s : 0.f;
        fin.kfinal = Kcum;
        fin.dead = ok ? 0 : 1;
        fin.pad = 0;
        A.fin[b] = fin;
      }
    }
  } else if (warp < nmain + 1 + NPROD) {
    // ============================ producers: x row -> e row ============================
    const int pw = warp - nmain - 1;
    constexpr int NQ = SCT > 0 ? (SCT * SCT + 31) / 32 : 1;
    const int B3 = B & 3, b3 = b & 3, V3 = V & 3;
    long long pw1 = 0, pw2 = 0, pw3 = 0, pt0 = clock64();
    const bool ptim = (A.dbg & 256) && b == 0 && pw == 0 && lane == 0;
    for (int f = pw; f < len; f += NPROD) {
      const int t = frame_t(f);
      const int rs = f % NRAW, es = f % NE;
      const int r = ((((t & 3) * B3 + b3) & 3) * V3) & 3;  // ((t*B + b) * V) mod 4
      const unsigned src = raw_a + (rs * rowf + r) * 4;
      const unsigned dst = eb_a + es * estride * 4;
      long long c1 = 0;
      {
        long long c0 = ptim ? clock64() : 0;
        mbar_wait(raw_full(rs), (f / NRAW) & 1);
        if (ptim) { c1 = clock64(); pw1 += c1 - c0; }
      }
      if (SCT > 0) {
        float xr[NQ];
        float mx = neg_inf();
#pragma unroll
        for (int q = 0; q < NQ; ++q) {
          const int v = q * 32 + lane;
          if (A.dbg & 16) { xr[q] = 0.f; continue; }
          xr[q] = (q * 32 + 31 < V || v < V) ? lds(src + v * 4) : neg_inf();
          mx = fmaxf(mx, xr[q]);
        }
        mx = warp_max(mx);  // every lane's loads have landed in registers here
        if (ptim) pw3 += clock64() - c1;
        if (lane == 0) mbar_arrive(raw_empty(rs));
        if (!(mx > neg_inf())) mx = 0.f;
        const float nmL = -mx * kLog2e;
        {
          long long c0 = ptim ? clock64() : 0;
          if (f >= NE) mbar_wait(e_empty(es), ((f / NE) - 1) & 1);
          if (ptim) pw2 += clock64() - c0;
        }
        // class v = c*S + l lands at c*PF + l: the same index when S is odd (PF = S).  The l = 0
        // positions hold the blank classes' e0; the main warps never read them as states.
#pragma unroll
        for (int q = 0; q < NQ; ++q) {
          if (A.dbg & 2) break;
          const int v = q * 32 + lane;
          const float e = ex2(fmaf(xr[q], kLog2e, nmL));
          const int vp = (SCT & 1) ? v : v + v / SCT;
          if (q * 32 + 31 < V || v < V) sts(dst + vp * 4, e);
        }
        if (lane == 0) sts(mb_a + es * 4, mx);
      } else {
        float mx = neg_inf();
        for (int v = lane; v < V; v += 32) mx = fmaxf(mx, lds(src + v * 4));
        mx = warp_max(mx);
        if (!(mx > neg_inf())) mx = 0.f;
        const float nmL = -mx * kLog2e;
        if (f >= NE) mbar_wait(e_empty(es), ((f / NE) - 1) & 1);
        for (int v = lane; v < V; v += 32) {
          const int c = v / S, l = v - c * S;
          const float e = ex2(fmaf(lds(src + v * 4), kLog2e, nmL));
          sts(dst + (c * PF + l) * 4, e);
        }
        __syncwarp();
        if (lane == 0) {
          mbar_arrive(raw_empty(rs));
          sts(mb_a + es * 4, mx);
        }
      }
      __syncwarp();
      if (lane == 0) mbar_arrive(e_full(es));
    }
    if (ptim) printf("sweep dir %d producer 0: total %lld  wait_raw_full %lld  wait_e_empty %lld  load+max %lld\n", DIR, clock64() - pt0, pw1, pw2, pw3);
  } else {
    // ============================ io lane: TMA row loads ============================
    if (lane == 0) {
      long long iw = 0, it0 = clock64();
      const bool itim = (A.dbg & 256) && b == 0;
      for (int f = 0; f < len; ++f) {
        const int rs = f % NRAW;
        {
          long long c0 = itim ? clock64() : 0;
          if (f >= NRAW) mbar_wait(raw_empty(rs), ((f / NRAW) - 1) & 1);
          if (itim) iw += clock64() - c0;
        }
        const RowGeo gq = row_geo(frame_t(f), b, B, V, A.total);
        const unsigned dst = raw_a + rs * rowf * 4;
        // <= 3 floats at the very end of the tensor do not fill a 16-byte chunk: plain stores, made
        // visible to the producers by the release of the arrive below
        for (int idx = gq.nbulk; idx < gq.r + V; ++idx) sts(dst + idx * 4, A.x[gq.w0 + idx]);
        if (A.dbg & 32) {
          mbar_arrive(raw_full(rs));
          continue;
        }
        fence_async_smem();
        mbar_expect_tx(raw_full(rs), gq.nbulk * 4);
        bulk_g2s(dst, A.x + gq.w0, gq.nbulk * 4, raw_full(rs));
      }
      if (itim) printf("sweep dir %d io: total %lld  wait_raw_empty %lld\n", DIR, clock64() - it0, iw);
    }
  }
}

template <int SCT, int MAXT>
__global__ void __launch_bounds__(MAXT, 1) bisweep_kernel(const BisweepArgs A) {
  extern __shared__ __align__(128) unsigned char smraw[];
  if (blockIdx.y == 0) sweep_body<SCT, 0>(A, smraw);
  else sweep_body<SCT, 1>(A, smraw);
}

// ---- combine: posterior rows of one K-frame segment of one utterance -----------------------
template <int SCT, int KMAX>
__global__ void __launch_bounds__(CTHREADS) bicombine_kernel(const BisweepArgs A) {
  const int S = SCT > 0 ? SCT : A.S;
  const int V = S * S;
  const int b = blockIdx.y, j = blockIdx.x, K = A.K, B = A.B, T = A.T, prowf = A.prowf;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int t0 = j * K;
  if (t0 >= T) return;
  const int len = min(A.lens[b], T);
  const int tz = min(t0 + K, T);
  const BiFin fin = A.fin[b];
  const int t1 = min(t0 + K, len);
  const int n = (fin.dead || t1 <= t0) ? 0 : t1 - t0;
  // rows beyond the utterance's end (and every row of an utterance with no accepting path) are zero
  for (int t = t0 + n; t < tz; ++t) {
    float* dst = A.grad + ((long long)t * B + b) * V;
    for (int v = tid; v < V; v += CTHREADS) dst[v] = 0.f;
  }
  if (n == 0) return;

  __shared__ float Ps[KMAX][RBW], Qs[KMAX][RBW];
  __shared__ float nmL_s[KMAX], scF_s[KMAX], scB_s[KMAX], cB_s[KMAX], D_s[KMAX];
  const float* spF = A.spill[0] + ((long long)b * T + t0) * prowf;
  const float* spB = A.spill[1] + ((long long)b * T + t0) * prowf;
  const float gsc = A.g_scale ? A.g_scale[b] : 1.f;
  // gamma = a^ b^ 2^E_t / mass with E_t = K_len - Ka_{t+1} - Kb_{t+1}; split into
  // cA = gscale / mass * 2^E1 (about 2^-TARGET, rides on the a recurrence) and cB_t = 2^(E_t - E1)
  const int E1 = max(-120, min(120, -TARGET - exponent_of(fin.inv_mass)));
  const float cA = fin.inv_mass * pow2i(E1) * gsc;
  for (int i = tid; i < n * RBW; i += CTHREADS) {
    const int f = i / RBW, c = i - f * RBW;
    Ps[f][c] = (c < S) ? spF[f * prowf + c] * cA : 0.f;
    Qs[f][c] = (c < S) ? spB[f * prowf + c] : 0.f;
  }
  if (tid < n) {
    const float* rf = spF + tid * prowf;
    const float* rq = spB + tid * prowf;
    nmL_s[tid] = -rf[S] * kLog2e;
    scF_s[tid] = pow2i((int)rf[S + 1]);
    scB_s[tid] = pow2i((int)rq[S + 1]);
    const int Et = fin.kfinal - ((int)rf[S + 2] + (int)rf[S + 1]) - (int)rq[S + 2];
    cB_s[tid] = pow2i(max(-126, min(126, Et - E1)));
  }
  __syncthreads();
  const bool loops = A.loops != 0, ctx = A.ctx != 0;
  if (!ctx) {  // one global blank: its posterior is the sum over the S blank states
    for (int f = warp; f < n; f += CTHREADS / 32) {
      const float cb = cB_s[f];
      float d = Ps[f][lane] * (Qs[f][lane] * cb) + Ps[f][lane + 32] * (Qs[f][lane + 32] * cb);
      d = warp_sum(d);
      if (lane == 0) D_s[f] = d;
    }
    __syncthreads();
  }
  const float* x0 = A.x + ((long long)t0 * B + b) * V;
  float* g0 = A.grad + ((long long)t0 * B + b) * V;
  const long long rstride = (long long)B * V;
  const float* ckA = A.ck[0] + ((long long)b * A.nck + j) * V;
  const float* ckF = A.ck[1] + ((long long)b * A.nck + j) * V;
  for (int v = tid; v < V; v += CTHREADS) {
    const int c = v / S, l = v - c * S;
    float e[KMAX], ah[KMAX];
#pragma unroll
    for (int f = 0; f < KMAX; ++f)
      if (f < n) e[f] = __ldg(x0 + f * rstride + v);
    if (l == 0) {
#pragma unroll
      for (int f = 0; f < KMAX; ++f) {
        if (f < n) {
          const float ef = ex2(fmaf(e[f], kLog2e, nmL_s[f])) * scF_s[f];
          float gm;
          if (ctx) gm = ef * Ps[f][c] * (Qs[f][c] * cB_s[f]);
          else gm = (v == 0) ? ef * D_s[f] : 0.f;
          g0[f * rstride + v] = gm;
        }
      }
      continue;
    }
    const bool m = loops && l != c;
    const float mf = m ? 1.f : 0.f;
    // the checkpoints hold the sweeps' scaled values; cA rides on the forward recurrence (Ps has it too)
    float a = (m && j > 0) ? ckA[v] * cA : 0.f;
    float fv = (m && t1 < len) ? ckF[v] : 0.f;
#pragma unroll
    for (int f = 0; f < KMAX; ++f) {
      if (f < n) {
        e[f] = ex2(fmaf(e[f], kLog2e, nmL_s[f]));
        ah[f] = e[f] * scF_s[f] * fmaf(mf, a, Ps[f][c]);  // a_{t+1}[c][l] * cA
        a = ah[f];
      }
    }
#pragma unroll
    for (int f = KMAX - 1; f >= 0; --f) {
      if (f < n) {
        const float bh = fmaf(mf, fv, Qs[f][l]);          // b_{t+1}[c][l]
        g0[f * rstride + v] = ah[f] * (bh * cB_s[f]);
        fv = e[f] * scB_s[f] * bh;                        // f_t[c][l]
      }
    }
  }
}

}  // namespace

// ---- host side ------------------------------------------------------------------
BisweepPlan bisweep_plan(int T, int B, int V, int S) {
  BisweepPlan p{};
  p.ok = 0;
  if (S < 4 || S > 64 || V != S * S || T < 1 || B < 1) return p;
  p.nmain = (S + 3) / 4;
  p.threads = (p.nmain + 1 + NPROD + 1) * 32;
  p.rowf = pad4(V + 3);
  p.prowf = pad4(S + 3);
  const int PF = S | 1;
  p.estride = pad4(S * PF) + 4;
  p.K = KSEG;
  p.nck = (T + p.K - 1) / p.K + 1;
  p.smem_bytes = sweep_smem(p.rowf, p.estride).bytes;
  if (p.smem_bytes > 227 * 1024 || p.threads > 704) return p;
  p.spill_bytes = (size_t)B * T * p.prowf * 4;
  p.ckpt_bytes = (size_t)B * p.nck * V * 4;
  p.ok = 1;
  return p;
}

cudaError_t launch_bisweep(const BisweepArgs& a_in, const BisweepPlan& p, cudaStream_t st) {
  cudaError_t e = cudaSuccess;
  BisweepArgs a = a_in;
  {
    static const int dbg = [] {
      const char* v = getenv("CTCG_SWEEP_DEBUG");
      return v ? atoi(v) : 0;
    }();
    a.dbg = dbg;  // timing experiments only: roles skip their work, results invalid
  }
  const dim3 grid(a.B, (a.grad && !(a.dbg & 4)) ? 2 : 1);
  {
    ProfScope prof("bisweep", st);
#define CTCG_SWEEP(SCT, MAXT)                                                                               \
  do {                                                                                                   \
    e = cudaFuncSetAttribute(bisweep_kernel<SCT, MAXT>, cudaFuncAttributeMaxDynamicSharedMemorySize,    \
                             (int)p.smem_bytes);                                                         \
    if (e != cudaSuccess) return e;                                                                      \
    bisweep_kernel<SCT, MAXT><<<grid, p.threads, p.smem_bytes, st>>>(a);                                \
  } while (0)
#define MAXT_OF(S_) ((((S_) + 3) / 4 + 1 + NPROD + 1) * 32)
    if (a.S == 49) CTCG_SWEEP(49, MAXT_OF(49));       // the WSJ character set (egs/wsj/vocabulary.txt)
    else if (a.S == 36) CTCG_SWEEP(36, MAXT_OF(36));  // BASELINE.json's nominal V = 1296
    else CTCG_SWEEP(0, 704);
#undef MAXT_OF
#undef CTCG_SWEEP
    ++g_launches;
    e = cudaGetLastError();
    if (e != cudaSuccess) return e;
  }
  if (!a.grad) return e;
  {
    ProfScope prof("bicombine", st);
    const dim3 cgrid((a.T + a.K - 1) / a.K, a.B);
    if (a.S == 49) bicombine_kernel<49, 8><<<cgrid, CTHREADS, 0, st>>>(a);
    else if (a.S == 36) bicombine_kernel<36, 8><<<cgrid, CTHREADS, 0, st>>>(a);
    else bicombine_kernel<0, 8><<<cgrid, CTHREADS, 0, st>>>(a);
    ++g_launches;
    e = cudaGetLastError();
  }
  return e;
}

}  // namespace ctcg
