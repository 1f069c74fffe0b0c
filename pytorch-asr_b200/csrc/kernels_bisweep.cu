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
#include <type_traits>

#include "device_utils.cuh"
#include "kernels.cuh"

namespace ctcg {
namespace {

constexpr int NPROD = 4;     // producer warps
#ifndef CTCG_NRAW
#define CTCG_NRAW 10
#endif
#ifndef CTCG_NE
#define CTCG_NE 6
#endif
#ifndef CTCG_SWEEP_MINB
#define CTCG_SWEEP_MINB 1
#endif
constexpr int NRAW = CTCG_NRAW;  // TMA landing rows in flight
constexpr int NE = CTCG_NE;      // e rows ahead of the recursion
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
__device__ __forceinline__ void sts4(unsigned a, float4 v) {
  asm volatile("st.shared.v4.f32 [%0], {%1,%2,%3,%4};" ::"r"(a), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
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
  int raw, eb, edb, rbuf, msum, kflag, mbar, bytes;
};
__host__ __device__ inline SwSmem sweep_smem(int rowf, int estride, int epad) {
  SwSmem s;
  int o = 0;
  s.raw = o;    o += NRAW * rowf * 4;
  s.eb = o;     o += (NE * estride + epad) * 4;   // e rows of NE frames, then zeros (over-reads of the last stage)
  s.edb = o;    o += NE * RBW * 4;                // per frame: e0[c] + e[c][c]
  s.rbuf = o;   o += 2 * RBW * 4;
  s.msum = o;   o += NPROD * 8;  // per-producer sums of the frame maxima (fp64)
  s.kflag = o;  o += 4 * 4;
  o = (o + 15) & ~15;
  s.mbar = o;   o += (2 * NRAW + 2 * NE + 2) * 8;
  s.bytes = o;
  return s;
}

template <int SCT>
struct SDims {
  int S, V, PF;
  __device__ __forceinline__ explicit SDims(const BisweepArgs& a) {
    S = SCT > 0 ? SCT : a.S;
    V = S * S;
    PF = S | 1;  // odd row pitch: the scalar e loads of the main warps are bank-conflict free in both directions
  }
};

// ---- one sweep (DIR 0: forward in time, P vectors; DIR 1: backward in time, Q vectors) ------
//
// Main-warp layout: context/letter group g = warp*8 + (lane & 7), k4 = lane >> 3; a thread owns the
// 16 states o = 32*(s >> 3) + 8*k4 + (s & 7), s = 0..15, of its group (DIR 0: state (c=o, l=g), DIR 1:
// state (c=g, l=o)); the group's sum is two xor-shuffles (8, 16) away.  States that do not exist are
// neutral without any masking: o >= S meets R[o] = 0, and the producers zero the blank column (c, 0)
// and the diagonal (c, c) of every e row after folding them into ed[c] = e0[c] + e[c][c]:
//     R'[g] = sum_o val'[o] + ed[g] * R[g]
// (the diagonal state has no explicit self-loop and the blank state none either, so both are functions
// of R[g] alone).
template <int SCT, int DIR>
__device__ __forceinline__ void sweep_body(const BisweepArgs& A, unsigned char* smraw) {
  const SDims<SCT> D(A);
  const int S = D.S, V = D.V, PF = D.PF;
  const int tid = threadIdx.x, lane = tid & 31;
  const int b = blockIdx.x + A.b0, B = A.B;
  const int len = min(A.lens[b], A.T);
  const int nmain = A.nmain;
  // Role order by hardware warp id: producers, io, main, scale.  The warp schedulers favour the
  // lower ids when several warps are ready, and the producers must never starve behind main warps
  // that are spinning on the very rows the producers have yet to deliver.
  const int hw = tid >> 5;
  const int warp = hw >= NPROD + 1 ? hw - (NPROD + 1) : hw + nmain + 1;
  const int rowf = A.rowf, estride = A.estride, prowf = A.prowf;
  const bool loops = A.loops != 0, ctx = A.ctx != 0;
  const SwSmem L = sweep_smem(rowf, estride, A.epad);
  const unsigned sm0 = smem_u32(smraw);
  const unsigned raw_a = sm0 + L.raw, eb_a = sm0 + L.eb, ed_a = sm0 + L.edb, rb_a = sm0 + L.rbuf,
                 kf_a = sm0 + L.kflag, bar_a = sm0 + L.mbar;
  auto raw_full = [&](int s) { return bar_a + s * 8; };
  auto raw_empty = [&](int s) { return bar_a + (NRAW + s) * 8; };
  auto e_full = [&](int s) { return bar_a + (2 * NRAW + s) * 8; };
  auto e_empty = [&](int s) { return bar_a + (2 * NRAW + NE + s) * 8; };
  const unsigned prod_done = bar_a + (2 * NRAW + 2 * NE) * 8;
  // Frame barrier of the recursion (main warps + scale warp) as an mbarrier, so that a warp can
  // ARRIVE as soon as its part of the new vector is stored and only WAIT after it has issued the
  // next frame's e loads: the prefetch leaves the per-frame chain.
  const unsigned frame_bar = bar_a + (2 * NRAW + 2 * NE + 1) * 8;
  float* spill = A.spill[DIR] + (long long)b * A.T * prowf;
  float* ck = A.ck[DIR] + (long long)b * A.nck * V;

  // ---- init ----
  {
    float* ebf = reinterpret_cast<float*>(smraw + L.eb);
    for (int i = tid; i < NE * estride + A.epad + NE * RBW; i += blockDim.x) ebf[i] = 0.f;  // e rows, pad, ed
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
        mbar_init(e_empty(s), nmain);
      }
      mbar_init(prod_done, NPROD);
      mbar_init(frame_bar, nmain + 1);
      asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
  }
  __syncthreads();
  auto frame_t = [&](int f) { return DIR == 0 ? f : len - 1 - f; };
  // Odd S: the e row keeps the layout of the landed 16-byte aligned window (class v at float r + v,
  // r = ((t*B + b) * V) mod 4), so the producers move it with 16-byte accesses; the pitch is S itself.
  constexpr bool WIN = (SCT > 0) && ((SCT & 1) != 0);
  const int B3_ = B & 3, b3_ = b & 3, V3_ = V & 3;
  auto row_r = [&](int t) { return ((((t & 3) * B3_ + b3_) & 3) * V3_) & 3; };

  if (warp < nmain) {
    // ============================ main warps: the recursion ============================
    const int g = warp * 8 + (lane & 7), k4 = lane >> 3;
    const bool gv = g < S;
    // byte offset of slot 0 inside an e stage; slot s adds the compile-time constant soff(s)
    const unsigned ebase = (DIR == 0) ? (unsigned)(8 * k4 * PF + g) * 4u : (unsigned)(g * PF + 8 * k4) * 4u;
    auto soff = [&](int s) { return (unsigned)((32 * (s >> 3) + (s & 7)) * (DIR == 0 ? PF : 1)) * 4u; };
    const unsigned edo = (unsigned)(gv ? g : 0) * 4u;
    const unsigned rc = (unsigned)(8 * k4) * 4u;  // this thread's R chunks: floats 8*k4 + {0..7} and + 32
    const float2 lf2 = splat2(loops ? 1.f : 0.f);
    float2 val[8];
#pragma unroll
    for (int h = 0; h < 8; ++h) val[h] = make_float2(0.f, 0.f);
    float Rown = (DIR == 0) ? (g == 0 ? pow2i(TARGET) : 0.f) : (gv ? pow2i(TARGET) : 0.f);
    float evA[16], evB[16], edA = 0.f, edB = 0.f;
    int es = 0;          // e stage of the frame being prefetched
    unsigned epar = 0;
    // steps until the values leaving a step are the checkpoint of the next frame (forward: frames 8j;
    // backward: frames with t = 8j + 7, the all-zero one at the sweep's first frame is never read)
    int ckc = !loops ? 0x7fffffff : (DIR == 0 ? KSEG : ((len & (KSEG - 1)) ? (len & (KSEG - 1)) : KSEG));
    int esc = 0;  // e stage of the frame being consumed
    int fpre = 0;        // frame being prefetched
    auto prefetch = [&](float (&ev)[16], float& ed) {
      mbar_wait(e_full(es), epar);
      const unsigned eb = eb_a + es * estride * 4 + ebase + (WIN ? row_r(frame_t(fpre)) * 4 : 0);
      ++fpre;
#pragma unroll
      for (int s = 0; s < 16; ++s) ev[s] = lds(eb + soff(s));
      ed = lds(ed_a + es * RBW * 4 + edo);
      if (++es == NE) { es = 0; epar ^= 1; }
    };
#ifdef CTCG_SWEEP_TIMERS
    long long mseg[5] = {0, 0, 0, 0, 0}, mprev = clock64();
    const bool mtim = (A.dbg & 256) && b == 0 && lane == 0 && warp == 0;
#endif
#ifdef CTCG_SWEEP_TIMERS
#define MSEG(i)                          \
  if (mtim) {                            \
    const long long now_ = clock64();    \
    mseg[i] += now_ - mprev;             \
    mprev = now_;                        \
  }
#else
#define MSEG(i)
#endif
    const unsigned rb_ld = rb_a + rc, rb_st = rb_a + g * 4;
    // PAR = f & 1 at compile time (the loop below is unrolled by two): buffer offsets become immediates
    auto step = [&](auto par_tag, int f, const float (&ev)[16], const float ed, float (&evn)[16], float& edn) {
      constexpr int PAR = decltype(par_tag)::value;
      if (f > 0) mbar_wait(frame_bar, PAR ^ 1);  // every group's entry of the vector entering frame f is stored
      MSEG(4)
      const unsigned rb = rb_ld + PAR * RBW * 4;
      const float4 r0 = lds4(rb), r1 = lds4(rb + 16), r2 = lds4(rb + 128), r3 = lds4(rb + 144);
      const int kk = ldsi(kf_a + (f & 3) * 4);
      MSEG(0)
      val[0] = fmul2(make_float2(ev[0], ev[1]), ffma2(lf2, val[0], make_float2(r0.x, r0.y)));
      val[1] = fmul2(make_float2(ev[2], ev[3]), ffma2(lf2, val[1], make_float2(r0.z, r0.w)));
      val[2] = fmul2(make_float2(ev[4], ev[5]), ffma2(lf2, val[2], make_float2(r1.x, r1.y)));
      val[3] = fmul2(make_float2(ev[6], ev[7]), ffma2(lf2, val[3], make_float2(r1.z, r1.w)));
      val[4] = fmul2(make_float2(ev[8], ev[9]), ffma2(lf2, val[4], make_float2(r2.x, r2.y)));
      val[5] = fmul2(make_float2(ev[10], ev[11]), ffma2(lf2, val[5], make_float2(r2.z, r2.w)));
      val[6] = fmul2(make_float2(ev[12], ev[13]), ffma2(lf2, val[6], make_float2(r3.x, r3.y)));
      val[7] = fmul2(make_float2(ev[14], ev[15]), ffma2(lf2, val[7], make_float2(r3.z, r3.w)));
      const float2 s2 = fadd2(fadd2(fadd2(val[0], val[1]), fadd2(val[2], val[3])),
                              fadd2(fadd2(val[4], val[5]), fadd2(val[6], val[7])));
      float tot = s2.x + s2.y;
      tot += __shfl_xor_sync(0xffffffffu, tot, 8);
      tot += __shfl_xor_sync(0xffffffffu, tot, 16);
      const float sc = pow2i(kk);  // power-of-two rescale decided two frames ago (1 almost always)
      const float Rn = fmaf(ed, Rown, tot) * sc;
      Rown = Rn;
      MSEG(2)
      if (k4 == 0 && gv) sts(rb_st + (PAR ^ 1) * RBW * 4, Rn);
      __syncwarp();
      if (lane == 0) {
        mbar_arrive(frame_bar);
        mbar_arrive(e_empty(esc));
      }
      MSEG(3)
      // ---- everything below is off the per-frame chain ----
      if (++esc == NE) esc = 0;
      if (kk != 0) {
        const float2 sc2 = splat2(sc);
#pragma unroll
        for (int h = 0; h < 8; ++h) val[h] = fmul2(val[h], sc2);
      }
      if (--ckc == 0) {  // the values entering frame f + 1 are a checkpoint (explicit self-loops only)
        ckc = KSEG;
        if (f + 1 < len) {
          const int t = frame_t(f + 1);
          float* cp = ck + (long long)(t / KSEG) * V;
#pragma unroll
          for (int s = 0; s < 16; ++s) {
            const int o = 32 * (s >> 3) + 8 * k4 + (s & 7);
            const bool valid = gv && o < S && (DIR == 0 ? g >= 1 : o >= 1);
            const int cls = DIR == 0 ? o * S + g : g * S + o;
            const float v = (s & 1) ? val[s >> 1].y : val[s >> 1].x;
            if (valid) cp[cls] = v;
          }
        }
      }
      if (f + 1 < len) prefetch(evn, edn);  // next frame's e values
      MSEG(1)
    };
#ifdef CTCG_SWEEP_TIMERS
    const long long mt0 = clock64();
#endif
    if (!(A.dbg & 1)) {
      if (len > 0) prefetch(evA, edA);
      int f = 0;
      for (; f + 1 < len; f += 2) {
        step(std::integral_constant<int, 0>{}, f, evA, edA, evB, edB);
        step(std::integral_constant<int, 1>{}, f + 1, evB, edB, evA, edA);
      }
      if (f < len) step(std::integral_constant<int, 0>{}, f, evA, edA, evB, edB);
    } else {  // timing experiment: hand-shakes and barriers only
      for (int f = 0; f < len; ++f) {
        mbar_wait(e_full(es), epar);
        if (++es == NE) { es = 0; epar ^= 1; }
        if (lane == 0) {
          mbar_arrive(e_empty(f % NE));
          mbar_arrive(frame_bar);
        }
        mbar_wait(frame_bar, f & 1);
      }
    }
#ifdef CTCG_SWEEP_TIMERS
    if (mtim) printf("sweep dir %d main: total %lld (len %d)  Rloads %lld  prefetch %lld  math %lld  store+arrive %lld  barrier %lld\n", DIR, clock64() - mt0, len, mseg[0], mseg[1], mseg[2], mseg[3], mseg[4]);
#endif
#undef MSEG
  } else if (warp == nmain) {
    // ============================ scale / spill warp ============================
    // Per frame: spill the vector entering the frame, decide the rescale of the frame after next from
    // its largest entry (one REDUX on the float bits: the entries are non-negative).  Nothing here
    // waits for anything but the frame barrier, so the main warps never wait for this warp.
    int kcur = 0, knext = 0, Kcum = TARGET;
    const float inf = __int_as_float(0x7f800000);
    bool dead = false;
    for (int f = 0; f < len; ++f) {
      if (f > 0) mbar_wait(frame_bar, (f - 1) & 1);
      if (!(A.dbg & 8)) {
        const int t = frame_t(f);
        const unsigned rb = rb_a + (f & 1) * RBW * 4;
        const float v0 = lds(rb + lane * 4), v1 = lds(rb + (lane + 32) * 4);
        const float mx = __int_as_float(__reduce_max_sync(0xffffffffu, max(__float_as_int(v0), __float_as_int(v1))));
        int knext2 = 0;
        if (mx > 0.f && mx < inf) {
          const int ex = exponent_of(mx) + kcur + knext;
          if (ex > TARGET + BAND || ex < TARGET - BAND) knext2 = max(-120, min(120, TARGET - ex));
        } else {
          dead = true;
        }
        if (lane == 0) stsi(kf_a + ((f + 2) & 3) * 4, knext2);
        __syncwarp();
        if (lane == 0) mbar_arrive(frame_bar);  // the vector is in registers: the spill below is off the chain
        float* sp = spill + (long long)t * prowf;
        if (lane < S) sp[lane] = v0;
        if (lane + 32 < S) sp[lane + 32] = v1;
        if (lane == 0) {
          sp[S + 1] = (float)kcur;
          sp[S + 2] = (float)Kcum;
        }
        Kcum += kcur;
        kcur = knext;
        knext = knext2;
      } else {
        __syncwarp();
        if (lane == 0) mbar_arrive(frame_bar);
      }
    }
    if (len > 0) mbar_wait(frame_bar, (len - 1) & 1);
    if (DIR == 0) {  // total mass = sum_c P_len[c]; cost = log(mass) - K ln2 + sum of the frame maxima
      const unsigned rb = rb_a + (len & 1) * RBW * 4;
      const float v0 = lds(rb + lane * 4), v1 = lds(rb + (lane + 32) * 4);
      const float s = warp_sum(v0 + v1);
      mbar_wait(prod_done, 0);
      if (lane == 0) {
        const double* mp = reinterpret_cast<const double*>(smraw + L.msum);
        double msum = 0.0;
        for (int i = 0; i < NPROD; ++i) msum += mp[i];
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
    double msum = 0.0;
    const bool skip_ld = (A.dbg & 16) != 0, skip_ex = (A.dbg & 2) != 0;
#ifdef CTCG_SWEEP_TIMERS
    long long seg[6] = {0, 0, 0, 0, 0, 0}, tprev = clock64();
    const bool ptim = (A.dbg & 256) && b == 0 && pw == 0 && lane == 0;
#endif
#ifdef CTCG_SWEEP_TIMERS
#define PSEG(i)                          \
  if (ptim) {                            \
    const long long now_ = clock64();    \
    seg[i] += now_ - tprev;              \
    tprev = now_;                        \
  }
#else
#define PSEG(i)
#endif
    for (int f = pw; f < len; f += NPROD) {
      const int t = frame_t(f);
      const int rs = f % NRAW, es = f % NE;
      PSEG(0)
      const int r = ((((t & 3) * B3 + b3) & 3) * V3) & 3;  // ((t*B + b) * V) mod 4
      const unsigned src = raw_a + (rs * rowf + r) * 4;
      const unsigned dst = eb_a + es * estride * 4;
      const int roff = WIN ? r : 0;
      mbar_wait(raw_full(rs), (f / NRAW) & 1);
      PSEG(1)
      float mx = neg_inf();
      if (WIN) {
        // 16-byte chunks of the window; floats that belong to the neighbouring rows become -inf (e = 0)
        constexpr int NQ4 = SCT > 0 ? ((SCT * SCT + 3 + 3) / 4 + 31) / 32 : 1;
        const unsigned srcw = raw_a + rs * rowf * 4;
        const int hi = r + V, nchw = (hi + 3) >> 2;
        float4 xq[NQ4];
#pragma unroll
        for (int q = 0; q < NQ4; ++q) {
          const int ch = q * 32 + lane;
          const float ninf = neg_inf();
          xq[q] = make_float4(ninf, ninf, ninf, ninf);
          if (!skip_ld && ch < nchw) xq[q] = lds4(srcw + ch * 16);
          if (q == 0 || q == NQ4 - 1) {
            const int w = 4 * ch;
            if (w < r || w >= hi) xq[q].x = ninf;
            if (w + 1 < r || w + 1 >= hi) xq[q].y = ninf;
            if (w + 2 < r || w + 2 >= hi) xq[q].z = ninf;
            if (w + 3 < r || w + 3 >= hi) xq[q].w = ninf;
          }
          mx = fmaxf(mx, fmaxf(fmaxf(xq[q].x, xq[q].y), fmaxf(xq[q].z, xq[q].w)));
        }
        mx = warp_max(mx);  // every lane's loads have landed in registers here
        if (lane == 0) mbar_arrive(raw_empty(rs));
        if (!(mx > neg_inf())) mx = 0.f;
        const float nmL = -mx * kLog2e;
        PSEG(2)
        if (f >= NE) mbar_wait(e_empty(es), ((f / NE) - 1) & 1);
        PSEG(3)
        if (!skip_ex) {
          const float2 l2 = splat2(kLog2e), n2 = splat2(nmL);
#pragma unroll
          for (int q = 0; q < NQ4; ++q) {
            const int ch = q * 32 + lane;
            const float2 y0 = ffma2(make_float2(xq[q].x, xq[q].y), l2, n2);
            const float2 y1 = ffma2(make_float2(xq[q].z, xq[q].w), l2, n2);
            if (ch < nchw) sts4(dst + ch * 16, make_float4(ex2(y0.x), ex2(y0.y), ex2(y1.x), ex2(y1.y)));
          }
        }
      } else if (SCT > 0) {
        float xr[NQ];
        if (!skip_ld) {
#pragma unroll
          for (int q = 0; q < NQ; ++q) {
            const int v = q * 32 + lane;
            xr[q] = (q * 32 + 31 < V || v < V) ? lds(src + v * 4) : neg_inf();
          }
        } else {
#pragma unroll
          for (int q = 0; q < NQ; ++q) xr[q] = 0.f;
        }
#pragma unroll
        for (int q = 0; q + 1 < NQ; q += 2) mx = fmaxf(mx, fmaxf(xr[q], xr[q + 1]));
        if (NQ & 1) mx = fmaxf(mx, xr[NQ - 1]);
        mx = warp_max(mx);  // every lane's loads have landed in registers here
        if (lane == 0) mbar_arrive(raw_empty(rs));
        if (!(mx > neg_inf())) mx = 0.f;
        const float nmL = -mx * kLog2e;
        PSEG(2)
        if (f >= NE) mbar_wait(e_empty(es), ((f / NE) - 1) & 1);
        PSEG(3)
        // class v = c*S + l lands at c*PF + l: the same index when S is odd (PF = S)
        if (!skip_ex)
#pragma unroll
        for (int q = 0; q < NQ; ++q) {
          const int v = q * 32 + lane;
          const float e = ex2(fmaf(xr[q], kLog2e, nmL));
          const int vp = (SCT & 1) ? v : v + v / (SCT > 0 ? SCT : 1);
          if (q * 32 + 31 < V || v < V) sts(dst + vp * 4, e);
        }
      } else {
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
        if (lane == 0) mbar_arrive(raw_empty(rs));
      }
      PSEG(4)
      // fold the blank column and the diagonal into ed[c] = e0[c] + e[c][c], then zero them in the row
      __syncwarp();
      float e0c[2], egg[2];
#pragma unroll
      for (int h = 0; h < 2; ++h) {
        const int c = lane + 32 * h;
        e0c[h] = egg[h] = 0.f;
        if (c < S) {
          e0c[h] = lds(dst + (roff + (ctx ? c * PF : 0)) * 4);
          if (c >= 1) egg[h] = lds(dst + (roff + c * PF + c) * 4);
        }
      }
      __syncwarp();
#pragma unroll
      for (int h = 0; h < 2; ++h) {
        const int c = lane + 32 * h;
        if (c < S) {
          sts(ed_a + (es * RBW + c) * 4, e0c[h] + egg[h]);
          sts(dst + (roff + c * PF) * 4, 0.f);
          sts(dst + (roff + c * PF + c) * 4, 0.f);
        }
      }
      if (lane == 0) {
        spill[(long long)t * prowf + S] = mx;  // frame maximum, read by the combine kernel
        msum += (double)mx;
      }
      __syncwarp();
      if (lane == 0) mbar_arrive(e_full(es));
      PSEG(5)
    }
#ifdef CTCG_SWEEP_TIMERS
    if (ptim) printf("sweep dir %d producer 0: loop-top %lld  wait_raw_full %lld  load+max %lld  wait_e_empty %lld  exp+store %lld  fixup+arrive %lld\n", DIR, seg[0], seg[1], seg[2], seg[3], seg[4], seg[5]);
#endif
#undef PSEG
    if (lane == 0) {
      reinterpret_cast<double*>(smraw + L.msum)[pw] = msum;
      mbar_arrive(prod_done);
    }
  } else {
    // ============================ io lane: TMA row loads ============================
    if (lane == 0) {
      const long long stepf = (DIR == 0 ? 1LL : -1LL) * (long long)B * V;
      long long i0 = ((long long)frame_t(0) * B + b) * V;
      for (int f = 0; f < len; ++f, i0 += stepf) {
        const int rs = f % NRAW;
        if (f >= NRAW) mbar_wait(raw_empty(rs), ((f / NRAW) - 1) & 1);
        const int r = (int)(i0 & 3);
        const long long w0 = i0 - r;
        const long long room = (A.total - w0) & ~3LL;
        const int nfl = pad4(r + V);
        const int nbulk = (int)(room < (long long)nfl ? room : (long long)nfl);
        const unsigned dst = raw_a + rs * rowf * 4;
        // <= 3 floats at the very end of the tensor do not fill a 16-byte chunk: plain stores, made
        // visible to the producers by the release of the arrive below
        for (int idx = nbulk; idx < r + V; ++idx) sts(dst + idx * 4, A.x[w0 + idx]);
        if (A.dbg & 32) {
          mbar_arrive(raw_full(rs));
          continue;
        }
        fence_async_smem();
        mbar_expect_tx(raw_full(rs), nbulk * 4);
        bulk_g2s(dst, A.x + w0, nbulk * 4, raw_full(rs));
      }
    }
  }
}

template <int SCT, int MAXT>
__global__ void __launch_bounds__(MAXT, CTCG_SWEEP_MINB) bisweep_kernel(const BisweepArgs A) {
  extern __shared__ __align__(128) unsigned char smraw[];
  if (blockIdx.y == 0) sweep_body<SCT, 0>(A, smraw);
  else sweep_body<SCT, 1>(A, smraw);
}

// ---- combine: posterior rows of one K-frame segment of one utterance -----------------------
// Per class v = (c, l) the K frames of the segment are two scalar recurrences (forward for a, backward
// for f) driven by the spilled vectors; every class reads P[pi] and Q[qi] with
//   l >= 1           pi = c, qi = l, explicit loop iff l != c
//   l == 0, ctx      pi = c, qi = c           (blank state of context c, class c*S)
//   l == 0, global   class 0: pi -> 1, qi -> sum_c P[c] Q[c];  classes c*S, c >= 1: pi -> 0 (never emitted)
// so the inner loops are branch free.  KW = floats per staged vector row.
constexpr int KW = 72;
template <int SCT, int KMAX>
__global__ void __launch_bounds__(CTHREADS, 2) bicombine_kernel(const BisweepArgs A) {
  const int S = SCT > 0 ? SCT : A.S;
  const int V = S * S;
  const int b = blockIdx.y + A.b0, j = blockIdx.x, K = A.K, B = A.B, T = A.T, prowf = A.prowf;
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

  // Numerator posteriors of this segment (compact, segment-major: the K frames of a label are one 32-byte
  // sector): requested first, so that their latency runs behind the staging below; they go to shared memory
  // with a class -> label-index table and leave in the store of the dense row (NLT labels at most; an
  // utterance with more takes the read-modify-write pass at the end of the kernel).
  constexpr int NLT = 256, VMAX = 64 * 64;
  __shared__ unsigned short dl[VMAX];
  __shared__ __align__(16) float gtab[NLT][KMAX];
  int fold_nl = 0, fold_lo = 0;
  if (A.num_gg) {
    const int gi = A.num_graphs == 1 ? 0 : b;
    fold_lo = A.num_lab_off[gi];
    fold_nl = A.num_lab_off[gi + 1] - fold_lo;
  }
  const bool fold_smem = fold_nl > 0 && fold_nl <= NLT && V <= VMAX;
  int my_lab = -1;
  float4 my_g0 = make_float4(0.f, 0.f, 0.f, 0.f), my_g1 = my_g0;
  if (fold_smem && tid < fold_nl) {
    my_lab = A.num_lab_list[fold_lo + tid];
    const float4* gp = reinterpret_cast<const float4*>(A.num_gg + (((long long)b * A.nck + j) * A.num_nlp + tid) * KMAX);
    my_g0 = __ldg(gp);
    my_g1 = __ldg(gp + 1);
  }
  if (fold_smem)
    for (int v = tid; v < V; v += CTHREADS) dl[v] = 0xffffu;
  __shared__ float Ps[KMAX][KW], Qs[KMAX][KW];
  __shared__ float nmL_s[KMAX], scF_s[KMAX], scB_s[KMAX], cB_s[KMAX];
  const float* spF = A.spill[0] + ((long long)b * T + t0) * prowf;
  const float* spB = A.spill[1] + ((long long)b * T + t0) * prowf;
  const float gsc = A.g_scale ? A.g_scale[b] : 1.f;
  // gamma = a^ b^ 2^E_t / mass with E_t = K_len - Ka_{t+1} - Kb_{t+1}; split into
  // cA = gscale / mass * 2^E1 (about 2^-TARGET, rides on the a recurrence) and cB_t = 2^(E_t - E1)
  const int E1 = max(-120, min(120, -TARGET - exponent_of(fin.inv_mass)));
  const float cA = fin.inv_mass * pow2i(E1) * gsc;
  for (int i = tid; i < KMAX * KW; i += CTHREADS) {
    const int f = i / KW, c = i - f * KW;
    float pv = 0.f, qv = 0.f;
    if (f < n && c < S) {
      pv = spF[f * prowf + c] * cA;
      qv = spB[f * prowf + c];
    }
    if (c == 64) pv = 1.f;
    Ps[f][c] = pv;
    Qs[f][c] = qv;
  }
  if (tid < KMAX) {
    float nm = 0.f, sF = 0.f, sB = 0.f, cb = 0.f;
    if (tid < n) {
      const float* rf = spF + tid * prowf;
      const float* rq = spB + tid * prowf;
      nm = -rf[S] * kLog2e;
      sF = pow2i((int)rf[S + 1]);
      sB = pow2i((int)rq[S + 1]);
      const int Et = fin.kfinal - ((int)rf[S + 2] + (int)rf[S + 1]) - (int)rq[S + 2];
      cb = pow2i(max(-126, min(126, Et - E1)));
    }
    nmL_s[tid] = nm;
    scF_s[tid] = sF;
    scB_s[tid] = sB;
    cB_s[tid] = cb;
  }
  __syncthreads();
  if (fold_smem && my_lab >= 0) {  // after the barrier: the table has been cleared
    dl[my_lab] = (unsigned short)tid;
    *reinterpret_cast<float4*>(&gtab[tid][0]) = my_g0;
    *reinterpret_cast<float4*>(&gtab[tid][4]) = my_g1;
  }
  const bool loops = A.loops != 0, ctx = A.ctx != 0;
  if (fold_smem && ctx) __syncthreads();  // (the global-blank branch below has a barrier of its own)
  if (!ctx) {  // one global blank: sum over the S blank states, read by class 0 through Qs[f][64]
    for (int f = warp; f < n; f += CTHREADS / 32) {
      float d = Ps[f][lane] * Qs[f][lane] + Ps[f][lane + 32] * Qs[f][lane + 32];
      d = warp_sum(d);
      if (lane == 0) Qs[f][64] = d;
    }
    __syncthreads();
  }
  float nmL[KMAX], scF[KMAX], scB[KMAX], cB[KMAX];
#pragma unroll
  for (int f = 0; f < KMAX; ++f) {
    nmL[f] = nmL_s[f];
    scF[f] = scF_s[f];
    scB[f] = scB_s[f];
    cB[f] = cB_s[f];
  }
  const float* x0 = A.x + ((long long)t0 * B + b) * V;
  float* g0 = A.grad + ((long long)t0 * B + b) * V;
  const long long rstride = (long long)B * V;
  const float* ckA = A.ck[0] + ((long long)b * A.nck + j) * V;
  const float* ckF = A.ck[1] + ((long long)b * A.nck + j) * V;
  const bool haveA = j > 0, haveF = t1 < len;
  // One class over the (at most KMAX) frames of the segment: e[] holds the logits on entry, the
  // posteriors on exit.
  auto walk = [&](int v, float (&e)[KMAX], const bool full) {
    const int c = v / S, l = v - c * S;
    int pi = c, qi = l;
    bool m = loops && l != c;
    if (l == 0) {
      m = false;
      qi = c;
      if (!ctx) {
        qi = 64;
        pi = (v == 0) ? 64 : 65;
      }
    }
    const float mf = m ? 1.f : 0.f;
    // the checkpoints hold the sweeps' scaled values; cA rides on the forward recurrence (Ps has it too)
    float a = (m && haveA) ? ckA[v] * cA : 0.f;
    float fv = (m && haveF) ? ckF[v] : 0.f;
    float ah[KMAX];
#pragma unroll
    for (int f = 0; f < KMAX; ++f) {
      if (full || f < n) {
        e[f] = ex2(fmaf(e[f], kLog2e, nmL[f]));
        ah[f] = (e[f] * scF[f]) * fmaf(mf, a, Ps[f][pi]);   // a_{t+1}[c][l] * cA
        a = ah[f];
      }
    }
#pragma unroll
    for (int f = KMAX - 1; f >= 0; --f) {
      if (full || f < n) {
        const float bh = fmaf(mf, fv, Qs[f][qi]);            // b_{t+1}[c][l]
        fv = (e[f] * scB[f]) * bh;                           // f_t[c][l]
        e[f] = ah[f] * (bh * cB[f]);
      }
    }
  };
  auto sub_num = [&](int v, float (&e)[KMAX]) {
    if (!fold_smem) return;
    const unsigned jl = dl[v];
    if (jl == 0xffffu) return;
    const float4 q0 = *reinterpret_cast<const float4*>(&gtab[jl][0]);
    const float4 q1 = *reinterpret_cast<const float4*>(&gtab[jl][4]);
    const float gv[8] = {q0.x, q0.y, q0.z, q0.w, q1.x, q1.y, q1.z, q1.w};
#pragma unroll
    for (int f = 0; f < KMAX; ++f) e[f] = fmaf(-gsc, gv[f], e[f]);  // frames >= n of a partial segment are not stored
  };
  auto scalar_class = [&](int v) {
    float e[KMAX];
#pragma unroll
    for (int f = 0; f < KMAX; ++f)
      if (f < n) e[f] = __ldg(x0 + v + f * rstride);
    walk(v, e, false);
    sub_num(v, e);
#pragma unroll
    for (int f = 0; f < KMAX; ++f)
      if (f < n) g0[v + f * rstride] = e[f];
  };
  // Rows start on arbitrary 4-byte boundaries.  When B*V is a multiple of 4 every frame of the
  // segment has the same offset inside its 16-byte window, so a thread can take the same aligned
  // 4-class chunk of all KMAX rows with 16-byte accesses; the <= 3 classes before the first and
  // after the last full chunk (and every partial segment) go through the scalar path.
  const long long i0 = ((long long)t0 * B + b) * V;
  const bool vec = (n == KMAX) && ((rstride & 3) == 0);
  const int v0 = (int)((4 - (i0 & 3)) & 3);
  const int nch = vec ? (V - v0) >> 2 : 0;
  if (!vec)
    for (int v = tid; v < V; v += CTHREADS) scalar_class(v);
  for (int k = tid; k < nch; k += CTHREADS) {
    const int v = v0 + 4 * k;
    float4 xv[KMAX];
#pragma unroll
    for (int f = 0; f < KMAX; ++f) {
      const float4* p = reinterpret_cast<const float4*>(x0 + v + f * rstride);
      asm volatile("ld.global.cs.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(xv[f].x), "=f"(xv[f].y), "=f"(xv[f].z), "=f"(xv[f].w) : "l"(p));
    }
    float e[KMAX];
#pragma unroll
    for (int f = 0; f < KMAX; ++f) e[f] = xv[f].x;
    walk(v, e, true);
    sub_num(v, e);
#pragma unroll
    for (int f = 0; f < KMAX; ++f) { xv[f].x = e[f]; e[f] = xv[f].y; }
    walk(v + 1, e, true);
    sub_num(v + 1, e);
#pragma unroll
    for (int f = 0; f < KMAX; ++f) { xv[f].y = e[f]; e[f] = xv[f].z; }
    walk(v + 2, e, true);
    sub_num(v + 2, e);
#pragma unroll
    for (int f = 0; f < KMAX; ++f) { xv[f].z = e[f]; e[f] = xv[f].w; }
    walk(v + 3, e, true);
    sub_num(v + 3, e);
#pragma unroll
    for (int f = 0; f < KMAX; ++f) {
      xv[f].w = e[f];
      float4* p = reinterpret_cast<float4*>(g0 + v + f * rstride);
      asm volatile("st.global.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(p), "f"(xv[f].x), "f"(xv[f].y), "f"(xv[f].z), "f"(xv[f].w) : "memory");
    }
  }
  if (vec) {
    if (tid < v0) scalar_class(tid);
    const int vt = v0 + 4 * nch + (tid - 32);
    if (tid >= 32 && vt < V) scalar_class(vt);
  }
  // Numerator posteriors: grad[t][b][lab_j] -= g_scale * gamma_num[t][j] for the ~140 labels of this utterance,
  // on the rows this CTA has just written (they are in L2: a read-modify-write that never reaches DRAM, where
  // the separate scatter kernel of round 1 moved 17 bytes of sectors per useful byte).  Segment-major compact
  // rows: the K frames of a label are one 32-byte sector.
  if (A.num_gg == nullptr || fold_smem) return;
  __syncthreads();  // every dense store of the segment is visible to the CTA
  const int gi = A.num_graphs == 1 ? 0 : b;
  const int lo = A.num_lab_off[gi], nl = A.num_lab_off[gi + 1] - lo;
  const float* ngg = A.num_gg + ((long long)b * A.nck + j) * A.num_nlp * KMAX;
  for (int jl = tid; jl < nl; jl += CTHREADS) {
    const int v = A.num_lab_list[lo + jl];
    const float4* gp = reinterpret_cast<const float4*>(ngg + jl * KMAX);
    float gv[KMAX], cur[KMAX];
#pragma unroll
    for (int h = 0; h < KMAX / 4; ++h) {
      const float4 q = __ldg(gp + h);
      gv[4 * h] = q.x, gv[4 * h + 1] = q.y, gv[4 * h + 2] = q.z, gv[4 * h + 3] = q.w;
    }
#pragma unroll
    for (int f = 0; f < KMAX; ++f)
      if (f < n) cur[f] = __ldcg(g0 + v + f * rstride);
#pragma unroll
    for (int f = 0; f < KMAX; ++f)
      if (f < n) g0[v + f * rstride] = fmaf(-gsc, gv[f], cur[f]);
  }
}

}  // namespace

// ---- host side ------------------------------------------------------------------
BisweepPlan bisweep_plan(int T, int B, int V, int S) {
  BisweepPlan p{};
  p.ok = 0;
  if (S < 4 || S > 64 || V != S * S || T < 1 || B < 1) return p;
  p.nmain = (S + 7) / 8;
  p.threads = (p.nmain + 1 + NPROD + 1) * 32;
  p.rowf = pad4(V + 3);
  p.prowf = pad4(S + 3);
  const int PF = S | 1;
  p.estride = pad4(S * PF + 3);
  p.epad = 16 * PF + 64;  // zeros behind the last stage: states o >= S read up to 15 rows past a stage
  p.K = KSEG;
  p.nck = (T + p.K - 1) / p.K + 1;
  p.smem_bytes = sweep_smem(p.rowf, p.estride, p.epad).bytes;
  if (p.smem_bytes > 227 * 1024 || p.threads > 448) return p;
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
  const dim3 grid(a.nb > 0 ? a.nb : a.B, (a.grad && !(a.dbg & 4)) ? 2 : 1);
  {
    ProfScope prof("bisweep", st);
#define CTCG_SWEEP(SCT, MAXT)                                                                               \
  do {                                                                                                   \
    static int attr_cache_##SCT[64] = {};                                                                 \
    e = smem_attr_once(bisweep_kernel<SCT, MAXT>, (int)p.smem_bytes, attr_cache_##SCT);                   \
    if (e != cudaSuccess) return e;                                                                      \
    bisweep_kernel<SCT, MAXT><<<grid, p.threads, p.smem_bytes, st>>>(a);                                \
  } while (0)
#define MAXT_OF(S_) ((((S_) + 7) / 8 + 1 + NPROD + 1) * 32)
    if (a.S == 49) CTCG_SWEEP(49, MAXT_OF(49));       // the WSJ character set (egs/wsj/vocabulary.txt)
    else if (a.S == 36) CTCG_SWEEP(36, MAXT_OF(36));  // BASELINE.json's nominal V = 1296
    else CTCG_SWEEP(0, 448);
#undef MAXT_OF
#undef CTCG_SWEEP
    ++g_launches;
    e = cudaGetLastError();
    if (e != cudaSuccess) return e;
  }
  return e;
}

cudaError_t launch_bicombine(const BisweepArgs& a, const BisweepPlan& p, cudaStream_t st) {
  (void)p;
  if (!a.grad) return cudaSuccess;
  ProfScope prof("bicombine", st);
  const dim3 cgrid((a.T + a.K - 1) / a.K, a.nb > 0 ? a.nb : a.B);
  if (a.S == 49) bicombine_kernel<49, 8><<<cgrid, CTHREADS, 0, st>>>(a);
  else if (a.S == 36) bicombine_kernel<36, 8><<<cgrid, CTHREADS, 0, st>>>(a);
  else bicombine_kernel<0, 8><<<cgrid, CTHREADS, 0, st>>>(a);
  ++g_launches;
  return cudaGetLastError();
}

}  // namespace ctcg
