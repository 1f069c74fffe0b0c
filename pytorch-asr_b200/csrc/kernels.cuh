// Launch wrappers of the CUDA kernels (definitions in kernels_*.cu).
#pragma once

#include <cuda_runtime.h>

#include "ctcg_internal.h"

namespace ctcg {

// Arguments of the generic log-semiring forward-backward over CSR graphs.
struct PathArgs {
  GraphDev g;
  const float* x;         // scores; element (t, b, l) at x[t*x_stride_t + b*x_stride_b + l]
                          // (split numerator: int2 pairs written by launch_gather(pairs), strides in entries)
  long long x_stride_t;
  long long x_stride_b;
  float x_scale;          // multiplies x on load (log2(e) for natural-log input, 1 if prescaled)
  const int32_t* lens;    // device, [B]
  int32_t T, B;
  float* alpha_ws;        // [B][T][row]: generic kernel: normalised alpha_t in log2 units + frame offset (2 floats);
                          // split numerator: (int32 integer part, fp32 fraction) per state
  int32_t alpha_row;      // floats per row: pad4(2 * max_states + 2)
  float* beta_ws;         // nullable; same shape as alpha_ws: beta rows of the split numerator (kernels_num.cu)
  double* cost;           // [B] natural-log path sum
  float* gout;            // posterior rows; element (t, b, l) at gout[t*g_stride_t + b*g_stride_b + l]
  long long g_stride_t;
  long long g_stride_b;
  int32_t g_width;        // row width written per frame (max_labels or V)
  int32_t g_seg;          // split numerator only: > 0 = posteriors stored segment-major, gout[b][t / g_seg][j][t % g_seg]
                          // (g_stride_b = entries per utterance), the layout the denominator's combine reads
  float g_sign;
  const float* g_scale;   // nullable, per-utterance multiplier
  int32_t write_grad;
  int32_t dbg;            // reserved for timing experiments; 0 in production
  int32_t alpha_rows_plus1;  // generic kernel: alpha_ws holds T + 1 rows per utterance (lets it take the transposed-table mode)
};

// compact = 1: x / gout are the per-utterance compact label columns produced by
// launch_gather (16-byte aligned rows, prescaled), staged through a cp.async ring.
// cudaFuncAttributeMaxDynamicSharedMemorySize once per (kernel instantiation, device, size): the attribute call
// costs about as much as a launch and the step has ~20 of them.  `cache` is a static int[64] of the call site,
// initialised to zero (the attribute is per device).
template <class Kernel>
inline cudaError_t smem_attr_once(Kernel kernel, int bytes, int* cache) {
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return e;
  if (dev < 0 || dev >= 64) return cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
  if (cache[dev] >= bytes + 1) return cudaSuccess;
  e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
  if (e == cudaSuccess) cache[dev] = bytes + 1;
  return e;
}

cudaError_t launch_path_fwd_bwd(const PathArgs& a, bool compact, cudaStream_t st);

// Numerator on the compact columns (kernels_num.cu): alpha and beta sweeps on separate CTAs at the same
// time, then a time-parallel combine; same arguments as the compact mode above.
bool num_split_supported(const PathArgs& a);
cudaError_t launch_num_sweeps(const PathArgs& a, cudaStream_t st);
// Posteriors from the spilled rows; with a.scatter_grad set they are subtracted from the dense gradient directly.
cudaError_t launch_num_combine(const PathArgs& a, cudaStream_t st);

// xg[b][t][j] = logits[t][b][lab_list_b[j]] for t < lens[b] (minus the context block's LSE when normalize == 2);
// pairs: as int2 {rint(x log2 e), bits of 2^frac} for the split numerator sweeps (8 bytes per entry)
cudaError_t launch_gather(const float* logits, int T, int B, int V, const int32_t* lens,
                          const GraphDev& num, float* xg, int nlp, int normalize, int S, bool pairs,
                          cudaStream_t st);

// Row pass: mode 1 grad = softmax(x) - scatter(gg), rowlse = logsumexp(x);
//           mode 0 grad = -scatter(gg).  Rows t >= lens[b] are zeroed.
cudaError_t launch_rows(const float* logits, int T, int B, int V, const int32_t* lens,
                        const GraphDev& num, const float* gg, int nlp, int mode,
                        const float* g_scale, float* grad, float* rowlse, cudaStream_t st);

// grad[t][b][lab_list_b[j]] -= scale_b * gg[b][t][j]
cudaError_t launch_ctx_norm(bool backward, const float* x, float* lp, float* grad, int T, int B, int V, int S,
                            const int32_t* lens, cudaStream_t st);
cudaError_t launch_rows_ctx(const float* logits, int T, int B, int V, int S, const int32_t* lens, const GraphDev& num,
                            const float* gg, int nlp, const float* g_scale, float* grad, cudaStream_t st);
cudaError_t launch_post(const float* in, long long rows, int V, int S, int mode, int transfer, int tile_normalize,
                        float* out, cudaStream_t st);
cudaError_t launch_scatter_sub(int T, int B, int V, const int32_t* lens, const GraphDev& num,
                               const float* gg, int nlp, const float* g_scale, float* grad,
                               cudaStream_t st);

// loss[b] = -cost_num[b] + cost_den[b] + sum_{t<len} rowlse[t][b]   (nullable terms skipped)
cudaError_t launch_final(int T, int B, const int32_t* lens, const double* cost_num,
                         const double* cost_den, const float* rowlse, float* loss, cudaStream_t st);

// grad[t][b][:] *= scale[b * stride] (stride 0: one scalar for the whole batch); rows with scale 1 are skipped
cudaError_t launch_scale_grad(float* grad, int T, int B, int V, const float* scale, int stride, cudaStream_t st);

cudaError_t launch_cost_to_float(int B, const double* cost, float* out, cudaStream_t st);

cudaError_t launch_viterbi(const float* x, int T, int B, int V, const int32_t* lens,
                           const GraphDev& g, int32_t* back_ws, int row, float* score,
                           int32_t* frame_labels, cudaStream_t st);

// Viterbi on the bi-char decoding graph, factorised (O(V) per frame); back_ws: B*T*V uint16
cudaError_t launch_biviterbi(const float* x, int T, int B, int S, const int32_t* lens, int loops, int ctx,
                             unsigned short* back_ws, float* score, int32_t* frame_labels, cudaStream_t st);

cudaError_t launch_path_count(int B, const int32_t* lens, const GraphDev& g, uint64_t* counts,
                              cudaStream_t st);

// Split-sweep bi-char denominator (kernels_bisweep.cu): forward and backward recursions on
// separate CTAs at the same time, then an elementwise combine over the whole GPU.
struct BiFin {
  float inv_mass;   // 1 / (scaled total mass); 0 when no path is accepted
  int32_t kfinal;   // exponent of the scaled total mass
  int32_t dead;     // no accepting path: cost -1e20, zero gradient
  int32_t pad;
};
struct BisweepArgs {
  const float* x;          // logits (T, B, V), raw
  long long total;         // T*B*V
  int32_t T, B, V, S;
  const int32_t* lens;     // device [B]
  int32_t loops, ctx;      // allow_nonblank_selfloops, use_contextual_blanks (after loop_using_symbol_repetitions)
  float* spill[2];         // [B][T][prowf]: vector entering frame t (P_t / Q_{t+1}), frame max, k_t, exponent
  float* ck[2];            // [B][nck][V]: per-state values every K frames (explicit self-loops only)
  int32_t prowf, nck, K;
  double* cost;            // [B] natural-log path sum over the decoding graph
  BiFin* fin;              // [B]
  float* grad;             // (T, B, V) nullable: receives g_scale[b] * gamma_den
  const float* g_scale;    // nullable
  int32_t nmain, rowf, estride, epad;
  int32_t dbg;             // timing experiments only (CTCG_SWEEP_DEBUG bit mask); 0 in production
  // a launch may cover the utterances [b0, b0 + nb) only (nb = 0: the whole batch): groups of a length-sorted
  // batch on streams of their own, so that a group's combine starts when ITS sweeps are done (abi.cu)
  int32_t b0, nb;
  // numerator posteriors folded into the combine (nullable): grad = g_scale * (gamma_den - gamma_num) in ONE store.
  const float* num_gg;            // compact posteriors, segment-major: gg[b][t / K][j][t % K], j indexes the utterance's labels
  int32_t num_nlp;                // labels per row (pitch)
  int32_t num_graphs;             // 1 (shared numerator graph) or B
  const int32_t* num_lab_off;     // GraphDev::lab_off / lab_list of the numerator graphs
  const int32_t* num_lab_list;
};
struct BisweepPlan {
  int32_t ok, K, nck, nmain, threads, rowf, prowf, estride, epad;
  size_t smem_bytes, spill_bytes, ckpt_bytes;
};
BisweepPlan bisweep_plan(int T, int B, int V, int S);
// The two sweeps (the backward one only when a.grad is set), then the combine as a launch of its own: with
// a.num_gg set it must be ordered after the numerator's compact rows are complete.
cudaError_t launch_bisweep(const BisweepArgs& a, const BisweepPlan& p, cudaStream_t st);
cudaError_t launch_bicombine(const BisweepArgs& a, const BisweepPlan& p, cudaStream_t st);

// Incremented by every launch_* above.
extern unsigned long long g_launches;

// Optional per-launch timing (CUDA events on the launch stream), off by default;
// bench.py turns it on for a few extra steps after the timed region to attribute
// the step time to kernels.  Never enabled inside a timed region.
struct ProfScope {
  ProfScope(const char* name, cudaStream_t st);
  ~ProfScope();
  cudaStream_t st_;
  int idx_;
};
void profile_enable(bool on);
bool profile_is_on();
int profile_collect(char* names, size_t cap, float* ms, int max_records);

}  // namespace ctcg
