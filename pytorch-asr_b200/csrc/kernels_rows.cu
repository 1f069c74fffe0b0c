// Streaming row kernels around the recursion: label-column gather, the fused
// log_softmax + gradient row pass, sparse numerator subtraction, loss combine.
//
// Reference counterparts: get_normalized_acts (modules/ctc_losses.py:29-43),
// the max-shift / mask / numerator-minus-denominator glue of
// FSTDecoder.get_fst_loss (modules/decoders/advanced_decoder.py:479-517) and the
// dense zero-initialised logprobs_grad + scatter_add_ of fst_utils.py:448,468-470.
#include <algorithm>

#include "device_utils.cuh"
#include "kernels.cuh"

namespace ctcg {

namespace {

constexpr int NT = 256;
constexpr int NWARP = NT / 32;

// ---------------------------------------------------------------------------
// xg[b][t][j] = normalise(logits[t][b][:])[lab_list_b[j]] (natural log); one warp per (t, b) row.
// normalize: 0/1 -> raw logits (a softmax normaliser is per-frame and is added
// to the loss by the row pass instead); 2 -> minus the LSE of the label's context block.
// PAIRS: the split numerator sweeps (kernels_num.cu) take every value as 2^(x log2 e) = f * 2^E, an int32
// exponent E = rint(x log2 e) and an fp32 mantissa f = 2^(x log2 e - E): xg holds int2 {E, bits of f}.
template <bool PAIRS>
__global__ void __launch_bounds__(NT) gather_kernel(const float* __restrict__ logits, int T, int B, int V,
                                                    const int32_t* __restrict__ lens, GraphDev g,
                                                    float* __restrict__ xg, int nlp, int normalize, int S) {
  const int lane = threadIdx.x & 31;
  const long long row = (long long)blockIdx.x * NWARP + (threadIdx.x >> 5);
  if (row >= (long long)T * B) return;
  const int t = (int)(row / B), b = (int)(row % B);
  if (t >= lens[b]) return;
  const int gi = g.num_graphs == 1 ? 0 : b;
  const int lo = g.lab_off[gi], nl = g.lab_off[gi + 1] - lo;
  const float* x = logits + row * V;
  float* dst = xg + ((long long)b * T + t) * nlp * (PAIRS ? 2 : 1);
  for (int j = lane; j < (PAIRS ? nlp : nl); j += 32) {
    if (PAIRS && j >= nl) {  // padding labels: 2^0
      reinterpret_cast<int2*>(dst)[j] = make_int2(0, __float_as_int(1.f));
      continue;
    }
    const int lab = g.lab_list[lo + j];
    float v = x[lab];
    if (normalize == 2) {
      const float* blk = x + (lab / S) * S;
      float m = neg_inf();
      for (int k = 0; k < S; ++k) m = fmaxf(m, blk[k]);
      float s = 0.f;
      for (int k = 0; k < S; ++k) s += __expf(blk[k] - m);
      v -= m + __logf(s);
    }
    if (PAIRS) {
      const float xs = fminf(fmaxf(v * kLog2e, -65536.f), 65536.f);  // T * 2^16 stays inside the sweeps' exponent range
      const float r = xs + 12582912.f;                               // 1.5 * 2^23: rint(xs) in the low mantissa bits
      reinterpret_cast<int2*>(dst)[j] = make_int2(__float_as_int(r) - 0x4B400000, __float_as_int(ex2(xs - (r - 12582912.f))));
    } else {
      dst[j] = v;
    }
  }
}

// ---------------------------------------------------------------------------
__device__ __forceinline__ float block_max(float v, float* red) {
  v = warp_max(v);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = v;
  __syncthreads();
  float m = red[0];
#pragma unroll
  for (int i = 1; i < NWARP; ++i) m = fmaxf(m, red[i]);
  __syncthreads();
  return m;
}
__device__ __forceinline__ float block_sum(float v, float* red) {
  v = warp_sum(v);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = v;
  __syncthreads();
  float s = 0.f;
#pragma unroll
  for (int i = 0; i < NWARP; ++i) s += red[i];
  __syncthreads();
  return s;
}

// One CTA per row at a time (grid-stride).  The row lives in registers
// (PER values per thread); delta[V] in shared memory carries the sparse
// numerator posteriors so that the row is read once and written once.
template <int PER>
__global__ void __launch_bounds__(NT) rows_kernel(const float* __restrict__ logits, int T, int B, int V,
                                                  const int32_t* __restrict__ lens, GraphDev g,
                                                  const float* __restrict__ gg, int nlp, int mode,
                                                  const float* __restrict__ g_scale, float* __restrict__ grad,
                                                  float* __restrict__ rowlse) {
  extern __shared__ __align__(16) float delta[];  // [V] + red[NWARP]
  float* red = delta + pad4(V);
  const int tid = threadIdx.x;
  for (int j = tid; j < V; j += NT) delta[j] = 0.f;
  __syncthreads();
  const long long rows = (long long)T * B;
  for (long long row = blockIdx.x; row < rows; row += gridDim.x) {
    const int t = (int)(row / B), b = (int)(row % B);
    float* out = grad ? grad + row * V : nullptr;
    if (t >= lens[b]) {
      if (out)
        for (int j = tid; j < V; j += NT) st_stream(out + j, 0.f);
      continue;
    }
    const float scale = g_scale ? g_scale[b] : 1.f;
    const int gi = g.num_graphs == 1 ? 0 : b;
    const int lo = g.lab_off[gi], nl = g.lab_off[gi + 1] - lo;
    const float* grow = gg + ((long long)b * T + t) * nlp;
    float v[PER];
    float inv = 0.f;
    if (mode == 1) {
      const float* x = logits + row * V;
      float m = neg_inf();
#pragma unroll
      for (int i = 0; i < PER; ++i) {
        const int j = tid + i * NT;
        v[i] = (j < V) ? ld_stream(x + j) : neg_inf();
        m = fmaxf(m, v[i]);
      }
      m = block_max(m, red);  // its barriers separate the previous row's delta clear from the fill below
      if (out)
        for (int j = tid; j < nl; j += NT) delta[g.lab_list[lo + j]] = grow[j];
      float s = 0.f;
#pragma unroll
      for (int i = 0; i < PER; ++i) {
        v[i] = ex2((v[i] - m) * kLog2e);
        s += v[i];
      }
      s = block_sum(s, red);  // its barriers also publish delta
      inv = 1.f / s;
      if (tid == 0) rowlse[row] = m + __logf(s);
    } else {
      __syncthreads();
      for (int j = tid; j < nl; j += NT) delta[g.lab_list[lo + j]] = grow[j];
      __syncthreads();
#pragma unroll
      for (int i = 0; i < PER; ++i) v[i] = 0.f;
    }
#pragma unroll
    for (int i = 0; i < PER; ++i) {
      const int j = tid + i * NT;
      if (out && j < V) st_stream(out + j, (v[i] * inv - delta[j]) * scale);
    }
    __syncthreads();
    for (int j = tid; j < nl; j += NT) delta[g.lab_list[lo + j]] = 0.f;
  }
}

// normalize_by_dim = 1 (get_normalized_acts, ctc_losses.py:36-40): log_softmax inside each
// context block of S labels.  With lp = x - LSE_block(x) the chain rule gives
//   dL/dx[v] = softmax_block(x)[v] * sum_{u in block(v)} gamma[u] - gamma[v],   L = -cost_num(lp).
// One CTA per row at a time, one warp per context block; gamma is sparse (delta[V] in shared memory).
__global__ void __launch_bounds__(NT) rows_ctx_kernel(const float* __restrict__ logits, int T, int B, int V, int S,
                                                      const int32_t* __restrict__ lens, GraphDev g,
                                                      const float* __restrict__ gg, int nlp,
                                                      const float* __restrict__ g_scale,
                                                      float* __restrict__ grad) {
  extern __shared__ __align__(16) float delta[];  // [V]
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  for (int j = tid; j < V; j += NT) delta[j] = 0.f;
  __syncthreads();
  const long long rows = (long long)T * B;
  const int nblk = V / S;
  for (long long row = blockIdx.x; row < rows; row += gridDim.x) {
    const int t = (int)(row / B), b = (int)(row % B);
    float* out = grad + row * V;
    if (t >= lens[b]) {
      for (int j = tid; j < V; j += NT) st_stream(out + j, 0.f);
      continue;
    }
    const float scale = g_scale ? g_scale[b] : 1.f;
    const int gi = g.num_graphs == 1 ? 0 : b;
    const int lo = g.lab_off[gi], nl = g.lab_off[gi + 1] - lo;
    const float* grow = gg + ((long long)b * T + t) * nlp;
    for (int j = tid; j < nl; j += NT) delta[g.lab_list[lo + j]] = grow[j];
    __syncthreads();
    const float* x = logits + row * V;
    for (int c = warp; c < nblk; c += NWARP) {
      const float* xb = x + c * S;
      float m = neg_inf(), G = 0.f;
      for (int l = lane; l < S; l += 32) {
        m = fmaxf(m, xb[l]);
        G += delta[c * S + l];
      }
      m = warp_max(m);
      G = warp_sum(G);
      float sum = 0.f;
      for (int l = lane; l < S; l += 32) sum += ex2((xb[l] - m) * kLog2e);
      sum = warp_sum(sum);
      const float w = G / sum;
      for (int l = lane; l < S; l += 32)
        st_stream(out + c * S + l, (ex2((xb[l] - m) * kLog2e) * w - delta[c * S + l]) * scale);
    }
    __syncthreads();
    for (int j = tid; j < nl; j += NT) delta[g.lab_list[lo + j]] = 0.f;
    __syncthreads();
  }
}

// normalize_by_dim = 1 in front of a loss that has a denominator (get_normalized_acts, ctc_losses.py:36-40,
// feeding advanced_decoder.py:488-500): the dense posteriors pass through the block softmax too, so the
// normalised scores are materialised once (FWD: lp = x - LSE_block(x), rows t < len only) and the gradient with
// respect to lp, g, is pulled back in place afterwards (BWD: dL/dx[v] = g[v] - exp(lp[v]) * sum_block g).
// One warp per context block; both passes stream 8 V bytes per frame.
template <bool BWD>
__global__ void __launch_bounds__(NT) ctx_norm_kernel(const float* __restrict__ x, float* __restrict__ lp,
                                                      float* __restrict__ grad, int T, int B, int V, int S,
                                                      const int32_t* __restrict__ lens) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int nblk = V / S;
  const long long blocks = (long long)T * B * nblk;
  for (long long i = (long long)blockIdx.x * NWARP + warp; i < blocks; i += (long long)gridDim.x * NWARP) {
    const long long row = i / nblk;
    const int b = (int)(row % B), t = (int)(row / B);
    if (t >= lens[b]) continue;  // never read by the loss; its gradient rows are zeros already
    const long long o = i * S;
    if (!BWD) {
      float m = neg_inf();
      for (int l = lane; l < S; l += 32) m = fmaxf(m, x[o + l]);
      m = warp_max(m);
      float sum = 0.f;
      for (int l = lane; l < S; l += 32) sum += ex2((x[o + l] - m) * kLog2e);
      sum = warp_sum(sum);
      const float lse = m + lg2(sum) * (float)kLn2;
      for (int l = lane; l < S; l += 32) lp[o + l] = x[o + l] - lse;
    } else {
      float G = 0.f;
      for (int l = lane; l < S; l += 32) G += grad[o + l];
      G = warp_sum(G);
      for (int l = lane; l < S; l += 32) grad[o + l] = fmaf(-ex2(lp[o + l] * kLog2e), G, grad[o + l]);
    }
  }
}

__global__ void __launch_bounds__(NT) scatter_sub_kernel(int T, int B, int V, const int32_t* __restrict__ lens,
                                                         GraphDev g, const float* __restrict__ gg, int nlp,
                                                         const float* __restrict__ g_scale,
                                                         float* __restrict__ grad) {
  const int lane = threadIdx.x & 31;
  const long long row = (long long)blockIdx.x * NWARP + (threadIdx.x >> 5);
  if (row >= (long long)T * B) return;
  const int t = (int)(row / B), b = (int)(row % B);
  if (t >= lens[b]) return;
  const int gi = g.num_graphs == 1 ? 0 : b;
  const int lo = g.lab_off[gi], nl = g.lab_off[gi + 1] - lo;
  const float scale = g_scale ? g_scale[b] : 1.f;
  const float* src = gg + ((long long)b * T + t) * nlp;
  float* out = grad + row * V;
  for (int j = lane; j < nl; j += 32) out[g.lab_list[lo + j]] -= src[j] * scale;
}

// One warp per utterance; fixed summation order, fp64 accumulation.
__global__ void __launch_bounds__(NT) final_kernel(int T, int B, const int32_t* __restrict__ lens,
                                                   const double* __restrict__ cost_num,
                                                   const double* __restrict__ cost_den,
                                                   const float* __restrict__ rowlse, float* __restrict__ loss) {
  const int lane = threadIdx.x & 31;
  const int b = blockIdx.x * NWARP + (threadIdx.x >> 5);
  if (b >= B) return;
  double acc = 0.0;
  if (rowlse) {
    const int len = min(lens[b], T);
    for (int t = lane; t < len; t += 32) acc += (double)rowlse[(long long)t * B + b];
    acc = warp_sum(acc);
  }
  if (lane == 0) {
    double l = acc;
    if (cost_num) l -= cost_num[b];
    if (cost_den) l += cost_den[b];
    loss[b] = (float)l;
  }
}

// grad[t][b][:] *= scale[b * stride]; a row whose scale is exactly 1 is left alone, so the usual
// loss.sum().backward() (grad_output = ones) costs a launch and no memory traffic.
__global__ void __launch_bounds__(NT) scale_grad_kernel(float* __restrict__ grad, long long rows, int B, int V,
                                                        const float* __restrict__ scale, int stride) {
  if (stride == 0 && scale[0] == 1.f) return;  // broadcast ones (.sum().backward()): one load per CTA
  for (long long row = blockIdx.x; row < rows; row += gridDim.x) {
    const float sc = scale[(long long)(row % B) * stride];
    if (sc == 1.f) continue;
    float* g = grad + row * V;
    for (int j = threadIdx.x; j < V; j += NT) g[j] *= sc;
  }
}

__global__ void cost_to_float_kernel(int B, const double* __restrict__ cost, float* __restrict__ out) {
  const int b = blockIdx.x * blockDim.x + threadIdx.x;
  if (b < B) out[b] = (float)cost[b];
}

}  // namespace

cudaError_t launch_gather(const float* logits, int T, int B, int V, const int32_t* lens, const GraphDev& num,
                          float* xg, int nlp, int normalize, int S, bool pairs, cudaStream_t st) {
  const long long rows = (long long)T * B;
  const unsigned grid = (unsigned)((rows + NWARP - 1) / NWARP);
  ProfScope prof("gather", st);
  if (pairs) gather_kernel<true><<<grid, NT, 0, st>>>(logits, T, B, V, lens, num, xg, nlp, normalize, S);
  else gather_kernel<false><<<grid, NT, 0, st>>>(logits, T, B, V, lens, num, xg, nlp, normalize, S);
  ++g_launches;
  return cudaGetLastError();
}

cudaError_t launch_rows(const float* logits, int T, int B, int V, const int32_t* lens, const GraphDev& num,
                        const float* gg, int nlp, int mode, const float* g_scale, float* grad, float* rowlse,
                        cudaStream_t st) {
  const size_t smem = (size_t)(pad4(V) + NWARP) * 4;
  const long long rows = (long long)T * B;
  const int per = (V + NT - 1) / NT;
  const unsigned grid = (unsigned)std::min<long long>(rows, 148LL * 8);
  ProfScope prof("rows", st);
#define CTCG_ROWS(P)                                                                                          \
  do {                                                                                                        \
    cudaError_t e = cudaFuncSetAttribute(rows_kernel<P>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem); \
    if (e != cudaSuccess) return e;                                                                           \
    rows_kernel<P><<<grid, NT, smem, st>>>(logits, T, B, V, lens, num, gg, nlp, mode, g_scale, grad, rowlse); \
  } while (0)
  if (per <= 1) CTCG_ROWS(1);
  else if (per <= 2) CTCG_ROWS(2);
  else if (per <= 4) CTCG_ROWS(4);
  else if (per <= 6) CTCG_ROWS(6);
  else if (per <= 10) CTCG_ROWS(10);
  else if (per <= 16) CTCG_ROWS(16);
  else if (per <= 32) CTCG_ROWS(32);
  else return cudaErrorInvalidConfiguration;
#undef CTCG_ROWS
  ++g_launches;
  return cudaGetLastError();
}

cudaError_t launch_rows_ctx(const float* logits, int T, int B, int V, int S, const int32_t* lens, const GraphDev& num,
                            const float* gg, int nlp, const float* g_scale, float* grad, cudaStream_t st) {
  const size_t smem = (size_t)pad4(V) * 4;
  const long long rows = (long long)T * B;
  const unsigned grid = (unsigned)std::min<long long>(rows, 148LL * 8);
  ProfScope prof("rows_ctx", st);
  cudaError_t e = cudaFuncSetAttribute(rows_ctx_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  if (e != cudaSuccess) return e;
  rows_ctx_kernel<<<grid, NT, smem, st>>>(logits, T, B, V, S, lens, num, gg, nlp, g_scale, grad);
  ++g_launches;
  return cudaGetLastError();
}

cudaError_t launch_ctx_norm(bool backward, const float* x, float* lp, float* grad, int T, int B, int V, int S,
                            const int32_t* lens, cudaStream_t st) {
  const long long blocks = (long long)T * B * (V / S);
  const unsigned grid = (unsigned)std::min<long long>((blocks + NWARP - 1) / NWARP, 148LL * 16);
  ProfScope prof(backward ? "ctx_norm_bwd" : "ctx_norm_fwd", st);
  if (backward) ctx_norm_kernel<true><<<grid, NT, 0, st>>>(x, lp, grad, T, B, V, S, lens);
  else ctx_norm_kernel<false><<<grid, NT, 0, st>>>(x, lp, grad, T, B, V, S, lens);
  ++g_launches;
  return cudaGetLastError();
}

cudaError_t launch_scatter_sub(int T, int B, int V, const int32_t* lens, const GraphDev& num, const float* gg,
                               int nlp, const float* g_scale, float* grad, cudaStream_t st) {
  const long long rows = (long long)T * B;
  const unsigned grid = (unsigned)((rows + NWARP - 1) / NWARP);
  ProfScope prof("scatter_sub", st);
  scatter_sub_kernel<<<grid, NT, 0, st>>>(T, B, V, lens, num, gg, nlp, g_scale, grad);
  ++g_launches;
  return cudaGetLastError();
}

cudaError_t launch_final(int T, int B, const int32_t* lens, const double* cost_num, const double* cost_den,
                         const float* rowlse, float* loss, cudaStream_t st) {
  const unsigned grid = (unsigned)((B + NWARP - 1) / NWARP);
  ProfScope prof("final", st);
  final_kernel<<<grid, NT, 0, st>>>(T, B, lens, cost_num, cost_den, rowlse, loss);
  ++g_launches;
  return cudaGetLastError();
}

cudaError_t launch_scale_grad(float* grad, int T, int B, int V, const float* scale, int stride, cudaStream_t st) {
  const long long rows = (long long)T * B;
  const unsigned grid = (unsigned)std::min<long long>(rows, 148LL * 8);
  ProfScope prof("scale_grad", st);
  scale_grad_kernel<<<grid, NT, 0, st>>>(grad, rows, B, V, scale, stride);
  ++g_launches;
  return cudaGetLastError();
}

cudaError_t launch_cost_to_float(int B, const double* cost, float* out, cudaStream_t st) {
  cost_to_float_kernel<<<(B + 127) / 128, 128, 0, st>>>(B, cost, out);
  ++g_launches;
  return cudaGetLastError();
}

}  // namespace ctcg
