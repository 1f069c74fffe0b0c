// Post-processing of the (T, B, V) log-probabilities that ctc_forward.py applies before it writes
// Kaldi archives (ctc_forward.py:96-126): '#' -> blank probability transfer, replication of a
// mono-char output to bi-char shape, per-context-block normalisation, marginalisation over contexts.
// Streaming row kernels: one warp per output row, fp32 like the reference's numpy code.
#include <algorithm>

#include "device_utils.cuh"
#include "kernels.cuh"

namespace ctcg {
namespace {

constexpr int NT = 256;
constexpr int NWARP = NT / 32;
constexpr float kEps = 1e-30f;  // EPSILON, ctc_forward.py:27

// x'[0], x'[3] after the transfer (ctc_forward.py:97-103); other columns unchanged
__device__ __forceinline__ float transferred(const float* x, int j, bool transfer) {
  if (!transfer || (j != 0 && j != 3)) return x[j];
  if (j == 3) return logf(kEps);
  return logf(expf(x[0]) + (expf(x[3]) - kEps));
}

// mode 0: copy (+transfer); 1: imitate_biphones (tile to V*V, normalise unless block_normalize);
// 2: block_normalize (V = S*S); 3: block_marginalize (V = S*S -> S columns)
__global__ void __launch_bounds__(NT) post_kernel(const float* __restrict__ in, long long rows, int V, int S, int mode,
                                                  int transfer, int tile_normalize, float* __restrict__ out) {
  const int lane = threadIdx.x & 31;
  const long long row = (long long)blockIdx.x * NWARP + (threadIdx.x >> 5);
  if (row >= rows) return;
  const float* x = in + row * V;
  if (mode == 0) {
    float* o = out + row * V;
    for (int j = lane; j < V; j += 32) o[j] = transferred(x, j, transfer != 0);
  } else if (mode == 1) {
    float z = 0.f;
    if (tile_normalize) {  // log(sum over the V tiles of sum_j exp(x'[j]) + EPS), ctc_forward.py:110-115
      for (int j = lane; j < V; j += 32) z += expf(transferred(x, j, transfer != 0));
      z = warp_sum(z);
      z = logf(z * (float)V + kEps);
    }
    float* o = out + row * (long long)V * V;
    for (int j = lane; j < V; j += 32) {
      const float v = transferred(x, j, transfer != 0) - z;
      for (int c = 0; c < V; ++c) o[(long long)c * V + j] = v;
    }
  } else if (mode == 2) {  // ctc_forward.py:116-120
    float* o = out + row * V;
    constexpr int G = 4;  // blocks in flight: independent shuffle chains hide each other's latency
    for (int c0 = 0; c0 < S; c0 += G) {
      float z[G];
#pragma unroll
      for (int g = 0; g < G; ++g) {
        z[g] = 0.f;
        if (c0 + g < S)
          for (int l = lane; l < S; l += 32) z[g] += expf(transferred(x, (c0 + g) * S + l, transfer != 0));
      }
#pragma unroll
      for (int off = 16; off > 0; off >>= 1) {
#pragma unroll
        for (int g = 0; g < G; ++g) z[g] += __shfl_xor_sync(0xffffffffu, z[g], off);
      }
#pragma unroll
      for (int g = 0; g < G; ++g) {
        if (c0 + g < S) {
          const float lz = logf(z[g] + kEps);
          for (int l = lane; l < S; l += 32)
            o[(c0 + g) * S + l] = transferred(x, (c0 + g) * S + l, transfer != 0) - lz;
        }
      }
    }
  } else {  // ctc_forward.py:121-128
    float* o = out + row * S;
    for (int l = lane; l < S; l += 32) {
      float p = 0.f;
      for (int c = 0; c < S; ++c) p += expf(transferred(x, c * S + l, transfer != 0));
      o[l] = logf(p / (float)S);
    }
  }
}

}  // namespace

cudaError_t launch_post(const float* in, long long rows, int V, int S, int mode, int transfer, int tile_normalize,
                        float* out, cudaStream_t st) {
  const unsigned grid = (unsigned)((rows + NWARP - 1) / NWARP);
  ProfScope prof("logits_post", st);
  post_kernel<<<grid, NT, 0, st>>>(in, rows, V, S, mode, transfer, tile_normalize, out);
  ++g_launches;
  return cudaGetLastError();
}

}  // namespace ctcg
