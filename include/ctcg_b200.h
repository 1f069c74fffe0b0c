/*
 * ctcg_b200.h -- C ABI of the B200-native CTC / CTC-G training-loss engine.
 *
 * This is the drop-in boundary for the one hot path of chorowski-lab/pytorch-asr
 * that this repository replaces: the log-semiring forward-backward over
 * per-utterance graphs plus the gradient with respect to the logits.  Every
 * entry point cites the reference interface it stands in for (paths relative
 * to the reference tree).  Plain pointers and sizes only; no torch types.
 *
 * Conventions
 *   - every function returns a ctcg_status (0 = OK) and never throws;
 *   - "device" pointers are CUDA global-memory pointers on the current device,
 *     "host" pointers are ordinary process memory;
 *   - `stream` is a cudaStream_t passed as void* (NULL = default stream);
 *     all kernels are stream-ordered, nothing synchronises the device except
 *     the *_host entry points;
 *   - logits / log-probs are fp32, time-major contiguous (T, B, V), exactly the
 *     layout FSTDecoder hands to path_reduction (advanced_decoder.py:488);
 *   - the caller owns all tensors, the engine owns graph handles.
 *
 * There is no CPU fallback behind this ABI: without a CUDA device the compute
 * entry points return CTCG_ECUDA.
 */
#ifndef CTCG_B200_H_
#define CTCG_B200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef enum ctcg_status {
  CTCG_OK = 0,
  CTCG_EINVAL = 1,       /* bad argument (shape mismatch, NULL, label out of range) */
  CTCG_ESTART = 2,       /* start state is not 0        (fst_utils.py:246-248 ValueError) */
  CTCG_ENONDET = 3,      /* graph not input-deterministic (fst_utils.py:250-252 ValueError) */
  CTCG_EEPS = 4,         /* epsilon input label          (fst_utils.py:272-274 ValueError) */
  CTCG_ECUDA = 5,        /* CUDA runtime error / no device */
  CTCG_ENOMEM = 6,       /* host allocation failed */
  CTCG_EWORKSPACE = 7,   /* workspace too small: call ctcg_workspace_bytes */
  CTCG_EUNSUPPORTED = 8, /* configuration outside what the kernels cover */
  CTCG_ENOTUPLOADED = 9, /* graph used on device before ctcg_graph_upload */
  CTCG_ESCRATCH = 10     /* problem larger than the built-in scratch of ctcg_viterbi_static */
} ctcg_status;

/* Opaque handle: one graph shared by a whole batch (a decoding / denominator
 * graph) or one graph per utterance (training / numerator graphs).  Holds CSR
 * by destination (for alpha) and by source (for beta): row pointers,
 * {other state, label, weight} per arc, final weights, and the per-graph table
 * of distinct labels used to gather only the logits an utterance touches. */
typedef struct ctcg_graph ctcg_graph;

/* Topology of the H o C decoding graph; mirrors CTCGraphGen(context_order,
 * graph_build_args, num_symbols) -- fst_utils.py:1053-1068 and the keyword
 * arguments of build_ctc_bigram_decoding_fst, fst_utils.py:729-734. */
typedef struct ctcg_topology {
  int32_t context_order;                 /* 1 = mono-char, 2 = bi-char */
  int32_t num_symbols;                   /* S (blank is symbol 0) */
  int32_t allow_nonblank_selfloops;      /* default 1 */
  int32_t use_contextual_blanks;         /* default 0 (CTC-GB sets 1) */
  int32_t loop_using_symbol_repetitions; /* default 0 */
} ctcg_topology;

typedef struct ctcg_graph_info {
  int32_t num_graphs;     /* 1 = shared by the batch */
  int32_t max_states;
  int32_t max_in_degree;
  int32_t max_out_degree;
  int32_t max_labels;     /* max distinct labels of any one graph */
  int32_t structured;     /* 1: bi-char decoding graph with factorised fast path */
  int32_t on_device;      /* device ordinal the graph was uploaded to, or -1 */
  int32_t reserved;
  int64_t total_states;
  int64_t total_arcs;
} ctcg_graph_info;

/* Normalisation applied to the logits before the path sums; mirrors
 * FSTDecoder.normalize_by_dim (advanced_decoder.py:444-452) through
 * get_normalized_acts (ctc_losses.py:29-43). */
#define CTCG_NORM_NONE 0    /* normalize_by_dim: None -> raw logits (CTC-G / CTC-GB) */
#define CTCG_NORM_SOFTMAX 1 /* normalize_by_dim: 0    -> log_softmax over all V       */
#define CTCG_NORM_CONTEXT 2 /* normalize_by_dim: 1    -> log_softmax inside each block of S */

/* ---- graph compilation (host only; callable from DataLoader workers) ------ */

/* Decoding (denominator) graph of a topology.
 * Replaces build_ctc_mono_decoding_fst / build_ctc_bigram_decoding_fst +
 * fst_to_matrices (fst_utils.py:679-835, :222-294) and
 * BaseGraphGen.get_decoding_matrices (fst_utils.py:662-676). */
int ctcg_decoding_graph(const ctcg_topology* topo, ctcg_graph** out);

/* Training (numerator) graphs: compose(decoding graph, chain(labels_b)) for B
 * transcripts; labels of utterance b are labels[offsets[b] .. offsets[b+1]).
 * Labels larger than S are reduced modulo S first (fst_utils.py:595-596).
 * Replaces BaseGraphGen.get_training_matrices(_batch) (fst_utils.py:603-613,
 * :647-660) + build_chain_fst (:195-219) + batch_training_graph_matrices
 * (:491-521). */
int ctcg_training_graphs(const ctcg_topology* topo, const int32_t* labels,
                         const int64_t* offsets, int32_t B, ctcg_graph** out);

/* ctcg_training_graphs keeps the composed graph of every (topology, transcript) it
 * has built (bounded, first in first out; CTCG_GRAPH_CACHE=<entries>, default
 * 16384, 0 = off): the reference recomputes a transcript's matrices every epoch
 * (data/kaldi_dataset.py:63-66, 230-232).  Counters since the last clear. */
void ctcg_graph_cache_info(uint64_t* hits, uint64_t* misses, uint64_t* entries);
void ctcg_graph_cache_clear(void);

/* Decoding graphs the caller builds arc by arc WITH output labels -- the
 * tri-char topologies (build_ctc_trigram_decoding_fst_v2, fst_utils.py:926-1050)
 * and anything else that is not one of the two built-in topologies.  Arcs in the
 * caller's insertion order (src, dst, class label, emitted symbol or 0, weight);
 * connect != 0 applies OpenFst's Connect first (what Fst.rmepsilon() does to a
 * graph without eps:eps arcs, :1025), then the arcs are stably sorted by output
 * label (:1030).  State 0 is the start state. */
typedef struct ctcg_decoder ctcg_decoder;
int ctcg_decoder_from_arcs(int32_t n_states, int64_t n_arcs, const int32_t* src,
                           const int32_t* dst, const int32_t* label,
                           const int32_t* olabel, const float* weight,
                           const float* final_w, int32_t connect, ctcg_decoder** out);
int32_t ctcg_decoder_num_states(const ctcg_decoder* dec);
/* Its decoding (denominator) graph; CTCG_ENONDET when it is not input-deterministic
 * (the reference raises there too, fst_utils.py:250-252 -- the default tri-char
 * graph with a shared blank class is such a graph: numerator only). */
int ctcg_decoder_graph(const ctcg_decoder* dec, ctcg_graph** out);
/* compose(decoder, chain(labels_b)) for B transcripts, like ctcg_training_graphs
 * (BaseGraphGen.get_training_fst, fst_utils.py:592-605); labels above `modulo`
 * are reduced modulo it when modulo > 0 (:595-596). */
int ctcg_decoder_training_graphs(const ctcg_decoder* dec, const int32_t* labels,
                                 const int64_t* offsets, int32_t B, int32_t modulo,
                                 ctcg_graph** out);
void ctcg_decoder_free(ctcg_decoder* dec);

/* Import the reference's own padded matrices (the 8-tuple FSTDecoder receives
 * in batch['graph_matrices'], advanced_decoder.py:457-459): G graphs of N
 * states, in-degree K_in, out-degree K_out; arcs with weight <= nc_weight/2 are
 * padding and dropped.  Validates start==0 reachability conventions only.  * A single graph that is arc for arc the bi-char decoding graph of
 * fst_utils.py:765-817 (what get_decoding_matrices exports) is recognised and
 * takes the factorised kernels, as a graph from ctcg_decoding_graph does. */
int ctcg_graph_from_padded(const int64_t* pred, const int64_t* lab_in,
                           const float* w_in, const float* final_w,
                           const int64_t* succ, const int64_t* lab_out,
                           const float* w_out, int32_t G, int32_t N,
                           int32_t K_in, int32_t K_out, float nc_weight,
                           ctcg_graph** out);

/* Generic arc-list import (any topology: grammar-composed, tri-char ...).
 * Arcs (src, dst, label, weight=log-prob) of G graphs concatenated;
 * arc_offsets[G+1], state_counts[G]; final_w concatenated per state
 * (-inf or <= -1e19 = not final).  Rejects label < 0 (CTCG_EEPS) and
 * non-deterministic states (CTCG_ENONDET), as fst_to_matrices does. */
int ctcg_graph_from_arcs(const int32_t* src, const int32_t* dst,
                         const int32_t* label, const float* weight,
                         const int64_t* arc_offsets, const int32_t* state_counts,
                         const float* final_w, int32_t G, ctcg_graph** out);

int ctcg_graph_get_info(const ctcg_graph* g, ctcg_graph_info* info);

/* Dimensions of graph `index` in the reference's padded encoding. */
int ctcg_graph_padded_dims(const ctcg_graph* g, int32_t index, int32_t* n_states,
                           int32_t* k_in, int32_t* k_out);

/* Export graph `index` in the exact encoding of fst_to_matrices
 * (fst_utils.py:280-288): per-state arc lists sorted by (state, label,
 * weight), padded with (0, 0, nc_weight); out_edges selects succ/pred form.
 * states/labels are (N, K) int64, weights (N, K) fp32, terminal (N) fp32. */
int ctcg_graph_export_padded(const ctcg_graph* g, int32_t index, int32_t out_edges,
                             float nc_weight, int64_t* states, int64_t* labels,
                             float* weights, float* terminal);

/* Copy the CSR arrays to the current CUDA device (idempotent per device).  * A handle holds ONE device copy: uploading it while another device is current
 * releases the first copy (stream-ordered on the stream it was uploaded on), so a
 * graph shared by several GPUs needs one handle per GPU -- what one process per GPU
 * gives.  The entry points that take lens_host copy them through a pinned staging
 * slot and may wait for an earlier call's copy: they are stream-ordered, but a
 * CALLER's stream capture may not contain them (ctcg_viterbi_static is the
 * allocation-free variant; ctcg_loss_fwd_bwd keeps CUDA graphs of its own for
 * repeated calls, see ctcg_step_graph_info). */
int ctcg_graph_upload(ctcg_graph* g, void* stream);

void ctcg_graph_free(ctcg_graph* g);

/* ---- loss + gradient (device) -------------------------------------------- */

/* Bytes of device workspace ctcg_loss_fwd_bwd / ctcg_path_sum need. */
size_t ctcg_workspace_bytes(int32_t T, int32_t B, int32_t V, const ctcg_graph* num,
                            const ctcg_graph* den);

/* The same, for a given `normalize`: CTCG_NORM_CONTEXT together with a
 * denominator keeps one normalised copy of the scores (T*B*V floats more);
 * every other combination equals ctcg_workspace_bytes. */
size_t ctcg_workspace_bytes_norm(int32_t T, int32_t B, int32_t V, const ctcg_graph* num,
                                 const ctcg_graph* den, int32_t normalize);

/* Fused training loss, replaces FSTDecoder.logits' normalisation +
 * get_fst_loss + the backward of PathLogSumExp (advanced_decoder.py:444-517,
 * fst_utils.py:400-488):
 *   loss[b]  = -cost_num_b(lp) + cost_den_b(lp)          (den != NULL)
 *            = -cost_num_b(lp)                           (den == NULL)
 *   grad     = d(sum_b grad_scale[b] * loss[b]) / d logits, written in full
 *              (rows t >= lens[b] are zero), grad_scale NULL = all ones.
 * lp = logits normalised per `normalize` (CTCG_NORM_CONTEXT together with a
 * denominator needs the workspace of ctcg_workspace_bytes_norm: the normalised
 * scores are materialised there).  lens_host: B int32 on the host
 * (the reference keeps act_lens on the CPU); need not be sorted.
 * grad may be NULL (loss only).  loss/grad_scale are device pointers. */
int ctcg_loss_fwd_bwd(const float* logits, int32_t T, int32_t B, int32_t V,
                      const int32_t* lens_host, const ctcg_graph* num,
                      const ctcg_graph* den, int32_t normalize,
                      int32_t num_symbols, const float* grad_scale, float* loss,
                      float* grad, void* workspace, size_t workspace_bytes,
                      void* stream);

/* path_reduction(log_probs, act_lens, graph_matrices, 'logsumexp')
 * (fst_utils.py:322-347): cost[b] = log-sum over accepting paths of length
 * lens[b]; grad (nullable) = d cost / d log_probs, dense (T,B,V). */
int ctcg_path_sum(const float* log_probs, int32_t T, int32_t B, int32_t V,
                  const int32_t* lens_host, const ctcg_graph* graph, float* cost,
                  float* grad, void* workspace, size_t workspace_bytes,
                  void* stream);

/* path_reduction(..., red_kind='viterbi') + the arg-min-of-gradient read-out of
 * FSTDecoder.decode (fst_utils.py:366-397, advanced_decoder.py:546-554):
 * score[b] = best accepting path weight, frame_labels (T,B) int32 = label
 * consumed at each frame on that path (0 for t >= lens[b]).  The bi-char
 * decoding graph of ctcg_decoding_graph (or one recognised by
 * ctcg_graph_from_padded) takes a factorised kernel, O(V) per frame instead of
 * the graph's 51 V arcs; ties go to the smallest source state id, the order
 * torch.max sees in the reference's sorted arc lists (fst_utils.py:285). */
int ctcg_viterbi(const float* log_probs, int32_t T, int32_t B, int32_t V,
                 const int32_t* lens_host, const ctcg_graph* graph, float* score,
                 int32_t* frame_labels, void* workspace, size_t workspace_bytes,
                 void* stream);

/* PathLogSumExp.backward (fst_utils.py:482-485): grad[t][b][:] *= scale[b *
 * scale_stride], scale a DEVICE pointer (scale_stride 0: one scalar for the
 * batch, the broadcast grad_output of loss.sum().backward()).  Rows whose scale
 * is exactly 1 are skipped on the device: the common case costs one launch and
 * no memory traffic, where the reference multiplies the whole (T,B,V) tensor. */
int ctcg_scale_grad(float* grad, int32_t T, int32_t B, int32_t V, const float* scale,
                    int32_t scale_stride, void* stream);

/* Drops the topology tag of a graph built by ctcg_decoding_graph (or recognised by
 * ctcg_graph_from_padded), so that it takes the generic CSR kernels like any
 * other graph: the cross-check of the factorised bi-char kernels
 * (fst_utils.py:400-488 on the arcs of fst_utils.py:765-817). */
void ctcg_graph_set_generic(ctcg_graph* graph);

/* ctcg_viterbi without any allocation on the call path.  NO COUNTERPART IN THE
 * REFERENCE: added for callers that cannot
 * call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses
 * that forbid it).  Same outputs as ctcg_viterbi; differences:
 *  - no workspace argument: the kernel's scratch (lengths + one back-pointer per
 *    frame and state) is a file-scope device array of CTCG_STATIC_SCRATCH_BYTES
 *    that every GPU which loads the library holds until exit;
 *  - lens_dev is a DEVICE pointer to B int32 (clamped to [0, T] on the device),
 *    so the call makes no host->device copy either;
 *  - the graph must already be uploaded to the current device
 *    (ctcg_graph_upload allocates: do it at set-up), else CTCG_ENOTUPLOADED;
 *  - 256*ceil(4B/256) + 4*B*T*pad4(max_states) bytes must fit the array, else
 *    CTCG_ESCRATCH (nothing is launched).
 * Calls no cudaMalloc / cudaFree / allocating library routine; launches two
 * kernels on `stream`.  All calls share the one array (it keeps the previous
 * call's contents; nothing in it needs to start at zero): calls that overlap on
 * the GPU -- other streams or host threads on the same device -- race. */
#define CTCG_STATIC_SCRATCH_BYTES (32u << 20)
int ctcg_viterbi_static(const float* log_probs, int32_t T, int32_t B, int32_t V,
                        const int32_t* lens_dev, const ctcg_graph* graph,
                        float* score, int32_t* frame_labels, void* stream);

/* Number of accepting paths of length lens[b], modulo 2^64 (the
 * integer counterpart of path_reduction on all-zero log-probs). */
int ctcg_path_count(int32_t B, const int32_t* lens_host, const ctcg_graph* graph,
                    uint64_t* counts, void* workspace, size_t workspace_bytes,
                    void* stream);

/* ---- end-to-end with host buffers ---------------------------------------- */

/* Same as ctcg_loss_fwd_bwd but every buffer is in host memory: copies the
 * logits to the device, runs the fused path, copies loss (B) and, if non-NULL,
 * grad (T*B*V) back, and synchronises.  Scratch device memory is cached in the
 * library between calls.  This is the call a non-torch host would bind. */
int ctcg_loss_host(const float* logits_host, int32_t T, int32_t B, int32_t V,
                   const int32_t* lens_host, ctcg_graph* num, ctcg_graph* den,
                   int32_t normalize, int32_t num_symbols, float* loss_host,
                   float* grad_host);

/* Host -> device copy of a (T, B, V) logits tensor that moves only the frames
 * t < lens_host[b] (the others are never read): for a batch sorted by
 * descending length -- the order the reference asserts, fst_utils.py:382,432 --
 * one strided copy per run of frames with the same number of live utterances,
 * sum(lens) * V * 4 bytes over the bus instead of T * B * V * 4; the padding of
 * `logits_dev` is zero-filled on the device.  An unsorted batch is copied
 * whole.  Asynchronous on `stream`; `logits_host` should be pinned.  Replaces
 * the `.to(device)` of a host-side producer (advanced_decoder.py:458 does the
 * same for the graph matrices); used by ctcg_loss_host. */
int ctcg_upload_logits(const float* logits_host, int32_t T, int32_t B, int32_t V,
                       const int32_t* lens_host, float* logits_dev, void* stream);

/* ---- logits post-processing of ctc_forward.py ------------------------------ */

#define CTCG_POST_TRANSFER_HASH 1   /* --transfer-hash-prob   (ctc_forward.py:96-103) */
#define CTCG_POST_IMITATE_BIPHONES 2 /* --imitate-biphones     (:107-115)  V -> V*V columns */
#define CTCG_POST_BLOCK_NORMALIZE 4 /* --block-normalize      (:116-120)  V = S*S */
#define CTCG_POST_BLOCK_MARGINALIZE 8 /* --block-marginalize  (:121-128)  V = S*S -> S columns */

/* Number of output columns for `flags` (V*V, V or sqrt(V)); 0 if V is not a square where one
 * is needed. */
int ctcg_logits_post_width(int32_t V, int32_t flags);

/* The array transformations ctc_forward.py applies to the model's (T, B, V) log-probabilities
 * before writing them out (ctc_forward.py:96-128), with the reference's precedence of the
 * flags (imitate_biphones, else block_normalize, else block_marginalize) and its EPSILON =
 * 1e-30; fp32 like the numpy code.  log_probs and out are device pointers, out holds
 * T*B*ctcg_logits_post_width(V, flags) floats and must not alias log_probs. */
int ctcg_logits_post(const float* log_probs, int32_t T, int32_t B, int32_t V, int32_t flags,
                     float* out, void* stream);

/* Captured steps: ctcg_loss_fwd_bwd replays a CUDA graph of its launch
 * sequence for a call whose every argument (pointers, sizes, lengths, graph
 * uploads, stream) equals an earlier call's; counters since load: calls
 * replayed, captures made, calls that took the ordinary path while the cache
 * was consulted.  CTCG_STEP_GRAPHS=0 in the environment turns the cache off.
 * No reference counterpart (the reference launches ~17 ATen kernels per frame). */
void ctcg_step_graph_info(uint64_t* replays, uint64_t* captures, uint64_t* ordinary);

const char* ctcg_strerror(int status);

/* Detail of the calling thread's most recent CTCG_ECUDA (which runtime call or
 * launch failed, source line, CUDA error name and string) or alignment refusal;
 * "" if there was none.  The pointer stays valid for the thread's lifetime; the
 * text is overwritten by the next failure.  (The reference raises Python
 * exceptions with a message; status codes alone would lose it.) */
const char* ctcg_last_error_detail(void);

/* Number of kernel launches issued by this library since load (bench.py's
 * gpu_launches claim is read from here, not guessed). */
uint64_t ctcg_launch_count(void);

/* Per-launch device timing for attribution (CUDA events around every kernel the
 * library launches while enabled).  collect() synchronises, writes up to
 * max_records durations in launch order and their ';'-separated kernel names,
 * clears the log and returns the number of records. */
void ctcg_profile_enable(int on);
int ctcg_profile_collect(char* names, size_t names_cap, float* ms, int max_records);

#ifdef __cplusplus
}
#endif
#endif /* CTCG_B200_H_ */
