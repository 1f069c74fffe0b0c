// Internal structures shared by the graph compiler, the C ABI and the kernels.
#pragma once

#include <cstdint>
#include <functional>
#include <vector>

#include "../../include/ctcg_b200.h"

namespace ctcg {

constexpr float kLog2e = 1.4426950408889634f;
constexpr double kLn2 = 0.6931471805599453;
constexpr float kNegInf = -__builtin_huge_valf();
constexpr float kDeadWeight = -1e19f;  // anything below is "no connection"

// Device view of a batch of graphs in CSR form (all pointers are device
// pointers into one allocation).  Graph g owns states
// [state_off[g], state_off[g+1]) of the per-state arrays and rows
// [state_off[g]+g, state_off[g+1]+g] of the row-pointer arrays; row pointers
// hold absolute arc indices.
struct GraphDev {
  int32_t num_graphs;
  int32_t max_states;
  int32_t max_labels;
  int32_t max_in_degree;
  int32_t max_out_degree;
  const int32_t* state_off;  // [G+1]
  const int32_t* lab_off;    // [G+1]
  const int32_t* lab_list;   // distinct labels of graph g, ascending
  const float* final_w;      // [total_states], natural log, -inf = not final
  const int32_t* in_ptr;     // CSR by destination
  const int32_t* in_src;
  const int32_t* in_lab;     // class id
  const int32_t* in_clab;    // index into the graph's lab_list
  const float* in_w;         // natural log
  const int32_t* out_ptr;    // CSR by source
  const int32_t* out_dst;
  const int32_t* out_lab;
  const int32_t* out_clab;
  const float* out_w;
  // arcs grouped by compact label (deterministic posterior sums): graph g owns rows
  // [lab_off[g]+g, lab_off[g+1]+g] of lab_ptr; entries are absolute arc indices
  const int32_t* lab_ptr;
  const int32_t* lab_src;
  const int32_t* lab_dst;
  const float* lab_w;
  // states grouped by the compact label on their in-arcs (graphs with in_label_uniform): graph g owns rows
  // [lab_off[g]+g, lab_off[g+1]+g] of slab_ptr; entries index slab_state, whose values are state ids inside
  // the graph (ascending within a label).  States without in-arcs are not listed (never occupied after frame 0).
  const int32_t* slab_ptr;
  const int32_t* slab_state;
  // Transposed arc tables of a wide shared graph (one graph, >= 16 arcs per state on average; null otherwise):
  // slot k of state s at [k * ell_np + s], so that one-state-per-thread loops read them coalesced.  In-arcs in CSR
  // order (slot = position among the state's in-arcs), padding: other state 0, label 0, weight -inf.
  int32_t ell_np, ell_kin, ell_kout;
  const int32_t* ell_in_src;
  const int32_t* ell_in_lab;
  const float* ell_in_w;   // natural log
  const int32_t* ell_out_dst;
  const int32_t* ell_out_lab;
  const float* ell_out_w;
  int32_t max_arcs;      // most arcs of any one graph
  int32_t unit_weights;  // every arc weight is 0 (semiring One)
  int32_t in_label_uniform;  // all arcs into a state carry the same label (true for every H o C o transcript graph)
};

// Factorised description of the bi-char decoding graph (fst_utils.py:765-817):
// alpha_new[(c2,l2)] = p[cls(c2,l2)] * (Q[c2] + loop(c2,l2) * alpha[(c2,l2)]),
// Q[c2] = [c2 != 0] * sum_c1 alpha[(c1,c2)] + alpha[(c2,0)].
struct Structured {
  int32_t kind;  // 0 none, 1 mono decoding, 2 bi-char decoding
  int32_t S;
  int32_t allow_nonblank_selfloops;
  int32_t use_contextual_blanks;
  int32_t loop_using_symbol_repetitions;
};

}  // namespace ctcg

struct ctcg_graph {
  int32_t num_graphs = 0;
  int32_t max_states = 0, max_in_degree = 0, max_out_degree = 0, max_labels = 0;
  int32_t max_label = -1;  // largest class id on any arc
  std::vector<int32_t> state_off, lab_off, lab_list;
  std::vector<float> final_w;
  std::vector<int32_t> in_ptr, in_src, in_lab, in_clab;
  std::vector<float> in_w;
  std::vector<int32_t> out_ptr, out_dst, out_lab, out_clab;
  std::vector<float> out_w;
  std::vector<int32_t> lab_ptr, lab_src, lab_dst;
  std::vector<int32_t> slab_ptr, slab_state;
  std::vector<float> lab_w;
  int32_t max_arcs = 0;
  int32_t unit_weights = 1;
  int32_t in_label_uniform = 1;
  ctcg::Structured structured{0, 0, 0, 0, 0};
  // device mirror
  int device = -1;
  void* dev_blob = nullptr;
  size_t dev_bytes = 0;
  void* dev_stream = nullptr;  // stream the blob was allocated on (stream-ordered free)
  bool dev_async = false;
  ctcg::GraphDev dev{};
  unsigned long long upload_uid = 0;  // changes with every device copy (identity of the captured-step cache)
};

namespace ctcg {

// One arc of the intermediate (pre-CSR) representation.
struct Arc {
  int32_t src, dst, label;
  float weight;   // log-prob
  int32_t olabel; // emitted symbol (0 = none); only used while composing
};

struct RawGraph {
  int32_t n = 0;
  std::vector<Arc> arcs;
  std::vector<float> final_w;  // -inf = not final
};

int build_decoding(const ctcg_topology& topo, RawGraph* out);
// first[s * (nol + 1) + v] = position in dec_out[s] of the first arc with output label >= v
// (dec_out[s] is sorted by output label), v = 0..nol; empty when the table would be too large
struct OlabelIndex {
  int32_t nol = 0;
  std::vector<int32_t> first;
};
void index_olabels(const RawGraph& dec, const std::vector<std::vector<int32_t>>& dec_out, OlabelIndex* out);
int compose_with_chain(const RawGraph& dec, const std::vector<std::vector<int32_t>>& dec_out,
                       const int32_t* labels, int64_t n_labels, RawGraph* out,
                       const OlabelIndex* oi = nullptr);
int finalize_graphs(const std::vector<RawGraph>& graphs, ctcg_graph** out);
// fn(0..n-1) on a few host threads (CTCG_HOST_THREADS, default min(8, hardware)); fn must not throw
void parallel_for(int32_t n, const std::function<void(int32_t)>& fn);

}  // namespace ctcg
