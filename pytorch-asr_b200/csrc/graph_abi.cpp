// extern "C" entry points of the graph compiler (host only, no CUDA calls, so
// DataLoader worker processes can use them without creating a CUDA context).
#include <algorithm>
#include <cmath>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <map>
#include <deque>
#include <unordered_map>
#include <string>
#include <memory>
#include <mutex>
#include <new>
#include <tuple>

#include "ctcg_internal.h"

using namespace ctcg;

namespace {

void index_out_arcs(const RawGraph& g, std::vector<std::vector<int32_t>>* out) {
  out->assign(g.n, {});
  for (int32_t i = 0; i < static_cast<int32_t>(g.arcs.size()); ++i) (*out)[g.arcs[i].src].push_back(i);
}

// Decoding graphs are a pure function of the topology: build each once per process
// (the S=49 bi-char graph has 120k arcs) and share it between calls and threads.
struct CachedDecoding {
  RawGraph g;
  std::vector<std::vector<int32_t>> out;
  OlabelIndex oi;
};
std::shared_ptr<const CachedDecoding> cached_decoding(const ctcg_topology& t, int* status) {
  static std::mutex mu;
  static std::map<std::tuple<int, int, int, int, int>, std::shared_ptr<const CachedDecoding>> cache;
  const auto key = std::make_tuple(t.context_order, t.num_symbols, t.allow_nonblank_selfloops != 0,
                                   t.use_contextual_blanks != 0, t.loop_using_symbol_repetitions != 0);
  std::lock_guard<std::mutex> lock(mu);
  auto it = cache.find(key);
  if (it != cache.end()) {
    *status = CTCG_OK;
    return it->second;
  }
  auto c = std::make_shared<CachedDecoding>();
  *status = build_decoding(t, &c->g);
  if (*status != CTCG_OK) return nullptr;
  index_out_arcs(c->g, &c->out);
  index_olabels(c->g, c->out, &c->oi);
  cache[key] = c;
  return c;
}

// Composed training graphs, per (topology, transcript): the same transcripts come back every epoch
// (kaldi_dataset.py:63-66, 230-232 recomputes them per item per epoch), so the product construction is
// done once.  Bounded, first-in first-out; CTCG_GRAPH_CACHE=<entries> (default 16384, about 12 KB each at
// U = 100; 0 turns it off).
struct TrainingCache {
  std::mutex mu;
  std::unordered_map<std::string, std::shared_ptr<const RawGraph>> map;
  std::deque<std::string> order;
  uint64_t hits = 0, misses = 0;
  size_t capacity = [] {
    const char* e = std::getenv("CTCG_GRAPH_CACHE");
    return e ? static_cast<size_t>(std::max(0L, std::atol(e))) : size_t{16384};
  }();
};
TrainingCache& training_cache() {
  static TrainingCache c;
  return c;
}
std::string cache_key(const ctcg_topology& t, const std::vector<int32_t>& lab) {
  const int32_t head[5] = {t.context_order, t.num_symbols, t.allow_nonblank_selfloops != 0,
                           t.use_contextual_blanks != 0, t.loop_using_symbol_repetitions != 0};
  std::string k(reinterpret_cast<const char*>(head), sizeof(head));
  k.append(reinterpret_cast<const char*>(lab.data()), lab.size() * sizeof(int32_t));
  return k;
}

void mark_structured(const ctcg_topology& t, ctcg_graph* g) {
  g->structured.kind = t.context_order;
  g->structured.S = t.num_symbols;
  bool loops = t.allow_nonblank_selfloops != 0;
  bool ctx = t.use_contextual_blanks != 0;
  if (t.loop_using_symbol_repetitions) {
    loops = false;
    ctx = true;
  }
  g->structured.allow_nonblank_selfloops = loops;
  g->structured.use_contextual_blanks = ctx;
  g->structured.loop_using_symbol_repetitions = t.loop_using_symbol_repetitions != 0;
}

// A single graph with S*S states, all final, unit weights, whose arcs are exactly those of the bi-char
// decoding graph of some flag combination (fst_utils.py:765-817) is that graph: tag it, so that matrices
// which went through the reference's padded format (get_decoding_matrices -> path_reduction, the route
// that leaves the reference's FSTDecoder untouched) still take the factorised kernels.  The padded format
// carries no output labels, so loop_using_symbol_repetitions cannot be told from (no self-loops,
// contextual blanks) -- and does not matter for a path sum.
void recognise_bichar(const RawGraph& r, ctcg_graph* g) {
  if (g->num_graphs != 1 || !g->unit_weights) return;
  int S = 2;
  while (S * S < r.n) ++S;
  if (S * S != r.n || S < 2 || S > 64) return;
  for (float f : r.final_w)
    if (f != 0.f) return;
  auto key = [](const Arc& a) { return std::make_tuple(a.src, a.dst, a.label); };
  std::vector<std::tuple<int32_t, int32_t, int32_t>> mine;
  mine.reserve(r.arcs.size());
  for (const Arc& a : r.arcs) mine.push_back(key(a));
  std::sort(mine.begin(), mine.end());
  for (int loops = 1; loops >= 0; --loops)
    for (int ctx = 0; ctx <= 1; ++ctx) {
      ctcg_topology t{};
      t.context_order = 2;
      t.num_symbols = S;
      t.allow_nonblank_selfloops = loops;
      t.use_contextual_blanks = ctx;
      t.loop_using_symbol_repetitions = 0;
      int st = CTCG_OK;
      const auto c = cached_decoding(t, &st);
      if (st != CTCG_OK || !c || c->g.arcs.size() != r.arcs.size()) continue;
      std::vector<std::tuple<int32_t, int32_t, int32_t>> want;
      want.reserve(c->g.arcs.size());
      for (const Arc& a : c->g.arcs) want.push_back(key(a));
      std::sort(want.begin(), want.end());
      if (want == mine) {
        mark_structured(t, g);
        return;
      }
    }
}

}  // namespace

extern "C" {

void ctcg_graph_cache_info(uint64_t* hits, uint64_t* misses, uint64_t* entries) {
  TrainingCache& tc = training_cache();
  std::lock_guard<std::mutex> lock(tc.mu);
  if (hits) *hits = tc.hits;
  if (misses) *misses = tc.misses;
  if (entries) *entries = tc.map.size();
}

void ctcg_graph_cache_clear(void) {
  TrainingCache& tc = training_cache();
  std::lock_guard<std::mutex> lock(tc.mu);
  tc.map.clear();
  tc.order.clear();
  tc.hits = tc.misses = 0;
}

void ctcg_graph_set_generic(ctcg_graph* g) {
  if (!g) return;
  g->structured = ctcg::Structured{0, 0, 0, 0, 0};
  // the handle now selects other kernels: steps captured with it (abi.cu, step graphs) must not be replayed
  g->upload_uid += 1ull << 40;
}

int ctcg_decoding_graph(const ctcg_topology* topo, ctcg_graph** out) {
  if (!topo || !out) return CTCG_EINVAL;
  try {
    std::vector<RawGraph> gs(1);
    int st = build_decoding(*topo, &gs[0]);
    if (st != CTCG_OK) return st;
    st = finalize_graphs(gs, out);
    if (st == CTCG_OK) mark_structured(*topo, *out);
    return st;
  } catch (const std::bad_alloc&) {
    return CTCG_ENOMEM;
  }
}

int ctcg_training_graphs(const ctcg_topology* topo, const int32_t* labels, const int64_t* offsets,
                         int32_t B, ctcg_graph** out) {
  if (!topo || !offsets || !out || B < 0) return CTCG_EINVAL;
  try {
    int st = CTCG_OK;
    const auto t_0 = std::chrono::steady_clock::now();
    const auto cached = cached_decoding(*topo, &st);
    if (st != CTCG_OK) return st;
    const auto t_1 = std::chrono::steady_clock::now();
    const RawGraph& dec = cached->g;
    const std::vector<std::vector<int32_t>>& dec_out = cached->out;
    std::vector<RawGraph> gs(B);
    std::vector<int> status(B, CTCG_OK);
    const int32_t S = topo->num_symbols;
    for (int32_t b = 0; b < B; ++b)
      if (offsets[b + 1] < offsets[b]) return CTCG_EINVAL;
    parallel_for(B, [&](int32_t b) {
      try {
        std::vector<int32_t> lab(labels + offsets[b], labels + offsets[b + 1]);
        // fst_utils.py:595-596: a transcript holding any id above S is taken modulo S
        bool any_big = false;
        for (int32_t v : lab) {
          if (v < 0) {
            status[b] = CTCG_EINVAL;
            return;
          }
          any_big |= v > S;
        }
        if (any_big)
          for (int32_t& v : lab) v %= S;
        TrainingCache& tc = training_cache();
        std::string key;
        if (tc.capacity > 0) {
          key = cache_key(*topo, lab);
          std::shared_ptr<const RawGraph> hit;
          {
            std::lock_guard<std::mutex> lock(tc.mu);
            auto it = tc.map.find(key);
            if (it != tc.map.end()) {
              hit = it->second;
              ++tc.hits;
            } else {
              ++tc.misses;
            }
          }
          if (hit) {
            gs[b] = *hit;
            return;
          }
        }
        status[b] = compose_with_chain(dec, dec_out, lab.data(), static_cast<int64_t>(lab.size()), &gs[b], &cached->oi);
        if (status[b] == CTCG_OK && tc.capacity > 0) {
          auto copy = std::make_shared<const RawGraph>(gs[b]);
          std::lock_guard<std::mutex> lock(tc.mu);
          if (tc.map.emplace(key, std::move(copy)).second) {
            tc.order.push_back(key);
            while (tc.order.size() > tc.capacity) {
              tc.map.erase(tc.order.front());
              tc.order.pop_front();
            }
          }
        }
      } catch (const std::bad_alloc&) {
        status[b] = CTCG_ENOMEM;
      }
    });
    for (int32_t b = 0; b < B; ++b)
      if (status[b] != CTCG_OK) return status[b];
    const auto t_2 = std::chrono::steady_clock::now();
    const int rc = finalize_graphs(gs, out);
    const auto t_3 = std::chrono::steady_clock::now();
    if (std::getenv("CTCG_HOST_TIMING"))
      std::fprintf(stderr, "cached %.3f compose %.3f finalize %.3f ms\n",
                   std::chrono::duration<double, std::milli>(t_1 - t_0).count(),
                   std::chrono::duration<double, std::milli>(t_2 - t_1).count(),
                   std::chrono::duration<double, std::milli>(t_3 - t_2).count());
    return rc;
  } catch (const std::bad_alloc&) {
    return CTCG_ENOMEM;
  }
}

// ---- decoding graphs given as arcs WITH output labels (tri-char and other topologies built by the caller) ----
}  // extern "C"

struct ctcg_decoder {
  RawGraph g;
  std::vector<std::vector<int32_t>> out;
  OlabelIndex oi;
};

extern "C" {

int ctcg_decoder_from_arcs(int32_t n_states, int64_t n_arcs, const int32_t* src, const int32_t* dst,
                           const int32_t* label, const int32_t* olabel, const float* weight,
                           const float* final_w, int32_t connect, ctcg_decoder** out) {
  if (n_states < 1 || n_arcs < 0 || !final_w || !out || (n_arcs > 0 && (!src || !dst || !label || !olabel))) return CTCG_EINVAL;
  try {
    std::unique_ptr<ctcg_decoder> d(new ctcg_decoder());
    RawGraph& r = d->g;
    r.n = n_states;
    r.final_w.resize(n_states);
    for (int32_t s = 0; s < n_states; ++s) r.final_w[s] = (final_w[s] > kDeadWeight) ? final_w[s] : kNegInf;
    r.arcs.reserve(static_cast<size_t>(n_arcs));
    for (int64_t i = 0; i < n_arcs; ++i) {
      if (src[i] < 0 || src[i] >= n_states || dst[i] < 0 || dst[i] >= n_states || olabel[i] < 0) return CTCG_EINVAL;
      if (label[i] < 0) return CTCG_EEPS;
      r.arcs.push_back({src[i], dst[i], label[i], weight ? weight[i] : 0.f, olabel[i]});
    }
    if (connect) {
      // OpenFst Connect (what Fst.rmepsilon() leaves of itself on a graph without eps:eps arcs,
      // fst_utils.py:1025): keep the states that are accessible from state 0 AND can reach a final
      // state, renumber them in increasing order, keep the arcs in their order
      const int32_t n = r.n;
      std::vector<int32_t> ptr(n + 1, 0), rptr(n + 1, 0);
      for (const Arc& a : r.arcs) {
        ptr[a.src + 1]++;
        rptr[a.dst + 1]++;
      }
      for (int32_t s = 0; s < n; ++s) {
        ptr[s + 1] += ptr[s];
        rptr[s + 1] += rptr[s];
      }
      std::vector<int32_t> fwd(r.arcs.size()), bwd(r.arcs.size()), fc(ptr.begin(), ptr.end() - 1), bc(rptr.begin(), rptr.end() - 1);
      for (const Arc& a : r.arcs) {
        fwd[fc[a.src]++] = a.dst;
        bwd[bc[a.dst]++] = a.src;
      }
      std::vector<char> acc(n, 0), co(n, 0);
      std::vector<int32_t> stack{0};
      acc[0] = 1;
      while (!stack.empty()) {
        const int32_t s = stack.back();
        stack.pop_back();
        for (int32_t i = ptr[s]; i < ptr[s + 1]; ++i)
          if (!acc[fwd[i]]) {
            acc[fwd[i]] = 1;
            stack.push_back(fwd[i]);
          }
      }
      for (int32_t s = 0; s < n; ++s)
        if (r.final_w[s] > kDeadWeight) {
          co[s] = 1;
          stack.push_back(s);
        }
      while (!stack.empty()) {
        const int32_t s = stack.back();
        stack.pop_back();
        for (int32_t i = rptr[s]; i < rptr[s + 1]; ++i)
          if (!co[bwd[i]]) {
            co[bwd[i]] = 1;
            stack.push_back(bwd[i]);
          }
      }
      std::vector<int32_t> remap(n, -1);
      int32_t m = 0;
      for (int32_t s = 0; s < n; ++s)
        if (acc[s] && co[s]) remap[s] = m++;
      if (m == 0 || remap[0] != 0) return CTCG_ESTART;
      RawGraph t;
      t.n = m;
      t.final_w.resize(m);
      for (int32_t s = 0; s < n; ++s)
        if (remap[s] >= 0) t.final_w[remap[s]] = r.final_w[s];
      for (const Arc& a : r.arcs)
        if (remap[a.src] >= 0 && remap[a.dst] >= 0) t.arcs.push_back({remap[a.src], remap[a.dst], a.label, a.weight, a.olabel});
      r = std::move(t);
    }
    // arcsort('olabel'), stable: the order compose visits the arcs in, hence the state numbering of the
    // training graphs (fst_utils.py:1030, :603-605)
    std::stable_sort(r.arcs.begin(), r.arcs.end(), [](const Arc& a, const Arc& b) {
      if (a.src != b.src) return a.src < b.src;
      return a.olabel < b.olabel;
    });
    index_out_arcs(r, &d->out);
    index_olabels(r, d->out, &d->oi);
    *out = d.release();
    return CTCG_OK;
  } catch (const std::bad_alloc&) {
    return CTCG_ENOMEM;
  }
}

int32_t ctcg_decoder_num_states(const ctcg_decoder* d) { return d ? d->g.n : 0; }

int ctcg_decoder_graph(const ctcg_decoder* d, ctcg_graph** out) {
  if (!d || !out) return CTCG_EINVAL;
  try {
    std::vector<RawGraph> gs(1, d->g);
    return finalize_graphs(gs, out);  // CTCG_ENONDET when two arcs of a state share a class (fst_utils.py:250-252)
  } catch (const std::bad_alloc&) {
    return CTCG_ENOMEM;
  }
}

int ctcg_decoder_training_graphs(const ctcg_decoder* d, const int32_t* labels, const int64_t* offsets, int32_t B,
                                 int32_t modulo, ctcg_graph** out) {
  if (!d || !offsets || !out || B < 0) return CTCG_EINVAL;
  try {
    std::vector<RawGraph> gs(B);
    std::vector<int> status(B, CTCG_OK);
    for (int32_t b = 0; b < B; ++b)
      if (offsets[b + 1] < offsets[b]) return CTCG_EINVAL;
    parallel_for(B, [&](int32_t b) {
      try {
        std::vector<int32_t> lab(labels + offsets[b], labels + offsets[b + 1]);
        bool any_big = false;  // fst_utils.py:595-596: a transcript holding any id above S is taken modulo S
        for (int32_t v : lab) {
          if (v < 0) {
            status[b] = CTCG_EINVAL;
            return;
          }
          any_big |= modulo > 0 && v > modulo;
        }
        if (any_big)
          for (int32_t& v : lab) v %= modulo;
        status[b] = compose_with_chain(d->g, d->out, lab.data(), static_cast<int64_t>(lab.size()), &gs[b], &d->oi);
      } catch (const std::bad_alloc&) {
        status[b] = CTCG_ENOMEM;
      }
    });
    for (int32_t b = 0; b < B; ++b)
      if (status[b] != CTCG_OK) return status[b];
    return finalize_graphs(gs, out);
  } catch (const std::bad_alloc&) {
    return CTCG_ENOMEM;
  }
}

void ctcg_decoder_free(ctcg_decoder* d) { delete d; }

int ctcg_graph_from_padded(const int64_t* pred, const int64_t* lab_in, const float* w_in,
                           const float* final_w, const int64_t* succ, const int64_t* lab_out,
                           const float* w_out, int32_t G, int32_t N, int32_t K_in, int32_t K_out,
                           float nc_weight, ctcg_graph** out) {
  if (!pred || !lab_in || !w_in || !final_w || !out || G < 1 || N < 1 || K_in < 0) return CTCG_EINVAL;
  (void)succ;
  (void)lab_out;
  (void)w_out;
  (void)K_out;  // the out-edge half carries the same arcs; the in-edge half is authoritative
  try {
    const float dead = nc_weight * 0.5f;
    std::vector<RawGraph> gs(G);
    for (int32_t g = 0; g < G; ++g) {
      RawGraph& r = gs[g];
      // trailing all-padding states of a batch-padded graph are kept: they are dead and cost nothing
      r.n = N;
      r.final_w.resize(N);
      for (int32_t d = 0; d < N; ++d) {
        const float f = final_w[static_cast<size_t>(g) * N + d];
        r.final_w[d] = (f > dead) ? f : kNegInf;
        for (int32_t k = 0; k < K_in; ++k) {
          const size_t i = (static_cast<size_t>(g) * N + d) * K_in + k;
          if (!(w_in[i] > dead)) continue;
          const int64_t s = pred[i], l = lab_in[i];
          if (s < 0 || s >= N) return CTCG_EINVAL;
          if (l < 0) return CTCG_EEPS;
          r.arcs.push_back({static_cast<int32_t>(s), d, static_cast<int32_t>(l), w_in[i], 0});
        }
      }
    }
    const int st = finalize_graphs(gs, out);
    if (st == CTCG_OK && G == 1) recognise_bichar(gs[0], *out);
    return st;
  } catch (const std::bad_alloc&) {
    return CTCG_ENOMEM;
  }
}

int ctcg_graph_from_arcs(const int32_t* src, const int32_t* dst, const int32_t* label,
                         const float* weight, const int64_t* arc_offsets,
                         const int32_t* state_counts, const float* final_w, int32_t G,
                         ctcg_graph** out) {
  if (!src || !dst || !label || !weight || !arc_offsets || !state_counts || !final_w || !out || G < 1)
    return CTCG_EINVAL;
  try {
    std::vector<RawGraph> gs(G);
    size_t so = 0;
    for (int32_t g = 0; g < G; ++g) {
      RawGraph& r = gs[g];
      r.n = state_counts[g];
      if (r.n < 1) return CTCG_EINVAL;
      r.final_w.assign(final_w + so, final_w + so + r.n);
      so += r.n;
      for (int64_t i = arc_offsets[g]; i < arc_offsets[g + 1]; ++i) {
        if (!(weight[i] > kDeadWeight)) continue;
        r.arcs.push_back({src[i], dst[i], label[i], weight[i], 0});
      }
    }
    return finalize_graphs(gs, out);
  } catch (const std::bad_alloc&) {
    return CTCG_ENOMEM;
  }
}

int ctcg_graph_get_info(const ctcg_graph* g, ctcg_graph_info* info) {
  if (!g || !info) return CTCG_EINVAL;
  info->num_graphs = g->num_graphs;
  info->max_states = g->max_states;
  info->max_in_degree = g->max_in_degree;
  info->max_out_degree = g->max_out_degree;
  info->max_labels = g->max_labels;
  info->structured = g->structured.kind == 2 ? 1 : 0;
  info->on_device = g->device;
  info->reserved = 0;
  info->total_states = g->state_off.back();
  info->total_arcs = static_cast<int64_t>(g->in_src.size());
  return CTCG_OK;
}

int ctcg_graph_padded_dims(const ctcg_graph* g, int32_t index, int32_t* n_states, int32_t* k_in,
                           int32_t* k_out) {
  if (!g || index < 0 || index >= g->num_graphs) return CTCG_EINVAL;
  const int32_t so = g->state_off[index], n = g->state_off[index + 1] - so, ro = so + index;
  int32_t ki = 0, ko = 0;
  for (int32_t s = 0; s < n; ++s) {
    ki = std::max(ki, g->in_ptr[ro + s + 1] - g->in_ptr[ro + s]);
    ko = std::max(ko, g->out_ptr[ro + s + 1] - g->out_ptr[ro + s]);
  }
  if (n_states) *n_states = n;
  if (k_in) *k_in = ki;
  if (k_out) *k_out = ko;
  return CTCG_OK;
}

int ctcg_graph_export_padded(const ctcg_graph* g, int32_t index, int32_t out_edges, float nc_weight,
                             int64_t* states, int64_t* labels, float* weights, float* terminal) {
  if (!g || index < 0 || index >= g->num_graphs || !states || !labels || !weights || !terminal)
    return CTCG_EINVAL;
  int32_t n, ki, ko;
  ctcg_graph_padded_dims(g, index, &n, &ki, &ko);
  const int32_t K = out_edges ? ko : ki;
  const int32_t so = g->state_off[index], ro = so + index;
  const int32_t* ptr = out_edges ? g->out_ptr.data() : g->in_ptr.data();
  const int32_t* other = out_edges ? g->out_dst.data() : g->in_src.data();
  const int32_t* lab = out_edges ? g->out_lab.data() : g->in_lab.data();
  const float* w = out_edges ? g->out_w.data() : g->in_w.data();
  for (int32_t s = 0; s < n; ++s) {
    terminal[s] = (g->final_w[so + s] > kDeadWeight) ? g->final_w[so + s] : nc_weight;
    const int32_t lo = ptr[ro + s], hi = ptr[ro + s + 1];
    for (int32_t k = 0; k < K; ++k) {
      const size_t o = static_cast<size_t>(s) * K + k;
      if (lo + k < hi) {
        states[o] = other[lo + k];
        labels[o] = lab[lo + k];
        weights[o] = w[lo + k];
      } else {
        states[o] = 0;
        labels[o] = 0;
        weights[o] = nc_weight;
      }
    }
  }
  return CTCG_OK;
}

const char* ctcg_strerror(int status) {
  switch (status) {
    case CTCG_OK: return "ok";
    case CTCG_EINVAL: return "invalid argument";
    case CTCG_ESTART: return "FST starting state is not 0";
    case CTCG_ENONDET: return "FST is not deterministic";
    case CTCG_EEPS: return "FST has eps-transitions";
    case CTCG_ECUDA: return "CUDA error (no device, or a runtime call failed)";
    case CTCG_ENOMEM: return "out of host memory";
    case CTCG_EWORKSPACE: return "workspace too small";
    case CTCG_EUNSUPPORTED: return "unsupported configuration";
    case CTCG_ENOTUPLOADED: return "graph not uploaded to the device";
    case CTCG_ESCRATCH: return "problem exceeds the built-in scratch (CTCG_STATIC_SCRATCH_BYTES) of ctcg_viterbi_static";
    default: return "unknown status";
  }
}

}  // extern "C"
