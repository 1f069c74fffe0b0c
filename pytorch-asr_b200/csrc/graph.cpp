// Host-side graph compiler: CTC topologies -> per-utterance training graphs ->
// CSR arc/state arrays with label and weight tables.  No OpenFst.
//
// What it replaces in the reference (att_speech/fst_utils.py):
//   build_ctc_mono_decoding_fst   :679-726      -> build_decoding (order 1)
//   build_ctc_bigram_decoding_fst :729-835      -> build_decoding (order 2)
//   build_chain_fst + fst.compose :195-219,:604 -> compose_with_chain
//   fst_to_matrices               :222-294      -> finalize_graphs (CSR instead
//                                                  of padded (N,K) matrices) and
//                                                  ctcg_graph_export_padded
//   batch_training_graph_matrices :491-521      -> concatenated CSR, no padding
#include <algorithm>
#include <cmath>
#include <cstring>
#include <deque>
#include <new>
#include <tuple>
#include <unordered_map>

#include "ctcg_internal.h"

namespace ctcg {

namespace {

// Stable per-state sort by output label == OpenFst ArcSort(olabel); keeps the
// breadth-first state numbering of compose identical to the oracle's.
void sort_by_olabel(RawGraph* g) {
  std::stable_sort(g->arcs.begin(), g->arcs.end(), [](const Arc& a, const Arc& b) {
    if (a.src != b.src) return a.src < b.src;
    return a.olabel < b.olabel;
  });
}

}  // namespace

int build_decoding(const ctcg_topology& t, RawGraph* out) {
  const int S = t.num_symbols;
  if (S < 1) return CTCG_EINVAL;
  out->arcs.clear();
  if (t.context_order == 1) {
    // fst_utils.py:694-712
    out->n = S;
    out->final_w.assign(S, 0.f);
    for (int s = 0; s < S; ++s) {
      out->arcs.push_back({s, s, s, 0.f, 0});
      for (int sn = 0; sn < S; ++sn) {
        if (sn == s) continue;
        out->arcs.push_back({s, sn, sn, 0.f, sn});
      }
    }
  } else if (t.context_order == 2) {
    // fst_utils.py:759-817
    if (S > 1024) return CTCG_EUNSUPPORTED;
    bool loops = t.allow_nonblank_selfloops != 0;
    bool ctx_blank = t.use_contextual_blanks != 0;
    const bool rep = t.loop_using_symbol_repetitions != 0;
    if (rep) {
      loops = false;
      ctx_blank = true;
    }
    auto in_class = [&](int c, int let) { return (let != 0 || ctx_blank) ? c * S + let : 0; };
    out->n = S * S;
    out->final_w.assign(static_cast<size_t>(S) * S, 0.f);
    out->arcs.reserve(static_cast<size_t>(S) * S * (S + 1));
    for (int s1 = 0; s1 < S * S; ++s1) {
      const int c1 = s1 / S, l1 = s1 % S;
      if (loops && l1 != 0 && c1 != l1) out->arcs.push_back({s1, s1, in_class(c1, l1), 0.f, 0});
      const int c2 = (l1 == 0) ? c1 : l1;
      for (int l2 = 0; l2 < S; ++l2) {
        const int s2 = c2 * S + l2;
        const int out_s = (l2 == 0 || s1 == s2 || (rep && l1 == l2)) ? 0 : l2;
        out->arcs.push_back({s1, s2, in_class(c2, l2), 0.f, out_s});
      }
    }
  } else {
    return CTCG_EUNSUPPORTED;  // tri-char graphs enter through ctcg_graph_from_arcs
  }
  sort_by_olabel(out);
  return CTCG_OK;
}

// Product construction over (decoding state, transcript position), breadth
// first from (0,0); then OpenFst's connect (drop states that cannot reach a
// final state).  See oracle/graphs.py:compose_with_chain for the same steps.
int compose_with_chain(const RawGraph& dec, const std::vector<std::vector<int32_t>>& dec_out,
                       const int32_t* labels, int64_t U, RawGraph* out) {
  std::unordered_map<int64_t, int32_t> ids;
  std::vector<std::pair<int32_t, int32_t>> order;
  std::vector<Arc> arcs;
  auto key = [&](int32_t s, int32_t pos) { return static_cast<int64_t>(s) * (U + 1) + pos; };
  ids.emplace(key(0, 0), 0);
  order.emplace_back(0, 0);
  for (size_t head = 0; head < order.size(); ++head) {
    const auto [s, pos] = order[head];
    const int32_t src = static_cast<int32_t>(head);
    for (int32_t ai : dec_out[s]) {
      const Arc& a = dec.arcs[ai];
      int32_t npos;
      if (a.olabel == 0) {
        npos = pos;
      } else if (pos < U && a.olabel == labels[pos]) {
        npos = pos + 1;
      } else {
        continue;
      }
      const int64_t k = key(a.dst, npos);
      auto it = ids.find(k);
      int32_t nid;
      if (it == ids.end()) {
        nid = static_cast<int32_t>(order.size());
        ids.emplace(k, nid);
        order.emplace_back(a.dst, npos);
      } else {
        nid = it->second;
      }
      arcs.push_back({src, nid, a.label, a.weight, a.olabel});
    }
  }
  const int32_t n = static_cast<int32_t>(order.size());
  std::vector<float> fin(n, kNegInf);
  for (int32_t i = 0; i < n; ++i) {
    if (order[i].second == U && dec.final_w[order[i].first] > kDeadWeight) fin[i] = dec.final_w[order[i].first];
  }
  // co-accessibility
  std::vector<std::vector<int32_t>> rev(n);
  for (const Arc& a : arcs) rev[a.dst].push_back(a.src);
  std::vector<char> co(n, 0);
  std::deque<int32_t> dq;
  for (int32_t i = 0; i < n; ++i)
    if (fin[i] > kDeadWeight) {
      co[i] = 1;
      dq.push_back(i);
    }
  while (!dq.empty()) {
    const int32_t s = dq.front();
    dq.pop_front();
    for (int32_t p : rev[s])
      if (!co[p]) {
        co[p] = 1;
        dq.push_back(p);
      }
  }
  std::vector<int32_t> remap(n, -1);
  int32_t m = 0;
  for (int32_t i = 0; i < n; ++i)
    if (co[i]) remap[i] = m++;
  out->n = m;
  out->final_w.assign(m, kNegInf);
  for (int32_t i = 0; i < n; ++i)
    if (co[i]) out->final_w[remap[i]] = fin[i];
  out->arcs.clear();
  for (const Arc& a : arcs)
    if (co[a.src] && co[a.dst]) out->arcs.push_back({remap[a.src], remap[a.dst], a.label, a.weight, a.olabel});
  if (m == 0 || remap[0] != 0) {
    // empty language (cannot happen for CTC topologies): keep a lone start state
    out->n = 1;
    out->final_w.assign(1, kNegInf);
    out->arcs.clear();
  }
  return CTCG_OK;
}

// Concatenate graphs into CSR by destination and by source.  Rows are sorted by
// (other state, label, weight), the order fst_to_matrices uses
// (fst_utils.py:285), which also fixes the Viterbi tie-break.
int finalize_graphs(const std::vector<RawGraph>& graphs, ctcg_graph** out) {
  ctcg_graph* g = new (std::nothrow) ctcg_graph();
  if (!g) return CTCG_ENOMEM;
  const int32_t G = static_cast<int32_t>(graphs.size());
  g->num_graphs = G;
  g->state_off.assign(G + 1, 0);
  g->lab_off.assign(G + 1, 0);
  int64_t total_states = 0, total_arcs = 0;
  for (int32_t i = 0; i < G; ++i) {
    total_states += graphs[i].n;
    total_arcs += static_cast<int64_t>(graphs[i].arcs.size());
    g->state_off[i + 1] = static_cast<int32_t>(total_states);
  }
  if (total_states + G > INT32_MAX || total_arcs > INT32_MAX) {
    delete g;
    return CTCG_EUNSUPPORTED;
  }
  g->final_w.resize(total_states);
  g->in_ptr.assign(total_states + G, 0);
  g->out_ptr.assign(total_states + G, 0);
  g->in_src.resize(total_arcs);
  g->in_lab.resize(total_arcs);
  g->in_clab.resize(total_arcs);
  g->in_w.resize(total_arcs);
  g->out_dst.resize(total_arcs);
  g->out_lab.resize(total_arcs);
  g->out_clab.resize(total_arcs);
  g->out_w.resize(total_arcs);
  g->lab_src.resize(total_arcs);
  g->lab_dst.resize(total_arcs);
  g->lab_w.resize(total_arcs);
  int32_t arc_base = 0;
  std::vector<int32_t> idx;
  for (int32_t gi = 0; gi < G; ++gi) {
    const RawGraph& r = graphs[gi];
    const int32_t n = r.n;
    const int32_t so = g->state_off[gi];
    const int32_t ro = so + gi;
    const int32_t na = static_cast<int32_t>(r.arcs.size());
    for (int32_t s = 0; s < n; ++s) g->final_w[so + s] = (r.final_w[s] > kDeadWeight) ? r.final_w[s] : kNegInf;
    // validation (fst_utils.py:250-252, :272-274) + label table
    std::vector<int32_t> labs;
    labs.reserve(na);
    for (const Arc& a : r.arcs) {
      if (a.label < 0) {
        delete g;
        return CTCG_EEPS;
      }
      if (a.src < 0 || a.src >= n || a.dst < 0 || a.dst >= n) {
        delete g;
        return CTCG_EINVAL;
      }
      labs.push_back(a.label);
    }
    std::sort(labs.begin(), labs.end());
    labs.erase(std::unique(labs.begin(), labs.end()), labs.end());
    g->lab_off[gi + 1] = g->lab_off[gi] + static_cast<int32_t>(labs.size());
    g->lab_list.insert(g->lab_list.end(), labs.begin(), labs.end());
    g->max_labels = std::max<int32_t>(g->max_labels, static_cast<int32_t>(labs.size()));
    if (!labs.empty()) g->max_label = std::max(g->max_label, labs.back());
    g->max_states = std::max(g->max_states, n);
    auto clab = [&](int32_t lab) {
      return static_cast<int32_t>(std::lower_bound(labs.begin(), labs.end(), lab) - labs.begin());
    };
    // by source
    idx.resize(na);
    for (int32_t i = 0; i < na; ++i) idx[i] = i;
    std::sort(idx.begin(), idx.end(), [&](int32_t x, int32_t y) {
      const Arc &a = r.arcs[x], &b = r.arcs[y];
      return std::tie(a.src, a.dst, a.label, a.weight) < std::tie(b.src, b.dst, b.label, b.weight);
    });
    {
      std::vector<int32_t> deg(n + 1, 0);
      for (const Arc& a : r.arcs) deg[a.src + 1]++;
      for (int32_t s = 0; s < n; ++s) {
        g->max_out_degree = std::max(g->max_out_degree, deg[s + 1]);
        deg[s + 1] += deg[s];
      }
      for (int32_t s = 0; s <= n; ++s) g->out_ptr[ro + s] = arc_base + deg[s];
      for (int32_t i = 0; i < na; ++i) {
        const Arc& a = r.arcs[idx[i]];
        g->out_dst[arc_base + i] = a.dst;
        g->out_lab[arc_base + i] = a.label;
        g->out_clab[arc_base + i] = clab(a.label);
        g->out_w[arc_base + i] = a.weight;
      }
      // input determinism: no two arcs of one state share a label
      std::vector<int32_t> row;
      for (int32_t s = 0; s < n; ++s) {
        row.clear();
        for (int32_t i = deg[s]; i < deg[s + 1]; ++i) row.push_back(r.arcs[idx[i]].label);
        std::sort(row.begin(), row.end());
        if (std::adjacent_find(row.begin(), row.end()) != row.end()) {
          delete g;
          return CTCG_ENONDET;
        }
      }
    }
    // by destination
    std::sort(idx.begin(), idx.end(), [&](int32_t x, int32_t y) {
      const Arc &a = r.arcs[x], &b = r.arcs[y];
      return std::tie(a.dst, a.src, a.label, a.weight) < std::tie(b.dst, b.src, b.label, b.weight);
    });
    {
      std::vector<int32_t> deg(n + 1, 0);
      for (const Arc& a : r.arcs) deg[a.dst + 1]++;
      for (int32_t s = 0; s < n; ++s) {
        g->max_in_degree = std::max(g->max_in_degree, deg[s + 1]);
        deg[s + 1] += deg[s];
      }
      for (int32_t s = 0; s <= n; ++s) g->in_ptr[ro + s] = arc_base + deg[s];
      for (int32_t i = 0; i < na; ++i) {
        const Arc& a = r.arcs[idx[i]];
        g->in_src[arc_base + i] = a.src;
        g->in_lab[arc_base + i] = a.label;
        g->in_clab[arc_base + i] = clab(a.label);
        g->in_w[arc_base + i] = a.weight;
      }
    }
    // by compact label (then source, destination): fixed summation order of the posteriors
    std::sort(idx.begin(), idx.end(), [&](int32_t x, int32_t y) {
      const Arc &a = r.arcs[x], &b = r.arcs[y];
      return std::tie(a.label, a.src, a.dst, a.weight) < std::tie(b.label, b.src, b.dst, b.weight);
    });
    {
      const int32_t nl = static_cast<int32_t>(labs.size());
      std::vector<int32_t> cnt(nl + 1, 0);
      for (const Arc& a : r.arcs) cnt[clab(a.label) + 1]++;
      for (int32_t j = 0; j < nl; ++j) cnt[j + 1] += cnt[j];
      for (int32_t j = 0; j <= nl; ++j) g->lab_ptr.push_back(arc_base + cnt[j]);
      for (int32_t i = 0; i < na; ++i) {
        const Arc& a = r.arcs[idx[i]];
        g->lab_src[arc_base + i] = a.src;
        g->lab_dst[arc_base + i] = a.dst;
        g->lab_w[arc_base + i] = a.weight;
        if (a.weight != 0.f) g->unit_weights = 0;
      }
    }
    g->max_arcs = std::max(g->max_arcs, na);
    arc_base += na;
  }
  *out = g;
  return CTCG_OK;
}

}  // namespace ctcg
