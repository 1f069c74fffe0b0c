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
#include <functional>
#include <new>
#include <tuple>
#include <atomic>
#include <cstdlib>
#include <thread>
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
void index_olabels(const RawGraph& dec, const std::vector<std::vector<int32_t>>& dec_out, OlabelIndex* out) {
  int32_t mx = 0;
  for (const Arc& a : dec.arcs) mx = std::max(mx, a.olabel);
  out->nol = mx + 1;
  out->first.clear();
  const int64_t cells = static_cast<int64_t>(dec.n) * (out->nol + 1);
  if (cells > (int64_t{16} << 20)) return;  // 64 MB: fall back to binary search
  out->first.assign(cells, 0);
  for (int32_t s = 0; s < dec.n; ++s) {
    const std::vector<int32_t>& row = dec_out[s];
    int32_t* f = out->first.data() + static_cast<int64_t>(s) * (out->nol + 1);
    int32_t q = 0;
    for (int32_t v = 0; v <= out->nol; ++v) {
      while (q < static_cast<int32_t>(row.size()) && dec.arcs[row[q]].olabel < v) ++q;
      f[v] = q;
    }
  }
}

int compose_with_chain(const RawGraph& dec, const std::vector<std::vector<int32_t>>& dec_out,
                       const int32_t* labels, int64_t U, RawGraph* out, const OlabelIndex* oi) {
  // scratch reused between calls of one thread: a training graph is a few hundred states,
  // so allocation would otherwise dominate
  struct Scratch {
    std::vector<int64_t> keys;   // open-addressing table: (state, position) -> product state
    std::vector<int32_t> vals;
    std::vector<std::pair<int32_t, int32_t>> order;
    std::vector<Arc> arcs;
    std::vector<float> fin;
    std::vector<int32_t> rptr, rsrc, stack, remap;
    std::vector<char> co;
  };
  thread_local Scratch sc;
  size_t cap = 256;
  while (cap < static_cast<size_t>(8 * (U + 1) + 16)) cap <<= 1;
  auto rehash = [&](size_t new_cap) {
    std::vector<int64_t> ok(new_cap, -1);
    std::vector<int32_t> ov(new_cap, 0);
    for (size_t i = 0; i < sc.keys.size(); ++i) {
      if (sc.keys[i] < 0) continue;
      size_t h = (static_cast<uint64_t>(sc.keys[i]) * 0x9E3779B97F4A7C15ull) >> 11 & (new_cap - 1);
      while (ok[h] >= 0) h = (h + 1) & (new_cap - 1);
      ok[h] = sc.keys[i];
      ov[h] = sc.vals[i];
    }
    sc.keys.swap(ok);
    sc.vals.swap(ov);
  };
  sc.keys.assign(cap, -1);
  sc.vals.assign(cap, 0);
  sc.order.clear();
  sc.arcs.clear();
  auto key = [&](int32_t s, int32_t pos) { return static_cast<int64_t>(s) * (U + 1) + pos; };
  // returns the id of (s, pos), creating it when new
  auto lookup = [&](int32_t s, int32_t pos) -> int32_t {
    const int64_t k = key(s, pos);
    size_t m = sc.keys.size() - 1;
    size_t h = (static_cast<uint64_t>(k) * 0x9E3779B97F4A7C15ull) >> 11 & m;
    while (sc.keys[h] >= 0) {
      if (sc.keys[h] == k) return sc.vals[h];
      h = (h + 1) & m;
    }
    const int32_t nid = static_cast<int32_t>(sc.order.size());
    sc.keys[h] = k;
    sc.vals[h] = nid;
    sc.order.emplace_back(s, pos);
    if (sc.order.size() * 2 > sc.keys.size()) rehash(sc.keys.size() * 2);
    return nid;
  };
  lookup(0, 0);
  for (size_t head = 0; head < sc.order.size(); ++head) {
    const auto [s, pos] = sc.order[head];
    const int32_t src = static_cast<int32_t>(head);
    // dec_out[s] is sorted by output label: the arcs that emit nothing come first, the ones
    // that emit labels[pos] are one contiguous run -- visit exactly those, in that order
    const std::vector<int32_t>& row = dec_out[s];
    int32_t n_eps, m0 = 0, m1 = 0;
    if (oi && !oi->first.empty()) {
      const int32_t* f = oi->first.data() + static_cast<int64_t>(s) * (oi->nol + 1);
      n_eps = f[1];
      if (pos < U && labels[pos] > 0 && labels[pos] < oi->nol) {
        m0 = f[labels[pos]];
        m1 = f[labels[pos] + 1];
      }
    } else {
      const auto olab_less = [&](int32_t ai, int32_t v) { return dec.arcs[ai].olabel < v; };
      n_eps = static_cast<int32_t>(std::lower_bound(row.begin(), row.end(), 1, olab_less) - row.begin());
      if (pos < U && labels[pos] > 0) {
        m0 = static_cast<int32_t>(std::lower_bound(row.begin() + n_eps, row.end(), labels[pos], olab_less) - row.begin());
        m1 = static_cast<int32_t>(std::lower_bound(row.begin() + m0, row.end(), labels[pos] + 1, olab_less) - row.begin());
      }
    }
    const int32_t n_take = n_eps + (m1 - m0);
    for (int32_t q = 0; q < n_take; ++q) {
      const Arc& a = dec.arcs[row[q < n_eps ? q : m0 + (q - n_eps)]];
      const int32_t npos = (a.olabel == 0) ? pos : pos + 1;
      const int32_t nid = lookup(a.dst, npos);
      sc.arcs.push_back({src, nid, a.label, a.weight, a.olabel});
    }
  }
  const int32_t n = static_cast<int32_t>(sc.order.size());
  sc.fin.assign(n, kNegInf);
  for (int32_t i = 0; i < n; ++i) {
    if (sc.order[i].second == U && dec.final_w[sc.order[i].first] > kDeadWeight)
      sc.fin[i] = dec.final_w[sc.order[i].first];
  }
  // co-accessibility over the reversed arcs (CSR by destination)
  sc.rptr.assign(n + 1, 0);
  for (const Arc& a : sc.arcs) sc.rptr[a.dst + 1]++;
  for (int32_t i = 0; i < n; ++i) sc.rptr[i + 1] += sc.rptr[i];
  sc.rsrc.resize(sc.arcs.size());
  sc.remap.assign(sc.rptr.begin(), sc.rptr.end() - 1);  // fill cursors
  for (const Arc& a : sc.arcs) sc.rsrc[sc.remap[a.dst]++] = a.src;
  sc.co.assign(n, 0);
  sc.stack.clear();
  for (int32_t i = 0; i < n; ++i)
    if (sc.fin[i] > kDeadWeight) {
      sc.co[i] = 1;
      sc.stack.push_back(i);
    }
  while (!sc.stack.empty()) {
    const int32_t s = sc.stack.back();
    sc.stack.pop_back();
    for (int32_t i = sc.rptr[s]; i < sc.rptr[s + 1]; ++i) {
      const int32_t p = sc.rsrc[i];
      if (!sc.co[p]) {
        sc.co[p] = 1;
        sc.stack.push_back(p);
      }
    }
  }
  sc.remap.assign(n, -1);
  int32_t m = 0;
  for (int32_t i = 0; i < n; ++i)
    if (sc.co[i]) sc.remap[i] = m++;
  out->n = m;
  out->final_w.assign(m, kNegInf);
  for (int32_t i = 0; i < n; ++i)
    if (sc.co[i]) out->final_w[sc.remap[i]] = sc.fin[i];
  out->arcs.clear();
  out->arcs.reserve(sc.arcs.size());
  for (const Arc& a : sc.arcs)
    if (sc.co[a.src] && sc.co[a.dst])
      out->arcs.push_back({sc.remap[a.src], sc.remap[a.dst], a.label, a.weight, a.olabel});
  if (m == 0 || sc.remap[0] != 0) {
    // empty language (cannot happen for CTC topologies): keep a lone start state
    out->n = 1;
    out->final_w.assign(1, kNegInf);
    out->arcs.clear();
  }
  return CTCG_OK;
}

// Runs fn(0..n-1) on up to CTCG_HOST_THREADS (default: min(8, hardware threads)) host
// threads; small batches stay on the calling thread.  fn must not throw.
void parallel_for(int32_t n, const std::function<void(int32_t)>& fn) {
  static const int max_threads = [] {
    const char* e = std::getenv("CTCG_HOST_THREADS");
    int v = e ? std::atoi(e) : 0;
    if (v <= 0) v = std::min(8u, std::max(1u, std::thread::hardware_concurrency()));
    return v;
  }();
  const int nt = std::min<int>(max_threads, n / 4);
  if (nt <= 1) {
    for (int32_t i = 0; i < n; ++i) fn(i);
    return;
  }
  std::atomic<int32_t> next(0);
  auto work = [&] {
    for (int32_t i = next.fetch_add(1); i < n; i = next.fetch_add(1)) fn(i);
  };
  std::vector<std::thread> pool;
  pool.reserve(nt - 1);
  for (int t = 1; t < nt; ++t) pool.emplace_back(work);
  work();
  for (std::thread& t : pool) t.join();
}

namespace {

// idx <- arc indices grouped by key (ptr = exclusive prefix sums of the key counts, size
// keys+1) and ordered inside each group by `less`; same result as one comparison sort on
// (key, ...) at a fraction of the cost when the groups are small (<= 3 arcs per state).
template <typename KeyFn, typename Less>
void bucket_sort(const std::vector<Arc>& arcs, const std::vector<int32_t>& ptr, KeyFn key, Less less,
                 std::vector<int32_t>* idx, std::vector<int32_t>* fill) {
  const int32_t na = static_cast<int32_t>(arcs.size());
  const int32_t nk = static_cast<int32_t>(ptr.size()) - 1;
  idx->resize(na);
  fill->assign(ptr.begin(), ptr.end() - 1);
  for (int32_t i = 0; i < na; ++i) (*idx)[(*fill)[key(arcs[i])]++] = i;
  for (int32_t k = 0; k < nk; ++k) {
    const int32_t lo = ptr[k], hi = ptr[k + 1];
    if (hi - lo < 2) continue;
    std::sort(idx->begin() + lo, idx->begin() + hi,
              [&](int32_t x, int32_t y) { return less(arcs[x], arcs[y]); });
  }
}

}  // namespace

// Concatenate graphs into CSR by destination and by source.  Rows are sorted by
// (other state, label, weight), the order fst_to_matrices uses
// (fst_utils.py:285), which also fixes the Viterbi tie-break.
//
// Two passes so that the per-graph work can run on several host threads: pass A finds each
// graph's distinct labels, a serial prefix sum fixes every graph's offsets, pass B fills the
// graph's slices of the shared arrays.
int finalize_graphs(const std::vector<RawGraph>& graphs, ctcg_graph** out) {
  ctcg_graph* g = new (std::nothrow) ctcg_graph();
  if (!g) return CTCG_ENOMEM;
  const int32_t G = static_cast<int32_t>(graphs.size());
  g->num_graphs = G;
  g->state_off.assign(G + 1, 0);
  g->lab_off.assign(G + 1, 0);
  std::vector<int32_t> arc_off(G + 1, 0);
  int64_t total_states = 0, total_arcs = 0;
  for (int32_t i = 0; i < G; ++i) {
    total_states += graphs[i].n;
    total_arcs += static_cast<int64_t>(graphs[i].arcs.size());
    if (total_states + G > INT32_MAX || total_arcs > INT32_MAX) {
      delete g;
      return CTCG_EUNSUPPORTED;
    }
    g->state_off[i + 1] = static_cast<int32_t>(total_states);
    arc_off[i + 1] = static_cast<int32_t>(total_arcs);
  }
  struct PerGraph {
    std::vector<int32_t> labs;
    int status = CTCG_OK;
    int32_t max_in = 0, max_out = 0;
    bool unit = true;
    bool in_uniform = true;  // all arcs into a state carry one label
  };
  std::vector<PerGraph> per(G);
  // ---- pass A: validation (fst_utils.py:272-274) + distinct labels ----
  parallel_for(G, [&](int32_t gi) {
    const RawGraph& r = graphs[gi];
    PerGraph& pg = per[gi];
    const int32_t n = r.n;
    // distinct labels through a per-thread "seen" table (labels are class ids, a few thousand)
    thread_local std::vector<int32_t> seen;
    for (const Arc& a : r.arcs) {
      if (a.label < 0) {
        pg.status = CTCG_EEPS;
        return;
      }
      if (a.src < 0 || a.src >= n || a.dst < 0 || a.dst >= n) {
        pg.status = CTCG_EINVAL;
        return;
      }
      if (a.label >= (1 << 22)) {  // huge label space: plain sort + unique
        pg.labs.clear();
        for (const Arc& b : r.arcs) pg.labs.push_back(b.label);
        std::sort(pg.labs.begin(), pg.labs.end());
        pg.labs.erase(std::unique(pg.labs.begin(), pg.labs.end()), pg.labs.end());
        return;
      }
      if (static_cast<size_t>(a.label) >= seen.size()) seen.resize(a.label + 1, 0);
      if (!seen[a.label]) {
        seen[a.label] = 1;
        pg.labs.push_back(a.label);
      }
    }
    for (int32_t l : pg.labs) seen[l] = 0;
    std::sort(pg.labs.begin(), pg.labs.end());
  });
  for (int32_t gi = 0; gi < G; ++gi) {
    if (per[gi].status != CTCG_OK) {
      const int st = per[gi].status;
      delete g;
      return st;
    }
    const int32_t nl = static_cast<int32_t>(per[gi].labs.size());
    g->lab_off[gi + 1] = g->lab_off[gi] + nl;
    g->max_labels = std::max(g->max_labels, nl);
    if (nl) g->max_label = std::max(g->max_label, per[gi].labs.back());
    g->max_states = std::max(g->max_states, graphs[gi].n);
    g->max_arcs = std::max(g->max_arcs, arc_off[gi + 1] - arc_off[gi]);
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
  g->lab_list.resize(g->lab_off[G]);
  g->lab_ptr.assign(static_cast<size_t>(g->lab_off[G]) + G, 0);
  g->slab_ptr.assign(static_cast<size_t>(g->lab_off[G]) + G, 0);
  g->slab_state.assign(static_cast<size_t>(total_states) + 1, 0);
  // ---- pass B: fill this graph's slices ----
  parallel_for(G, [&](int32_t gi) {
    const RawGraph& r = graphs[gi];
    PerGraph& pg = per[gi];
    const std::vector<int32_t>& labs = pg.labs;
    const int32_t n = r.n;
    const int32_t so = g->state_off[gi];
    const int32_t ro = so + gi;
    const int32_t arc_base = arc_off[gi];
    const int32_t na = static_cast<int32_t>(r.arcs.size());
    const int32_t nl = static_cast<int32_t>(labs.size());
    std::vector<int32_t> idx, ptr, fill, clab(na), row;
    for (int32_t s = 0; s < n; ++s) g->final_w[so + s] = (r.final_w[s] > kDeadWeight) ? r.final_w[s] : kNegInf;
    std::copy(labs.begin(), labs.end(), g->lab_list.begin() + g->lab_off[gi]);
    {
      thread_local std::vector<int32_t> compact;  // label -> index in labs
      const bool table = nl > 0 && labs.back() < (1 << 22);
      if (table) {
        if (static_cast<size_t>(labs.back()) >= compact.size()) compact.resize(labs.back() + 1);
        for (int32_t j = 0; j < nl; ++j) compact[labs[j]] = j;
      }
      for (int32_t i = 0; i < na; ++i) {
        clab[i] = table ? compact[r.arcs[i].label]
                        : static_cast<int32_t>(std::lower_bound(labs.begin(), labs.end(), r.arcs[i].label) - labs.begin());
        if (r.arcs[i].weight != 0.f) pg.unit = false;
      }
    }
    // by source
    ptr.assign(n + 1, 0);
    for (const Arc& a : r.arcs) ptr[a.src + 1]++;
    for (int32_t s = 0; s < n; ++s) {
      pg.max_out = std::max(pg.max_out, ptr[s + 1]);
      ptr[s + 1] += ptr[s];
    }
    bucket_sort(r.arcs, ptr, [](const Arc& a) { return a.src; }, [](const Arc& a, const Arc& b) {
      return std::tie(a.dst, a.label, a.weight) < std::tie(b.dst, b.label, b.weight);
    }, &idx, &fill);
    for (int32_t s = 0; s <= n; ++s) g->out_ptr[ro + s] = arc_base + ptr[s];
    for (int32_t i = 0; i < na; ++i) {
      const Arc& a = r.arcs[idx[i]];
      g->out_dst[arc_base + i] = a.dst;
      g->out_lab[arc_base + i] = a.label;
      g->out_clab[arc_base + i] = clab[idx[i]];
      g->out_w[arc_base + i] = a.weight;
    }
    // input determinism (fst_utils.py:250-252): no two arcs of one state share a label
    for (int32_t s = 0; s < n && pg.status == CTCG_OK; ++s) {
      if (ptr[s + 1] - ptr[s] < 2) continue;
      row.clear();
      for (int32_t i = ptr[s]; i < ptr[s + 1]; ++i) row.push_back(r.arcs[idx[i]].label);
      std::sort(row.begin(), row.end());
      if (std::adjacent_find(row.begin(), row.end()) != row.end()) pg.status = CTCG_ENONDET;
    }
    // by destination
    ptr.assign(n + 1, 0);
    for (const Arc& a : r.arcs) ptr[a.dst + 1]++;
    for (int32_t s = 0; s < n; ++s) {
      pg.max_in = std::max(pg.max_in, ptr[s + 1]);
      ptr[s + 1] += ptr[s];
    }
    bucket_sort(r.arcs, ptr, [](const Arc& a) { return a.dst; }, [](const Arc& a, const Arc& b) {
      return std::tie(a.src, a.label, a.weight) < std::tie(b.src, b.label, b.weight);
    }, &idx, &fill);
    for (int32_t s = 0; s <= n; ++s) g->in_ptr[ro + s] = arc_base + ptr[s];
    for (int32_t s = 0; s < n; ++s)
      for (int32_t i = ptr[s] + 1; i < ptr[s + 1]; ++i)
        if (r.arcs[idx[i]].label != r.arcs[idx[ptr[s]]].label) pg.in_uniform = false;
    {  // states by the compact label of their (first) in-arc: counting sort, states ascending within a label
      std::vector<int32_t> cnt(nl + 1, 0);
      for (int32_t s = 0; s < n; ++s)
        if (ptr[s] < ptr[s + 1]) cnt[clab[idx[ptr[s]]] + 1]++;
      for (int32_t j = 0; j < nl; ++j) cnt[j + 1] += cnt[j];
      for (int32_t j = 0; j <= nl; ++j) g->slab_ptr[g->lab_off[gi] + gi + j] = so + cnt[j];
      std::vector<int32_t> pos(cnt.begin(), cnt.end() - 1);
      for (int32_t s = 0; s < n; ++s)
        if (ptr[s] < ptr[s + 1]) g->slab_state[so + pos[clab[idx[ptr[s]]]]++] = s;
    }
    for (int32_t i = 0; i < na; ++i) {
      const Arc& a = r.arcs[idx[i]];
      g->in_src[arc_base + i] = a.src;
      g->in_lab[arc_base + i] = a.label;
      g->in_clab[arc_base + i] = clab[idx[i]];
      g->in_w[arc_base + i] = a.weight;
    }
    // by compact label (then source, destination): fixed summation order of the posteriors
    ptr.assign(nl + 1, 0);
    for (int32_t i = 0; i < na; ++i) ptr[clab[i] + 1]++;
    for (int32_t j = 0; j < nl; ++j) ptr[j + 1] += ptr[j];
    const Arc* base = r.arcs.data();
    bucket_sort(r.arcs, ptr, [&](const Arc& a) { return clab[&a - base]; }, [](const Arc& a, const Arc& b) {
      return std::tie(a.src, a.dst, a.weight) < std::tie(b.src, b.dst, b.weight);
    }, &idx, &fill);
    for (int32_t j = 0; j <= nl; ++j) g->lab_ptr[g->lab_off[gi] + gi + j] = arc_base + ptr[j];
    for (int32_t i = 0; i < na; ++i) {
      const Arc& a = r.arcs[idx[i]];
      g->lab_src[arc_base + i] = a.src;
      g->lab_dst[arc_base + i] = a.dst;
      g->lab_w[arc_base + i] = a.weight;
    }
  });
  for (int32_t gi = 0; gi < G; ++gi) {
    if (per[gi].status != CTCG_OK) {
      const int st = per[gi].status;
      delete g;
      return st;
    }
    g->max_in_degree = std::max(g->max_in_degree, per[gi].max_in);
    g->max_out_degree = std::max(g->max_out_degree, per[gi].max_out);
    if (!per[gi].unit) g->unit_weights = 0;
    if (!per[gi].in_uniform) g->in_label_uniform = 0;
  }
  *out = g;
  return CTCG_OK;
}

}  // namespace ctcg
