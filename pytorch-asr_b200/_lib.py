"""ctypes binding of ``libctcg_b200.so`` (the C ABI declared in ``include/ctcg_b200.h``).

There is deliberately no fallback: if the shared library is missing, or a
compute entry point is called without a CUDA device, this raises.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "libctcg_b200.so")

OK, EINVAL, ESTART, ENONDET, EEPS, ECUDA, ENOMEM, EWORKSPACE, EUNSUPPORTED, ENOTUPLOADED = range(10)
NORM_NONE, NORM_SOFTMAX, NORM_CONTEXT = 0, 1, 2

# every symbol include/ctcg_b200.h declares
EXPORTS = (
    "ctcg_decoding_graph", "ctcg_training_graphs", "ctcg_graph_from_padded",
    "ctcg_graph_from_arcs", "ctcg_graph_get_info", "ctcg_graph_padded_dims",
    "ctcg_graph_export_padded", "ctcg_graph_upload", "ctcg_graph_free",
    "ctcg_workspace_bytes", "ctcg_workspace_bytes_norm", "ctcg_loss_fwd_bwd", "ctcg_path_sum", "ctcg_viterbi", "ctcg_viterbi_static", "ctcg_graph_set_generic", "ctcg_scale_grad",
    "ctcg_decoder_from_arcs", "ctcg_decoder_num_states", "ctcg_decoder_graph", "ctcg_decoder_training_graphs",
    "ctcg_decoder_free", "ctcg_graph_cache_info", "ctcg_graph_cache_clear",
    "ctcg_path_count", "ctcg_loss_host", "ctcg_upload_logits", "ctcg_strerror", "ctcg_last_error_detail", "ctcg_step_graph_info", "ctcg_launch_count",
    "ctcg_profile_enable", "ctcg_profile_collect", "ctcg_logits_post", "ctcg_logits_post_width",
)


class Topology(C.Structure):
    _fields_ = [("context_order", C.c_int32), ("num_symbols", C.c_int32),
                ("allow_nonblank_selfloops", C.c_int32),
                ("use_contextual_blanks", C.c_int32),
                ("loop_using_symbol_repetitions", C.c_int32)]


class GraphInfo(C.Structure):
    _fields_ = [("num_graphs", C.c_int32), ("max_states", C.c_int32),
                ("max_in_degree", C.c_int32), ("max_out_degree", C.c_int32),
                ("max_labels", C.c_int32), ("structured", C.c_int32),
                ("on_device", C.c_int32), ("reserved", C.c_int32),
                ("total_states", C.c_int64), ("total_arcs", C.c_int64)]


_lib = None


def lib():
    """Load the shared library (once).  Fails loudly when it has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(
            "libctcg_b200.so is not built (%s missing): run `python -c 'import "
            "__graft_entry__ as g; g.build()'`.  There is no CPU fallback." % LIB_PATH)
    L = C.CDLL(LIB_PATH)
    vp, i32, i64, f32p = C.c_void_p, C.c_int32, C.c_int64, C.c_void_p
    L.ctcg_decoding_graph.argtypes = [C.POINTER(Topology), C.POINTER(vp)]
    L.ctcg_training_graphs.argtypes = [C.POINTER(Topology), vp, vp, i32, C.POINTER(vp)]
    L.ctcg_graph_from_padded.argtypes = [vp] * 7 + [i32] * 4 + [C.c_float, C.POINTER(vp)]
    L.ctcg_graph_from_arcs.argtypes = [vp] * 7 + [i32, C.POINTER(vp)]
    L.ctcg_graph_get_info.argtypes = [vp, C.POINTER(GraphInfo)]
    L.ctcg_graph_padded_dims.argtypes = [vp, i32] + [C.POINTER(i32)] * 3
    L.ctcg_graph_export_padded.argtypes = [vp, i32, i32, C.c_float, vp, vp, vp, vp]
    L.ctcg_graph_upload.argtypes = [vp, vp]
    L.ctcg_graph_free.argtypes = [vp]
    L.ctcg_graph_free.restype = None
    L.ctcg_graph_set_generic.argtypes = [vp]
    L.ctcg_graph_set_generic.restype = None
    L.ctcg_workspace_bytes.argtypes = [i32, i32, i32, vp, vp]
    L.ctcg_workspace_bytes.restype = C.c_size_t
    L.ctcg_workspace_bytes_norm.argtypes = [i32, i32, i32, vp, vp, i32]
    L.ctcg_workspace_bytes_norm.restype = C.c_size_t
    L.ctcg_loss_fwd_bwd.argtypes = [f32p, i32, i32, i32, vp, vp, vp, i32, i32, vp, vp, vp, vp,
                                    C.c_size_t, vp]
    L.ctcg_path_sum.argtypes = [f32p, i32, i32, i32, vp, vp, vp, vp, vp, C.c_size_t, vp]
    L.ctcg_viterbi.argtypes = [f32p, i32, i32, i32, vp, vp, vp, vp, vp, C.c_size_t, vp]
    L.ctcg_viterbi_static.argtypes = [f32p, i32, i32, i32, vp, vp, vp, vp, vp]
    L.ctcg_path_count.argtypes = [i32, vp, vp, vp, vp, C.c_size_t, vp]
    L.ctcg_scale_grad.argtypes = [vp, i32, i32, i32, vp, i32, vp]
    L.ctcg_decoder_from_arcs.argtypes = [i32, i64, vp, vp, vp, vp, vp, vp, i32, C.POINTER(vp)]
    L.ctcg_decoder_num_states.argtypes = [vp]
    L.ctcg_decoder_num_states.restype = i32
    L.ctcg_decoder_graph.argtypes = [vp, C.POINTER(vp)]
    L.ctcg_decoder_training_graphs.argtypes = [vp, vp, vp, i32, i32, C.POINTER(vp)]
    L.ctcg_decoder_free.argtypes = [vp]
    L.ctcg_decoder_free.restype = None
    L.ctcg_graph_cache_info.argtypes = [C.POINTER(C.c_uint64)] * 3
    L.ctcg_graph_cache_info.restype = None
    L.ctcg_graph_cache_clear.argtypes = []
    L.ctcg_graph_cache_clear.restype = None
    L.ctcg_loss_host.argtypes = [vp, i32, i32, i32, vp, vp, vp, i32, i32, vp, vp]
    L.ctcg_upload_logits.argtypes = [vp, i32, i32, i32, vp, vp, vp]
    L.ctcg_logits_post_width.argtypes = [i32, i32]
    L.ctcg_logits_post_width.restype = i32
    L.ctcg_logits_post.argtypes = [f32p, i32, i32, i32, i32, vp, vp]
    L.ctcg_strerror.argtypes = [C.c_int]
    L.ctcg_strerror.restype = C.c_char_p
    L.ctcg_step_graph_info.argtypes = [C.POINTER(C.c_uint64)] * 3
    L.ctcg_step_graph_info.restype = None
    L.ctcg_last_error_detail.argtypes = []
    L.ctcg_last_error_detail.restype = C.c_char_p
    L.ctcg_launch_count.argtypes = []
    L.ctcg_launch_count.restype = C.c_uint64
    L.ctcg_profile_enable.argtypes = [C.c_int]
    L.ctcg_profile_enable.restype = None
    L.ctcg_profile_collect.argtypes = [C.c_char_p, C.c_size_t, vp, C.c_int]
    for name in EXPORTS:
        getattr(L, name)
    _lib = L
    return L


def check(status):
    """Turn a ctcg_status into the exception the reference would raise."""
    if status == OK:
        return
    msg = lib().ctcg_strerror(status).decode()
    if status in (ESTART, ENONDET, EEPS):   # fst_utils.py:246-252, :272-274 raise ValueError
        raise ValueError(msg)
    if status == EINVAL:
        raise AssertionError("ctcg: " + msg)  # the reference asserts on shape mismatches
    detail = lib().ctcg_last_error_detail().decode() if status in (ECUDA, EUNSUPPORTED) else ""
    raise RuntimeError("ctcg: %s (status %d)%s" % (msg, status, ": " + detail if detail else ""))


def launch_count():
    return int(lib().ctcg_launch_count())


def profile(enable):
    lib().ctcg_profile_enable(int(bool(enable)))


def profile_collect(max_records=4096):
    """[(kernel name, milliseconds)] for every launch since profiling was enabled."""
    names = C.create_string_buffer(max_records * 32)
    ms = np.zeros(max_records, dtype=np.float32)
    n = lib().ctcg_profile_collect(names, len(names), _np_ptr(ms), max_records)
    return list(zip(names.value.decode().split(";")[:n], ms[:n].tolist()))


def _np_ptr(a):
    return a.ctypes.data_as(C.c_void_p)


def _stream_ptr(device):
    return C.c_void_p(torch.cuda.current_stream(device).cuda_stream)


class _on_device(object):
    """``with torch.cuda.device(d)`` costs ~15 us per use even when ``d`` is already current; the loss path enters
    it several times per step.  This switches only when it has to."""
    __slots__ = ("idx", "prev")

    def __init__(self, device):
        self.idx = device.index if device.index is not None else torch.cuda.current_device()
        self.prev = -1

    def __enter__(self):
        cur = torch.cuda.current_device()
        if cur != self.idx:
            self.prev = cur
            torch.cuda.set_device(self.idx)

    def __exit__(self, *exc):
        if self.prev >= 0:
            torch.cuda.set_device(self.prev)
        return False


def _require_cuda(t, what):
    if not (torch.is_tensor(t) and t.is_cuda):
        raise RuntimeError("ctcg: %s must be a CUDA tensor -- this engine has no CPU path" % what)


class Graph(object):
    """Owns one ``ctcg_graph*`` (a shared graph or one graph per utterance)."""

    def __init__(self, ptr):
        self._ptr = C.c_void_p(ptr) if not isinstance(ptr, C.c_void_p) else ptr
        self._info = None

    def __del__(self):
        try:
            if self._ptr and _lib is not None:
                _lib.ctcg_graph_free(self._ptr)
        except Exception:
            pass
        self._ptr = None

    # -- constructors --------------------------------------------------
    @staticmethod
    def _topology(context_order, num_symbols, allow_nonblank_selfloops=True,
                  use_contextual_blanks=False, loop_using_symbol_repetitions=False,
                  eval_repeats_in_context=False):
        assert not eval_repeats_in_context  # fst_utils.py:759-760
        return Topology(int(context_order), int(num_symbols), int(bool(allow_nonblank_selfloops)),
                        int(bool(use_contextual_blanks)), int(bool(loop_using_symbol_repetitions)))

    @classmethod
    def decoding(cls, context_order, num_symbols, **build_args):
        topo = cls._topology(context_order, num_symbols, **build_args)
        out = C.c_void_p()
        check(lib().ctcg_decoding_graph(C.byref(topo), C.byref(out)))
        return cls(out)

    @classmethod
    def training(cls, context_order, num_symbols, label_seqs, **build_args):
        """``label_seqs``: list of 1-D integer sequences (one transcript per utterance)."""
        topo = cls._topology(context_order, num_symbols, **build_args)
        seqs = [np.asarray(s, dtype=np.int64).reshape(-1) for s in label_seqs]
        offsets = np.zeros(len(seqs) + 1, dtype=np.int64)
        for i, s in enumerate(seqs):
            offsets[i + 1] = offsets[i] + len(s)
        flat = (np.concatenate(seqs) if seqs else np.zeros(0, np.int64)).astype(np.int32)
        flat = np.ascontiguousarray(flat)
        out = C.c_void_p()
        check(lib().ctcg_training_graphs(C.byref(topo), _np_ptr(flat), _np_ptr(offsets),
                                         len(seqs), C.byref(out)))
        return cls(out)

    @classmethod
    def from_padded(cls, matrices, nc_weight=-1e20):
        """From the reference's 4- or 8-tuple of padded (G,N,K) matrices."""
        mats = [m.detach().cpu().numpy() if torch.is_tensor(m) else np.asarray(m) for m in matrices]
        pred = np.ascontiguousarray(mats[0], dtype=np.int64)
        lab = np.ascontiguousarray(mats[1], dtype=np.int64)
        w = np.ascontiguousarray(mats[2], dtype=np.float32)
        fin = np.ascontiguousarray(mats[3], dtype=np.float32).reshape(pred.shape[0], pred.shape[1])
        assert pred.ndim == 3 and lab.shape == pred.shape and w.shape == pred.shape
        G, N, K = pred.shape
        out = C.c_void_p()
        check(lib().ctcg_graph_from_padded(_np_ptr(pred), _np_ptr(lab), _np_ptr(w), _np_ptr(fin),
                                           None, None, None, G, N, K, 0, float(nc_weight),
                                           C.byref(out)))
        return cls(out)

    def set_generic(self):
        """Drop the topology tag: the graph takes the generic CSR kernels (cross-checks)."""
        lib().ctcg_graph_set_generic(self.ptr)
        return self

    @classmethod
    def from_arcs(cls, graphs):
        """``graphs``: list of ``(n_states, arcs[(src,dst,label,weight)], final[n])``."""
        src, dst, lab, w, fin, counts, offs = [], [], [], [], [], [], [0]
        for (n, arcs, final) in graphs:
            counts.append(n)
            for a in arcs:
                src.append(a[0]); dst.append(a[1]); lab.append(a[2]); w.append(a[3])
            offs.append(len(src))
            fin.extend(np.asarray(final, dtype=np.float32).tolist())
        a32 = lambda x: np.ascontiguousarray(np.asarray(x, dtype=np.int32))  # noqa: E731
        src, dst, lab, counts = a32(src), a32(dst), a32(lab), a32(counts)
        w = np.ascontiguousarray(np.asarray(w, dtype=np.float32))
        fin = np.ascontiguousarray(np.asarray(fin, dtype=np.float32))
        offs = np.ascontiguousarray(np.asarray(offs, dtype=np.int64))
        out = C.c_void_p()
        check(lib().ctcg_graph_from_arcs(_np_ptr(src), _np_ptr(dst), _np_ptr(lab), _np_ptr(w),
                                         _np_ptr(offs), _np_ptr(counts), _np_ptr(fin),
                                         len(graphs), C.byref(out)))
        return cls(out)

    # -- inspection ----------------------------------------------------
    @property
    def ptr(self):
        return self._ptr

    def info(self):
        gi = GraphInfo()
        check(lib().ctcg_graph_get_info(self._ptr, C.byref(gi)))
        return gi

    @property
    def num_graphs(self):
        return self.info().num_graphs

    def padded_dims(self, index=0):
        n, ki, ko = C.c_int32(), C.c_int32(), C.c_int32()
        check(lib().ctcg_graph_padded_dims(self._ptr, index, C.byref(n), C.byref(ki), C.byref(ko)))
        return n.value, ki.value, ko.value

    def export_padded(self, index=0, out_edges=False, nc_weight=-1e20):
        """Graph ``index`` in the exact encoding of ``fst_to_matrices`` (numpy)."""
        n, ki, ko = self.padded_dims(index)
        k = ko if out_edges else ki
        states = np.zeros((n, k), dtype=np.int64)
        labels = np.zeros((n, k), dtype=np.int64)
        weights = np.zeros((n, k), dtype=np.float32)
        terminal = np.zeros((n, 1), dtype=np.float32)
        check(lib().ctcg_graph_export_padded(self._ptr, index, int(out_edges), float(nc_weight),
                                             _np_ptr(states), _np_ptr(labels), _np_ptr(weights),
                                             _np_ptr(terminal)))
        return states, labels, weights, terminal

    def upload(self, device=None):
        if device is not None and getattr(self, "_on", None) == device:   # already there (one device copy per handle)
            return self
        if not torch.cuda.is_available():
            raise RuntimeError("ctcg: no CUDA device -- this engine has no CPU path")
        device = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
        with _on_device(device):
            check(lib().ctcg_graph_upload(self._ptr, _stream_ptr(device)))
        self._on = device
        return self


# ---------------------------------------------------------------------------
_workspaces = {}


def step_graph_info():
    """``(replays, captures, ordinary)`` of the captured-step cache of ``ctcg_loss_fwd_bwd``."""
    a, b, c = C.c_uint64(), C.c_uint64(), C.c_uint64()
    lib().ctcg_step_graph_info(C.byref(a), C.byref(b), C.byref(c))
    return int(a.value), int(b.value), int(c.value)


def graph_cache_info():
    """``(hits, misses, entries)`` of the per-transcript cache of composed training graphs."""
    h, m, n = C.c_uint64(), C.c_uint64(), C.c_uint64()
    lib().ctcg_graph_cache_info(C.byref(h), C.byref(m), C.byref(n))
    return int(h.value), int(m.value), int(n.value)


def graph_cache_clear():
    lib().ctcg_graph_cache_clear()


class Decoder(object):
    """A decoding graph given arc by arc with output labels (``ctcg_decoder_*``): the tri-char
    topologies and anything else the two built-in topologies do not cover."""

    def __init__(self, n_states, arcs, final, connect=True):
        """``arcs``: (n, 5) integers ``src, dst, class label, emitted symbol (0 = none), 0`` in insertion
        order, or a tuple of five arrays (the fifth, weights, may be None); ``final``: n floats."""
        if isinstance(arcs, tuple):
            src, dst, lab, olab, w = arcs
        else:
            a = np.asarray(arcs, dtype=np.int64).reshape(-1, 5)
            src, dst, lab, olab, w = a[:, 0], a[:, 1], a[:, 2], a[:, 3], None
        a32 = lambda x: np.ascontiguousarray(np.asarray(x, dtype=np.int32))  # noqa: E731
        src, dst, lab, olab = a32(src), a32(dst), a32(lab), a32(olab)
        w = None if w is None else np.ascontiguousarray(np.asarray(w, dtype=np.float32))
        fin = np.ascontiguousarray(np.asarray(final, dtype=np.float32))
        assert fin.shape[0] == n_states
        out = C.c_void_p()
        check(lib().ctcg_decoder_from_arcs(int(n_states), int(src.shape[0]), _np_ptr(src), _np_ptr(dst), _np_ptr(lab),
                                           _np_ptr(olab), _np_ptr(w) if w is not None else None, _np_ptr(fin),
                                           int(bool(connect)), C.byref(out)))
        self._ptr = out

    @property
    def num_states(self):
        return int(lib().ctcg_decoder_num_states(self._ptr))

    def graph(self):
        """The decoding (denominator) graph; ``ValueError`` when it is not input-deterministic."""
        out = C.c_void_p()
        check(lib().ctcg_decoder_graph(self._ptr, C.byref(out)))
        return Graph(out)

    def training(self, label_seqs, modulo=0):
        seqs = [np.asarray(s, dtype=np.int64).reshape(-1) for s in label_seqs]
        offsets = np.zeros(len(seqs) + 1, dtype=np.int64)
        for i, s in enumerate(seqs):
            offsets[i + 1] = offsets[i] + len(s)
        flat = np.ascontiguousarray((np.concatenate(seqs) if seqs else np.zeros(0, np.int64)).astype(np.int32))
        out = C.c_void_p()
        check(lib().ctcg_decoder_training_graphs(self._ptr, _np_ptr(flat), _np_ptr(offsets), len(seqs), int(modulo),
                                                 C.byref(out)))
        return Graph(out)

    def __del__(self):
        try:
            if getattr(self, "_ptr", None):
                lib().ctcg_decoder_free(self._ptr)
                self._ptr = None
        except Exception:
            pass


def _workspace(device, nbytes):
    key = (device.index, torch.cuda.current_stream(device).cuda_stream)
    ws = _workspaces.get(key)
    if ws is None or ws.numel() < nbytes:
        ws = torch.empty(max(int(nbytes), 1 << 20), dtype=torch.uint8, device=device)
        _workspaces[key] = ws
    return ws


def _lens_array(act_lens, B):
    if torch.is_tensor(act_lens):
        act_lens = act_lens.detach().cpu().numpy()
    lens = np.ascontiguousarray(np.asarray(act_lens, dtype=np.int32).reshape(-1))
    assert lens.shape[0] == B, "act_lens must hold one length per utterance"
    return lens


def _prep_logits(x):
    _require_cuda(x, "logits / log_probs")
    assert x.dim() == 3, "expected a (T, B, V) tensor"
    x = x.detach()
    if x.dtype != torch.float32:
        x = x.float()
    x = x.contiguous()
    if x.data_ptr() % 16:        # a view into a larger buffer: the TMA row loads want 16-byte aligned bases
        x = x.clone()
    return x


def loss_fwd_bwd(logits, act_lens, num, den=None, normalize=NORM_NONE, num_symbols=0,
                 grad_scale=None, need_grad=True):
    """``ctcg_loss_fwd_bwd``: returns ``(loss (B,), grad (T,B,V) or None)``."""
    x = _prep_logits(logits)
    T, B, V = x.shape
    lens = _lens_array(act_lens, B)
    dev = x.device
    with _on_device(dev):
        num.upload(dev)
        if den is not None:
            den.upload(dev)
        nbytes = lib().ctcg_workspace_bytes_norm(T, B, V, num.ptr, den.ptr if den is not None else None, int(normalize))
        ws = _workspace(dev, nbytes)
        loss = torch.empty(B, dtype=torch.float32, device=dev)
        grad = torch.empty_like(x) if need_grad else None
        gs = None
        if grad_scale is not None:
            gs = grad_scale.detach().to(device=dev, dtype=torch.float32).contiguous()
        check(lib().ctcg_loss_fwd_bwd(
            x.data_ptr(), T, B, V, _np_ptr(lens), num.ptr, den.ptr if den is not None else None,
            int(normalize), int(num_symbols), gs.data_ptr() if gs is not None else None,
            loss.data_ptr(), grad.data_ptr() if grad is not None else None,
            ws.data_ptr(), ws.numel(), _stream_ptr(dev)))
    return loss, grad


def path_sum(log_probs, act_lens, graph, need_grad=True):
    x = _prep_logits(log_probs)
    T, B, V = x.shape
    lens = _lens_array(act_lens, B)
    dev = x.device
    with torch.cuda.device(dev):
        graph.upload(dev)
        ws = _workspace(dev, lib().ctcg_workspace_bytes(T, B, V, graph.ptr, None))
        cost = torch.empty(B, dtype=torch.float32, device=dev)
        grad = torch.empty_like(x) if need_grad else None
        check(lib().ctcg_path_sum(x.data_ptr(), T, B, V, _np_ptr(lens), graph.ptr, cost.data_ptr(),
                                  grad.data_ptr() if grad is not None else None,
                                  ws.data_ptr(), ws.numel(), _stream_ptr(dev)))
    return cost, grad


def viterbi(log_probs, act_lens, graph):
    x = _prep_logits(log_probs)
    T, B, V = x.shape
    lens = _lens_array(act_lens, B)
    dev = x.device
    with torch.cuda.device(dev):
        graph.upload(dev)
        ws = _workspace(dev, lib().ctcg_workspace_bytes(T, B, V, graph.ptr, None))
        score = torch.empty(B, dtype=torch.float32, device=dev)
        labels = torch.empty((T, B), dtype=torch.int32, device=dev)
        check(lib().ctcg_viterbi(x.data_ptr(), T, B, V, _np_ptr(lens), graph.ptr, score.data_ptr(),
                                 labels.data_ptr(), ws.data_ptr(), ws.numel(), _stream_ptr(dev)))
    return score, labels


def scale_grad(grad, grad_output):
    """In place ``grad[t, b, :] *= grad_output[b]`` on the device (``PathLogSumExp.backward``,
    fst_utils.py:482-485); rows whose factor is exactly 1 cost nothing."""
    _require_cuda(grad, "grad")
    T, B, V = grad.shape
    go = grad_output
    assert go.numel() == B or go.numel() == 1
    stride = 0 if (go.numel() == 1 or go.stride() == (0,)) else 1
    go = go if stride == 0 else go.contiguous()
    go = go.to(device=grad.device, dtype=torch.float32)
    with _on_device(grad.device):
        check(lib().ctcg_scale_grad(grad.data_ptr(), T, B, V, go.data_ptr(), stride, _stream_ptr(grad.device)))
    return grad


STATIC_SCRATCH_BYTES = 32 << 20


def viterbi_static(log_probs, lens_dev, graph, score, labels):
    """``ctcg_viterbi_static``: Viterbi with no allocation on the call path (the scratch is a device
    array built into the library).  Everything is passed in by the caller: ``lens_dev`` int32 (B,) on
    the device, ``graph`` already uploaded, ``score`` float32 (B,) and ``labels`` int32 (T, B)
    outputs.  Calls that overlap on the GPU race on the shared scratch."""
    _require_cuda(log_probs, "log_probs")
    T, B, V = log_probs.shape
    assert log_probs.dtype == torch.float32 and log_probs.is_contiguous()
    assert lens_dev.is_cuda and lens_dev.dtype == torch.int32 and lens_dev.numel() == B
    assert score.is_cuda and score.dtype == torch.float32 and score.numel() == B
    assert labels.is_cuda and labels.dtype == torch.int32 and tuple(labels.shape) == (T, B) and labels.is_contiguous()
    dev = log_probs.device
    with torch.cuda.device(dev):
        check(lib().ctcg_viterbi_static(log_probs.data_ptr(), T, B, V, lens_dev.data_ptr(), graph.ptr,
                                        score.data_ptr(), labels.data_ptr(), _stream_ptr(dev)))
    return score, labels


def path_count(act_lens, graph, device=None):
    if not torch.cuda.is_available():
        raise RuntimeError("ctcg: no CUDA device -- this engine has no CPU path")
    dev = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
    lens = np.ascontiguousarray(np.asarray(act_lens, dtype=np.int32).reshape(-1))
    B = lens.shape[0]
    with torch.cuda.device(dev):
        graph.upload(dev)
        ws = _workspace(dev, lib().ctcg_workspace_bytes(1, B, 1, None, None))
        counts = torch.zeros(B, dtype=torch.int64, device=dev)
        check(lib().ctcg_path_count(B, _np_ptr(lens), graph.ptr, counts.data_ptr(),
                                    ws.data_ptr(), ws.numel(), _stream_ptr(dev)))
    return counts.cpu().numpy().view(np.uint64)


POST_TRANSFER_HASH, POST_IMITATE_BIPHONES, POST_BLOCK_NORMALIZE, POST_BLOCK_MARGINALIZE = 1, 2, 4, 8


def logits_post(log_probs, flags):
    """``ctcg_logits_post``: the array transformations of ctc_forward.py:96-128 on the device."""
    x = _prep_logits(log_probs)
    T, B, V = x.shape
    W = lib().ctcg_logits_post_width(V, int(flags))
    if W < 1:
        raise ValueError("ctcg: %d classes are not a square number of symbols" % V)
    out = torch.empty((T, B, W), dtype=torch.float32, device=x.device)
    with torch.cuda.device(x.device):
        check(lib().ctcg_logits_post(x.data_ptr(), T, B, V, int(flags), out.data_ptr(), _stream_ptr(x.device)))
    return out


def upload_logits(logits_host, act_lens, device=None, out=None):
    """``ctcg_upload_logits``: (T, B, V) host logits -> device, moving only the frames t < act_lens[b] of a
    length-sorted batch (the padding of the result is zeros).  Asynchronous on the current stream; pass a pinned
    tensor.  Returns the device tensor."""
    if not torch.cuda.is_available():
        raise RuntimeError("ctcg: no CUDA device -- this engine has no CPU path")
    x = logits_host
    assert torch.is_tensor(x) and not x.is_cuda and x.dtype == torch.float32 and x.is_contiguous() and x.dim() == 3
    T, B, V = x.shape
    lens = _lens_array(act_lens, B)
    dev = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
    if out is None:
        out = torch.empty((T, B, V), dtype=torch.float32, device=dev)
    assert out.is_cuda and out.is_contiguous() and out.shape == x.shape and out.dtype == torch.float32
    with torch.cuda.device(out.device):
        check(lib().ctcg_upload_logits(x.data_ptr(), T, B, V, _np_ptr(lens), out.data_ptr(), _stream_ptr(out.device)))
    return out


def loss_host(logits_host, act_lens, num, den=None, normalize=NORM_NONE, num_symbols=0,
              loss_out=None, grad_out=None, want_grad=True):
    """``ctcg_loss_host``: host buffers in, host buffers out (copies inside)."""
    if not torch.cuda.is_available():
        raise RuntimeError("ctcg: no CUDA device -- this engine has no CPU path")
    x = logits_host
    assert torch.is_tensor(x) and not x.is_cuda and x.dtype == torch.float32 and x.is_contiguous()
    T, B, V = x.shape
    lens = _lens_array(act_lens, B)
    if loss_out is None:
        loss_out = torch.empty(B, dtype=torch.float32)
    if want_grad and grad_out is None:
        grad_out = torch.empty_like(x)
    num._on = None        # the C call uploads to the current device itself: drop the wrapper's "already there" marks
    if den is not None:
        den._on = None
    check(lib().ctcg_loss_host(x.data_ptr(), T, B, V, _np_ptr(lens), num.ptr,
                               den.ptr if den is not None else None, int(normalize), int(num_symbols),
                               loss_out.data_ptr(), grad_out.data_ptr() if want_grad else None))
    return loss_out, (grad_out if want_grad else None)
