"""B200-native CTC / CTC-G training-loss engine for pytorch-asr's hot path.

The directory name ``pytorch-asr_b200`` is not a Python identifier; import it as
``pytorch_asr_b200`` (the shim module of that name at the repository root loads
this package).  Layout:

* ``csrc/``          CUDA kernels (sm_100a), the C++ graph compiler, the C ABI
* ``_lib.py``        ctypes binding of ``libctcg_b200.so`` (``include/ctcg_b200.h``)
* ``fst_utils.py``   mirror of ``att_speech/fst_utils.py`` (path_reduction, CTCGraphGen)
* ``ctc_losses.py``  mirror of the FST losses + ``FSTDecoder`` loss glue
* ``ctc_forward.py`` mirror of the logits post-processing of the reference's ctc_forward.py
* ``parallel.py``    utterance-sharded multi-GPU driver (one process per GPU)
"""
from . import _lib  # noqa: F401
from ._lib import Graph, NORM_CONTEXT, NORM_NONE, NORM_SOFTMAX, upload_logits  # noqa: F401
from .fst_utils import (CTCGraphGen, PackedGraphs, batch_training_graph_matrices,  # noqa: F401
                        collapse_frame_labels, make_full_ngram_table, padded_matrices, path_logsumexp, path_reduction)
from .ctc_forward import postprocess_logprobs  # noqa: F401
from .ctc_losses import (FSTLoss, FSTLossMixin, ctc_fst_global_loss, ctc_fst_loss,  # noqa: F401
                         ctc_raw_loss, ctc_raw_loss_batch, fst_loss, get_normalized_acts)

__all__ = [
    "Graph", "upload_logits", "CTCGraphGen", "PackedGraphs", "path_reduction", "path_logsumexp",
    "batch_training_graph_matrices", "padded_matrices", "make_full_ngram_table", "collapse_frame_labels",
    "fst_loss", "ctc_fst_loss", "ctc_fst_global_loss", "ctc_raw_loss", "ctc_raw_loss_batch", "get_normalized_acts",
    "FSTLoss", "FSTLossMixin", "NORM_NONE", "NORM_SOFTMAX", "NORM_CONTEXT", "postprocess_logprobs",
]
