"""Device version of the array post-processing in the reference's ``ctc_forward.py`` (:96-128):
what happens to the model's ``(T, B, V)`` log-probabilities between ``model.decoder.logits`` and the
Kaldi archive writer.  Model loading, the dataset loop and the Kaldi IO stay in the reference."""
from __future__ import annotations

from . import _lib


def postprocess_logprobs(logprobs, transfer_hash_prob=False, imitate_biphones=False,
                         block_normalize=False, block_marginalize=False):
    """Same flags and precedence as the command line of ctc_forward.py (:52-63, :96-128):

    * ``transfer_hash_prob``: move the probability of class 3 (``#``) onto class 0 (blank);
    * ``imitate_biphones``: tile a mono-char output to ``V*V`` columns, renormalised unless
      ``block_normalize`` is also given;
    * else ``block_normalize``: normalise inside each context block of ``S`` classes (``V = S*S``);
    * else ``block_marginalize``: average over contexts, ``S`` columns out.

    ``logprobs``: CUDA fp32 ``(T, B, V)``; returns a new CUDA tensor ``(T, B, W)``.
    Utterance ``i`` is then written as ``out[:encoded_lens[i], i, :]`` (:129-131)."""
    flags = ((_lib.POST_TRANSFER_HASH if transfer_hash_prob else 0) |
             (_lib.POST_IMITATE_BIPHONES if imitate_biphones else 0) |
             (_lib.POST_BLOCK_NORMALIZE if block_normalize else 0) |
             (_lib.POST_BLOCK_MARGINALIZE if block_marginalize else 0))
    return _lib.logits_post(logprobs, flags)
