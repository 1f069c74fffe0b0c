"""CPU restatement of the array post-processing in the reference's ``ctc_forward.py``.

TEST INFRASTRUCTURE ONLY (see ``oracle/__init__.py``).  The reference performs these steps inline in
``main()`` between the model call and the Kaldi writer, so they cannot be imported; the numpy
statements below are the reference's own, line for line cited, with the dtype as a parameter
(the reference runs them in float32)."""
from __future__ import annotations

import numpy as np

EPSILON = 1e-30  # ctc_forward.py:27


def postprocess(logprobs, transfer_hash_prob=False, imitate_biphones=False, block_normalize=False,
                block_marginalize=False, dtype=np.float64):
    logprobs = np.array(logprobs, dtype=dtype, copy=True)
    if transfer_hash_prob:                                      # ctc_forward.py:96-103
        blank_probs = np.exp(logprobs[:, :, 0])
        hash_probs = np.exp(logprobs[:, :, 3])
        blank_probs += hash_probs - EPSILON
        hash_probs = EPSILON
        logprobs[:, :, 0] = np.log(blank_probs)
        logprobs[:, :, 3] = np.log(hash_probs)
    t, bsz, num_classes = logprobs.shape                        # :105
    if imitate_biphones:                                        # :107-115
        logprobs = np.tile(logprobs, (1, 1, num_classes))
        num_classes = num_classes ** 2
        if not block_normalize:
            z = np.exp(logprobs).sum(axis=2, keepdims=True)
            logprobs -= np.log(z + EPSILON)
    elif block_normalize:                                       # :116-120
        num_mono = int(np.round(num_classes ** 0.5))
        z = np.exp(logprobs).reshape(t, bsz, num_mono, num_mono)
        z = z.sum(axis=3).repeat(num_mono, axis=2)
        logprobs -= np.log(z + EPSILON)
    elif block_marginalize:                                     # :121-128
        num_symbols = int(np.round(num_classes ** 0.5))
        probs = np.exp(logprobs)
        probs = (probs.reshape(t, bsz, num_symbols, num_symbols).sum(axis=2) / num_symbols)
        logprobs = np.log(probs)
        assert not np.any(np.isnan(logprobs))
    return logprobs
