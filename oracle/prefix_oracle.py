"""CPU oracle for kernel (3), part 1: the CTC prefix scorer.

TEST INFRASTRUCTURE ONLY -- never imported by the product package.

numpy float32 restatement of the reference's `CTCPrefixScore`
(src/ctc.py:9-101): `blank_only_state` = init_state (ctc.py:19-27),
`extend_prefix` = cheap_compute (ctc.py:65-101) and `extend_prefix_all` =
full_compute (ctc.py:29-63).  The reference departs from the ESPnet scorer it
cites (ctc.py:5-6) in ways that change decoded tokens, all reproduced here
(SURVEY.md section 3.4):
  (i)   for a repeated label (candidate == last prefix token) the BLANK-ending
        predecessor is dropped and the non-blank one kept (ctc.py:87-91);
  (ii)  <eos> gets no special score (ctc.py:100 is commented out);
  (iii) "log zero" is the finite sentinel -1e8 and takes part in ordinary
        float32 arithmetic (ctc.py:11);
  (iv)  all arithmetic is numpy float32 `logaddexp`, evaluated frame by frame
        in increasing t.
State layout: r[t,0] = log-prob that the prefix ends in a NON-blank at frame t,
r[t,1] = ends in blank (ctc.py:20).
"""
import numpy as np

LOGZERO = np.float32(-100000000.0)   # ctc.py:11
BLANK = 0                             # ctc.py:12


def blank_only_state(x):
    """x [T,V] float32 log-posteriors -> r [T,2] for the empty prefix (ctc.py:19-27)."""
    x = np.asarray(x, dtype=np.float32)
    T = x.shape[0]
    r = np.full((T, 2), LOGZERO, dtype=np.float32)
    acc = np.float32(0.0)
    for t in range(T):
        acc = np.float32(x[t, BLANK]) if t == 0 else np.float32(acc + x[t, BLANK])
        r[t, 1] = acc
    return r


def extend_prefix(x, prefix, r_prev, cands):
    """Score prefix+[c] for every c in cands (ctc.py:65-101).

    Returns psi [K] float32 (cumulative prefix log-prob) and r [K,T,2] float32.
    """
    x = np.asarray(x, dtype=np.float32)
    T = x.shape[0]
    cands = [int(c) for c in cands]
    K = len(cands)
    last = int(prefix[-1]) if len(prefix) > 0 else 0
    r = np.full((T, 2, K), LOGZERO, dtype=np.float32)
    first = max(1, len(prefix))
    if len(prefix) == 0:
        r[0, 0, :] = x[0, cands]
    psi = r[first - 1, 0, :]          # NB: a view, as upstream; logaddexp below rebinds it
    same = np.asarray([c == last for c in cands])
    for t in range(first, T):
        from_blank = np.where(same, LOGZERO, np.float32(r_prev[t - 1, 1])).astype(np.float32)
        from_label = np.full(K, r_prev[t - 1, 0], dtype=np.float32)
        phi = np.logaddexp(from_label, from_blank)
        emit = x[t, cands]
        r[t, 0, :] = np.logaddexp(r[t - 1, 0, :], phi) + emit
        r[t, 1, :] = np.logaddexp(r[t - 1, 1, :], r[t - 1, 0, :]) + x[t, BLANK]
        psi = np.logaddexp(psi, phi + emit)
    return psi, np.ascontiguousarray(np.moveaxis(r, 2, 0))


def extend_prefix_all(x, prefix, r_prev):
    """Same recursion over the whole vocabulary (ctc.py:29-63)."""
    V = np.asarray(x).shape[1]
    return extend_prefix(x, prefix, r_prev, list(range(V)))
