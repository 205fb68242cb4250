"""CPU oracle for the attention branch's sequence loss.

TEST INFRASTRUCTURE ONLY -- never imported by the product package.  Only tests/,
__graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may import this file.

Restates reference src/solver.py:90,149-155
    seq_loss = CrossEntropyLoss(ignore_index=0, reduction='none')
    att_loss = seq_loss(att_pred.view(b*t,c), label.view(-1))
    att_loss = torch.sum(att_loss.view(b,t),dim=-1) / torch.sum(y!=0,dim=-1).float()
    att_loss = torch.mean(att_loss)
and the argmax + accuracy of src/postprocess.py:121-133 in float64 numpy with the log-softmax written out
(the arithmetic itself lives in the `torch` dependency; published definition: CE_n = logsumexp(x_n) - x_n[y_n],
0 where y_n == ignore_index).  Pinned by tests/test_oracle_golden.py against tests/golden/seqloss.pt, which
tools/make_golden.py produced by running the reference's own call sequence and its own cal_acc.
"""
import numpy as np


def rows(logits, labels, ignore_index=0):
    """logits [N,V], labels [N] -> (ce [N] float64, softmax [N,V] float64)."""
    x = np.asarray(logits, dtype=np.float64)
    y = np.asarray(labels).reshape(-1)
    m = x.max(axis=-1, keepdims=True)
    e = np.exp(x - m)
    s = e.sum(axis=-1, keepdims=True)
    lse = (m + np.log(s))[:, 0]
    ce = lse - x[np.arange(x.shape[0]), np.clip(y, 0, x.shape[1] - 1)]
    ce[y == ignore_index] = 0.0
    return ce, e / s


def seq_loss_mean(att_pred, label, tgt_len, ignore_index=0, upstream=1.0):
    """solver.py:149-155 -> (att_loss, rows [B,L], d(upstream*att_loss)/d att_pred [B,L,V], argmax [B,L])."""
    x = np.asarray(att_pred, dtype=np.float64)
    B, L, V = x.shape
    y = np.asarray(label).reshape(B, L)
    n = np.asarray(tgt_len, dtype=np.float64).reshape(B)
    ce, sm = rows(x.reshape(B * L, V), y.reshape(-1), ignore_index)
    ce = ce.reshape(B, L)
    loss = float(np.mean(ce.sum(axis=-1) / n))
    onehot = np.zeros((B * L, V))
    onehot[np.arange(B * L), np.clip(y.reshape(-1), 0, V - 1)] = 1.0
    g = (sm - onehot).reshape(B, L, V) * (upstream / (n * B))[:, None, None]
    g[y == ignore_index] = 0.0
    return loss, ce, g, np.argmax(x, axis=-1)


def cal_acc(argmax, label):
    """postprocess.py:121-133: per utterance, correct / counted up to the first 0 label; mean over utterances."""
    accs = []
    for p, l in zip(np.asarray(argmax), np.asarray(label)):
        correct, total = 0.0, 0
        for pp, ll in zip(p, l):
            if ll == 0:
                break
            correct += int(pp == ll)
            total += 1
        accs.append(correct / total)
    return sum(accs) / len(accs)
