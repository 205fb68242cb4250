"""Write a tiny synthetic corpus in the reference's on-disk formats so that
`python3 main.py --config ...` runs end to end without a real corpus:
  TIMIT layout       <dir>/mapping.pkl, <set>_x.pkl (list of [T,D] arrays), <set>_y.pkl (list of int lists)
  LibriSpeech layout <dir>/mapping.pkl, <set>.csv (file_path,length,label) + one .npy per utterance
usage: python tools/make_synthetic_data.py timit|libri <dir> [n_utts]"""
import os
import pickle
import sys

import numpy as np


TIMIT_PHONES = ("h# b bcl d dcl g gcl p pcl t tcl k kcl dx q jh ch s sh z zh f th v dh m n ng em en eng nx l r w y hh hv el "
                "iy ih eh ey ae aa aw ay ah ao oy ow uh uw ux er ax ix axr ax-h pau epi").split()


def mapping(vocab, unit):
    """token -> index with <sos>=0, <eos>=1 (postprocess.py:11-22 auto-detects the unit from the symbols:
    '#' (TIMIT's h#) -> phone, fewer than 50 symbols -> char)."""
    m = {'<sos>': 0, '<eos>': 1}
    for i in range(2, vocab):
        m[TIMIT_PHONES[i - 2] if unit == 'phone' else chr(ord('a') + (i - 2) % 26) * (1 + (i - 2) // 26)] = i
    return m


def utterances(n, d, vocab, t_lo, t_hi, rng, l_div=(32, 12)):
    out = []
    for _ in range(n):
        T = int(rng.integers(t_lo, t_hi + 1))
        L = int(rng.integers(max(2, T // l_div[0]), max(3, T // l_div[1])))
        y = [0] + [int(v) for v in rng.integers(2, vocab, L)] + [1]
        # weakly label-dependent features so that a few training steps move the loss
        x = rng.standard_normal((T, d)).astype(np.float32)
        seg = np.linspace(0, L, T, endpoint=False).astype(int)
        x[:, :8] += np.asarray(y[1:-1])[seg][:, None] / vocab
        out.append((x, y))
    return out


def make_timit(root, n=48, d=39, vocab=40, seed=0):
    os.makedirs(root, exist_ok=True)
    rng = np.random.default_rng(seed)
    pickle.dump(mapping(vocab, 'phone'), open(os.path.join(root, 'mapping.pkl'), 'wb'))
    for s, k in (('train', n), ('dev', max(4, n // 6)), ('test', max(4, n // 6))):
        u = utterances(k, d, vocab, 300, 500, rng, l_div=(8, 6))   # >30 labels: greedy dev decoding runs 30 extra steps
        pickle.dump([x for x, _ in u], open(os.path.join(root, s + '_x.pkl'), 'wb'))
        pickle.dump([y for _, y in u], open(os.path.join(root, s + '_y.pkl'), 'wb'))


def make_libri(root, n=48, d=40, vocab=30, seed=0):
    os.makedirs(root, exist_ok=True)
    rng = np.random.default_rng(seed)
    pickle.dump(mapping(vocab, 'char'), open(os.path.join(root, 'mapping.pkl'), 'wb'))
    for s, k in (('train-clean-100', n), ('dev-clean', max(4, n // 6)), ('test-clean', max(4, n // 6))):
        os.makedirs(os.path.join(root, s), exist_ok=True)
        rows = ['file_path,length,label']
        for i, (x, y) in enumerate(utterances(k, d, vocab, 80, 200, rng)):
            rel = os.path.join(s, '%04d.npy' % i)
            np.save(os.path.join(root, rel), x)
            rows.append('%s,%d,%s' % (rel, len(x), '_'.join(str(v) for v in y)))
        open(os.path.join(root, s + '.csv'), 'w').write('\n'.join(rows) + '\n')


if __name__ == '__main__':
    kind, root = sys.argv[1], sys.argv[2]
    n = int(sys.argv[3]) if len(sys.argv) > 3 else 48
    {'timit': make_timit, 'libri': make_libri}[kind](root, n)
    print('wrote', root)
