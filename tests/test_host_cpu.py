"""CPU: host-side mirrors of the reference's data / post-processing / config plumbing (no CUDA calls)."""
import importlib
import os
import subprocess
import sys

import numpy as np
import pytest
import torch
import yaml

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tools"))


def _m(name):
    return importlib.import_module("end-to-end-asr-pytorch_b200." + name)


@pytest.fixture(scope="module")
def corpora(tmp_path_factory):
    import make_synthetic_data as msd
    root = tmp_path_factory.mktemp("corpus")
    msd.make_libri(str(root / "libri"), n=20)
    msd.make_timit(str(root / "timit"), n=20)
    return root


def _solver_cfg(path, dataset, sets):
    return dict(data_path=str(path), batch_size=8, max_timestep=3000 if dataset == 'librispeech' else 0,
                max_label_len=400 if dataset == 'librispeech' else 0, use_gpu=False, n_jobs=0, dataset=dataset,
                train_set=[sets[0]], dev_set=[sets[1]], test_set=[sets[2]], dev_batch_size=4, decode_beam_size=5)


def test_libri_dataset_buckets_sorted_and_padded(corpora):
    ds = _m("dataset")
    kw = _solver_cfg(corpora / "libri", 'librispeech', ['train-clean-100', 'dev-clean', 'test-clean'])
    dl = ds.LoadDataset('train', False, **kw)
    seen = 0
    for x, y in dl:
        assert x.dim() == 4 and y.dim() == 3 and x.shape[0] == 1 and x.shape[1] == y.shape[1]
        lens = (x[0].sum(-1) != 0).sum(-1).tolist()
        assert lens == sorted(lens, reverse=True) and lens[0] == x.shape[2]       # pack_padded_sequence needs this
        assert (y[0, :, 0] == 0).all() and ((y[0] == 1).sum(-1) == 1).all()
        seen += x.shape[1]
    assert seen == 20
    test = ds.LoadDataset('test', False, **kw)
    assert all(x.shape[1] == 1 for x, _ in test)                                  # beam > 1 forces batch 1 (dataset.py:132)
    text = ds.LoadDataset('text', True, **kw)
    for y in text:
        assert y.dim() == 3
        break


def test_timit_dataset_and_mapper(corpora):
    ds, pp = _m("dataset"), _m("postprocess")
    kw = _solver_cfg(corpora / "timit", 'timit', ['train', 'dev', 'test'])
    dl = ds.LoadDataset('dev', False, **kw)
    x, y = next(iter(dl))
    assert x.shape[1] == 4 and x.shape[-1] == 39
    m = pp.Mapper(str(corpora / "timit"))
    assert m.unit == 'phone' and m.get_dim() == 40
    m2 = pp.Mapper(str(corpora / "libri"))
    assert m2.unit == 'char' and m2.translate([0, 2, 3, 1, 5], return_string=True) == 'ab'


def test_half_batch_rule():
    ds = _m("dataset")
    assert ds.HALF_BATCHSIZE_TIME == 800 and ds.HALF_BATCHSIZE_LABEL == 150


def test_cal_acc_matches_reference_loop():
    pp = _m("postprocess")
    g = torch.Generator().manual_seed(0)
    pred = torch.randn(5, 9, 7, generator=g)
    label = torch.randint(1, 7, (5, 9), generator=g)
    label[0, 4:] = 0
    label[3, 1:] = 0
    hyp = pred.argmax(-1)
    accs = []
    for p, l in zip(hyp, label):
        c = t = 0
        for pp_, ll in zip(p, l):
            if ll == 0:
                break
            c += int(pp_ == ll)
            t += 1
        accs.append(c / t)
    assert abs(pp.cal_acc(pred, label) - sum(accs) / len(accs)) < 1e-12


def test_hypothesis_addtopk_semantics():
    pp = _m("postprocess")
    emb = torch.nn.Embedding(6, 3)
    h = pp.Hypothesis('S', emb, output_seq=[torch.tensor(4)], output_scores=[torch.tensor(-1.0)], ctc_prob=0.5)
    topi = torch.tensor([[3, 1, 5]])
    topv = torch.tensor([[-0.1, -0.2, -0.7]])
    st = np.arange(3 * 4 * 2, dtype=np.float32).reshape(3, 4, 2)
    final, kids = h.addTopk(topi, topv, 'S2', att_map='A', lm_state='L', ctc_state=st, ctc_prob=np.array([1., 2., 3.]),
                            ctc_candidates=[5, 3, 1])
    assert final is h and h.outIndex == [4, 1] and float(h.output_scores[-1]) == pytest.approx(-0.2)
    assert [k.outIndex for k in kids] == [[4, 3], [4, 5]]
    assert kids[0].ctc_prob == 2.0 and np.array_equal(kids[1].ctc_state, st[0]) and kids[0].decoder_state == 'S2'
    assert float(kids[1].avgScore()) == pytest.approx((-1.0 - 0.7) / 2)
    assert h.last_char_idx.tolist() == [[1]] and h.last_char.shape == (1, 3)


def test_edit_distance_and_eval_cli(tmp_path):
    pp = _m("postprocess")
    assert pp.edit_distance("kitten", "sitting") == 3 and pp.edit_distance([], [1, 2]) == 2
    assert pp.edit_distance("", "") == 0 and pp.edit_distance("abc", "") == 3 and pp.edit_distance(["a b", "c"], ["c"]) == 1

    def dp(a, b):                                   # textbook Wagner-Fischer, the definition editdistance implements
        prev = list(range(len(b) + 1))
        for i, x in enumerate(a, 1):
            cur = [i]
            for j, y in enumerate(b, 1):
                cur.append(min(prev[j] + 1, cur[j - 1] + 1, prev[j - 1] + (x != y)))
            prev = cur
        return prev[-1]
    rng = np.random.default_rng(0)
    for _ in range(200):
        a = rng.integers(0, 4, rng.integers(0, 30)).tolist()
        b = rng.integers(0, 4, rng.integers(0, 30)).tolist()
        assert pp.edit_distance(a, b) == dp(a, b) == pp.edit_distance(b, a)
    f = tmp_path / "decode.txt"
    f.write_text("the cat sat\tthe cat sat\nhello world\thello word\n")
    out = subprocess.run([sys.executable, os.path.join(ROOT, "eval.py"), "--file", str(f)], capture_output=True, text=True)
    assert out.returncode == 0, out.stderr
    assert "CER : 0.045455" in out.stdout and "WER : 0.250000" in out.stdout


@pytest.mark.parametrize("name", ["timit_example", "libri_example", "libri_rnnlm960"])
def test_shipped_configs_match_the_reference_schema(name):
    mine = yaml.load(open(os.path.join(ROOT, "config", name + ".yaml")), Loader=yaml.FullLoader)
    ref_path = os.path.join("/root/reference/config", name + ".yaml")
    if os.path.exists(ref_path):
        ref = yaml.load(open(ref_path), Loader=yaml.FullLoader)
        ref.get('solver', {}).pop('apex', None)
        got = yaml.load(open(os.path.join(ROOT, "config", name + ".yaml")), Loader=yaml.FullLoader)
        got.get('solver', {}).pop('apex', None)
        assert got == ref
    if 'asr_model' in mine:
        assert set(mine['asr_model']) == {'optimizer', 'encoder', 'attention', 'decoder'}
        assert mine['clm']['enable'] is False


def test_src_shim_exposes_the_reference_import_paths():
    sys.path.insert(0, ROOT)
    for k in [k for k in sys.modules if k == 'src' or k.startswith('src.')]:
        del sys.modules[k]
    from src.asr import Attention, Listener, Seq2Seq, Speller  # noqa: F401
    from src.ctc import CTCPrefixScore  # noqa: F401
    from src.postprocess import Hypothesis, Mapper, cal_acc, cal_cer, draw_att  # noqa: F401
    from src.rnnlm import RNN_LM  # noqa: F401
    from src.solver import RNNLM_Trainer, Tester, Trainer  # noqa: F401
    from src.dataset import LoadDataset  # noqa: F401
    assert Seq2Seq.__module__ == "end-to-end-asr-pytorch_b200.asr"
