"""GPU: the reference's command line end to end on a synthetic corpus in the reference's on-disk format:
`main.py --config C` trains a few hybrid CTC/attention steps through the sm_100a kernels and checkpoints,
`main.py --config C --test` loads the checkpoint and runs the joint CTC/attention beam decode, `eval.py` reads
the result file.  Run in subprocesses, exactly as a user would."""
import os
import subprocess
import sys

import pytest
import yaml

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tools"))


def _config(data, dataset):
    libri = dataset == 'librispeech'
    return {
        'asr_model': {
            'optimizer': {'type': 'Adadelta', 'learning_rate': 1.0, 'joint_ctc': 0.5},
            'encoder': {'enc_type': 'BiRNN', 'sample_rate': '2_2_1', 'sample_style': 'drop' if libri else 'concat',
                        'dim': '32_32_32', 'dropout': '0_0_0', 'rnn_cell': 'LSTM'},
            'attention': {'att_mode': 'loc' if libri else 'dot', 'dim': 24, 'proj': True, 'num_head': 1},
            'decoder': {'dim': 32, 'layer': 1, 'dropout': 0.0, 'rnn_cell': 'LSTMCell'}},
        'clm': {'enable': False},
        'solver': {
            'dataset': dataset, 'data_path': data, 'n_jobs': 0, 'max_timestep': 3000 if libri else 0,
            'max_label_len': 400 if libri else 0,
            'train_set': ['train-clean-100'] if libri else ['train'], 'batch_size': 8, 'apex': False, 'total_steps': 6,
            'tf_start': 1.0 if libri else 0.9, 'tf_end': 1.0 if libri else 0.7,
            'dev_set': ['dev-clean'] if libri else ['dev'], 'dev_batch_size': 4, 'dev_step': 3,
            'test_set': ['test-clean'] if libri else ['test'], 'max_decode_step_ratio': 0.1, 'decode_beam_size': 4,
            'decode_ctc_weight': 0.5, 'decode_lm_path': '', 'decode_lm_weight': 0.0}}


def _run(args, cwd):
    env = dict(os.environ, PYTHONPATH=ROOT)
    r = subprocess.run([sys.executable] + args, cwd=cwd, env=env, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-4000:]
    return r.stdout


@pytest.mark.parametrize("dataset", ["librispeech", "timit"])
def test_train_then_beam_decode_then_eval(cuda, tmp_path, dataset):
    import make_synthetic_data as msd
    data = str(tmp_path / "data")
    (msd.make_libri if dataset == 'librispeech' else msd.make_timit)(data, n=24)
    cfg = tmp_path / "exp.yaml"
    cfg.write_text(yaml.safe_dump(_config(data, dataset)))
    common = ["--config", str(cfg), "--ckpdir", str(tmp_path / "result"), "--logdir", str(tmp_path / "log"), "--name", "t"]
    out = _run([os.path.join(ROOT, "main.py")] + common, str(tmp_path))
    assert "Best val er" in out
    ckp = tmp_path / "result" / "t"
    assert (ckp / "asr").exists() and (ckp / "best_hyp.txt").exists()
    out = _run([os.path.join(ROOT, "main.py")] + common + ["--test"], str(tmp_path))
    assert "Decode done" in out
    best = ckp / "decode_beam_4_len_0.1_ctc0.5.txt"
    nbest = ckp / "decode_beam_4_len_0.1_ctc0.5_nbest.txt"
    assert best.exists() and nbest.exists() and (ckp / "dev_att_decode.txt").exists() and (ckp / "dev_ctc_decode.txt").exists()
    lines = best.read_text().splitlines()
    assert len(lines) == 4 and all(len(l.split('\t')) == 2 for l in lines)      # one "gt<TAB>hyp" line per test utterance
    assert len(nbest.read_text().splitlines()) >= 4
    out = _run([os.path.join(ROOT, "eval.py"), "--file", str(best)], str(tmp_path))
    assert "CER :" in out and "WER :" in out
    # --njobs 3: three utterances in flight per GPU (beam.beam_decode_many) write byte-identical result files
    one_by_one = (best.read_text(), nbest.read_text())
    best.unlink(); nbest.unlink()
    out = _run([os.path.join(ROOT, "main.py")] + common + ["--test", "--njobs", "3"], str(tmp_path))
    assert "Decode done" in out
    assert (best.read_text(), nbest.read_text()) == one_by_one
