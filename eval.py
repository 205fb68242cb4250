#!/usr/bin/env python
"""Character / word error rate of a decode result file ("truth<TAB>hypothesis" per line, the format
Tester.write_hyp produces; reference eval.py:1-22).  `editdistance` is not installed here, so the Levenshtein
distance comes from the package's postprocess module."""
import argparse
import importlib
import os
import sys

import pandas as pd

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
edit_distance = importlib.import_module("end-to-end-asr-pytorch_b200.postprocess").edit_distance

parser = argparse.ArgumentParser(description='Evaluate decoding result.')
parser.add_argument('--file', type=str, help='Path to decode result file.')


def main(argv=None):
    paras = parser.parse_args(argv)
    table = pd.read_csv(paras.file, sep='\t', header=None, keep_default_na=False, dtype=str, quoting=3)
    cer, wer = [], []
    for gt, hyp in zip(table[0].tolist(), table[1].tolist()):
        wer.append(edit_distance(hyp.split(' '), gt.split(' ')) / len(gt.split(' ')))
        cer.append(edit_distance(hyp, gt) / len(gt))
    print('CER : {:.6f}'.format(sum(cer) / len(cer)))
    print('WER : {:.6f}'.format(sum(wer) / len(wer)))
    print('p.s. for phoneme sequences, WER=Phone Error Rate and CER is meaningless.')
    return sum(cer) / len(cer), sum(wer) / len(wer)


if __name__ == '__main__':
    main()
