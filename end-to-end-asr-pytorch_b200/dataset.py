"""Host-side mirror of the reference's src/dataset.py: pre-bucketed datasets whose items are whole batches
(`DataLoader(batch_size=1)` yields x [1,B,T,D], y [1,B,L]).  On-disk formats are the reference's
(SURVEY.md section 8b): TIMIT `<set>_x.pkl` / `<set>_y.pkl`; LibriSpeech `<set>.csv` with columns
file_path,length,label (+ one .npy per utterance).  Plain host code; nothing here is accelerated."""
import os
import pickle

import numpy as np
import pandas as pd
import torch
from torch.nn.utils.rnn import pad_sequence
from torch.utils.data import DataLoader
from torch.utils.data.dataset import Dataset

from .preprocess import target_padding, zero_padding

HALF_BATCHSIZE_TIME = 800      # dataset.py:12
HALF_BATCHSIZE_LABEL = 150     # dataset.py:13


class TimitDataset(Dataset):
    """dataset.py:23-54: everything in memory, sorted by length (longest first), bucketed."""

    def __init__(self, file_path, sets, bucket_size, max_timestep=0, max_label_len=0):
        x, y = [], []
        for s in sets:
            with open(os.path.join(file_path, s + '_x.pkl'), 'rb') as fp:
                x += pickle.load(fp)
            with open(os.path.join(file_path, s + '_y.pkl'), 'rb') as fp:
                y += pickle.load(fp)
        assert len(x) == len(y)
        order = list(reversed(np.argsort([len(t) for t in x])))
        xs, ys = [x[i] for i in order], [y[i] for i in order]
        self.X, self.Y = [], []
        for b in range(int(np.ceil(len(xs) / bucket_size))):
            lo, hi = b * bucket_size, min((b + 1) * bucket_size, len(xs))
            self.X.append(zero_padding(xs[lo:hi], min(max_timestep, len(xs[lo]))))
            self.Y.append(target_padding(ys[lo:hi], min(max_label_len, max(len(v) for v in ys[lo:hi]))))

    def __getitem__(self, index):
        return self.X[index], self.Y[index]

    def __len__(self):
        return len(self.X)


class LibriDataset(Dataset):
    """dataset.py:56-116: csv index sorted by length, features loaded lazily, half-size buckets for long items."""

    def __init__(self, file_path, sets, bucket_size, max_timestep=0, max_label_len=0, drop=False, text_only=False):
        self.root = file_path
        tables = [pd.read_csv(os.path.join(file_path, s + '.csv')) for s in sets]
        self.table = pd.concat(tables, ignore_index=True).sort_values(by=['length'], ascending=False)
        self.text_only = text_only
        if drop and max_timestep > 0 and not text_only:
            self.table = self.table[self.table.length < max_timestep]
        if drop and max_label_len > 0:
            self.table = self.table[self.table.label.str.count('_') + 1 < max_label_len]
        X = self.table['file_path'].tolist()
        X_lens = self.table['length'].tolist()
        Y = [list(map(int, label.split('_'))) for label in self.table['label'].tolist()]
        if text_only:
            Y.sort(key=len, reverse=True)
        self.X, self.Y = [], []
        bx, bl, by = [], [], []
        for x, n, y in zip(X, X_lens, Y):
            bx.append(x); bl.append(n); by.append(y)
            if len(bx) == bucket_size:
                if bucket_size >= 2 and (max(bl) > HALF_BATCHSIZE_TIME or max(len(v) for v in by) > HALF_BATCHSIZE_LABEL):
                    h = bucket_size // 2
                    self.X += [bx[:h], bx[h:]]
                    self.Y += [by[:h], by[h:]]
                else:
                    self.X.append(bx)
                    self.Y.append(by)
                bx, bl, by = [], [], []
        if bx:
            self.X.append(bx)
            self.Y.append(by)

    def __getitem__(self, index):
        y = target_padding(list(self.Y[index]), max(len(v) for v in self.Y[index]))
        if self.text_only:
            return y
        x = [torch.FloatTensor(np.load(os.path.join(self.root, f))) for f in self.X[index]]
        return pad_sequence(x, batch_first=True), y

    def __len__(self):
        return len(self.Y)


def LoadDataset(split, text_only, data_path, batch_size, max_timestep, max_label_len, use_gpu, n_jobs,
                dataset, train_set, dev_set, test_set, dev_batch_size, decode_beam_size, **kwargs):
    """dataset.py:119-157; every yaml key under `solver:` arrives here as a keyword."""
    if split == 'train':
        bs, shuffle, sets, drop = batch_size, True, train_set, True
    elif split == 'dev':
        bs, shuffle, sets, drop = dev_batch_size, False, dev_set, True
    elif split == 'test':
        bs, shuffle, sets, drop = (1 if decode_beam_size > 1 else dev_batch_size), False, test_set, False
        n_jobs = 1
    elif split == 'text':
        bs, shuffle, sets, drop = batch_size, True, train_set, True
    else:
        raise NotImplementedError
    if dataset.upper() == 'TIMIT':
        assert not text_only, 'TIMIT does not support text only.'
        ds = TimitDataset(file_path=data_path, sets=sets, max_timestep=max_timestep, max_label_len=max_label_len,
                          bucket_size=bs)
    elif dataset.upper() == 'LIBRISPEECH':
        ds = LibriDataset(file_path=data_path, sets=sets, max_timestep=max_timestep, text_only=text_only,
                          max_label_len=max_label_len, bucket_size=bs, drop=drop)
    else:
        raise ValueError('Unsupported Dataset: ' + dataset)
    return DataLoader(ds, batch_size=1, shuffle=shuffle, drop_last=False, num_workers=n_jobs, pin_memory=use_gpu)
