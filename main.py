#!/usr/bin/env python
# coding: utf-8
"""Entry point with the reference's command line (main.py:13-25):

    python3 main.py --config config/libri_example.yaml            # train (hybrid CTC/attention)
    python3 main.py --config config/libri_example.yaml --test     # joint CTC/attention/LM beam decode
    python3 main.py --config config/libri_rnnlm960.yaml --rnnlm   # RNN-LM training (stock kernels)

Multi-GPU: launch the same command under torchrun (one process per GPU); training all-reduces gradients
over NCCL, decoding shards utterances.  `--cpu` is rejected: the hot path has no CPU fallback.
"""
import argparse
import os
import random

import numpy as np
import torch
import yaml

parser = argparse.ArgumentParser(description='Training E2E asr.')
parser.add_argument('--config', type=str, help='Path to experiment config.')
parser.add_argument('--name', default=None, type=str, help='Name for logging.')
parser.add_argument('--logdir', default='log/', type=str, help='Logging path.')
parser.add_argument('--ckpdir', default='result/', type=str, help='Checkpoint/Result path.')
parser.add_argument('--load', default=None, type=str, help='Load pre-trained model')
parser.add_argument('--seed', default=0, type=int, help='Random seed for reproducable results.')
parser.add_argument('--njobs', default=1, type=int, help='Utterances decoded concurrently per GPU (the reference: worker processes).')
parser.add_argument('--cpu', action='store_true', help='Disable GPU training (unsupported in this build).')
parser.add_argument('--test', action='store_true', help='Test the model.')
parser.add_argument('--no-msg', action='store_true', help='Hide all messages.')
parser.add_argument('--rnnlm', action='store_true', help='Option for training RNNLM.')


def main(argv=None):
    paras = parser.parse_args(argv)
    paras.gpu = not paras.cpu
    paras.verbose = not paras.no_msg
    with open(paras.config, 'r') as f:
        config = yaml.load(f, Loader=yaml.FullLoader)

    world = int(os.environ.get('WORLD_SIZE', '1'))
    if torch.cuda.is_available():
        torch.cuda.set_device(int(os.environ.get('LOCAL_RANK', '0')))
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group('nccl', device_id=torch.device('cuda', torch.cuda.current_device()))

    random.seed(paras.seed)
    np.random.seed(paras.seed)
    torch.manual_seed(paras.seed)
    if torch.cuda.is_available():
        torch.cuda.manual_seed_all(paras.seed)

    if paras.rnnlm:
        from src.solver import RNNLM_Trainer as Solver
    elif paras.test:
        from src.solver import Tester as Solver
    else:
        from src.solver import Trainer as Solver
    solver = Solver(config, paras)
    solver.load_data()
    solver.set_model()
    solver.exec()
    if world > 1:
        import torch.distributed as dist
        dist.destroy_process_group()
    return solver


if __name__ == '__main__':
    main()
