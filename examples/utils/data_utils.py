"""Data sources of the example scripts: the CIFAR-10 python-pickle batches and the synthetic generators
the reference's examples import from `examples.utils.data_utils` (reference:
examples/utils/data_utils.py:7-79 CIFAR-10, :191-224 adding problem, :227-243 Gaussian clusters), so that
`from examples.utils.data_utils import get_CIFAR10_data` in an unmodified script resolves here.

Written against the published CIFAR-10 "python version" format (data_batch_1..5 and test_batch: a pickled
dict with 'data' -- uint8 (10000, 3072), each row R, G, B planes of 32x32 -- and 'labels'), not against the
reference's loader: no SciPy image reader is needed for it.
"""
from __future__ import absolute_import

import os
import pickle

import numpy as np


def load_CIFAR_batch(filename):
    """One batch file -> (X float64 (n, 32, 32, 3) channels-last, y int (n,))."""
    with open(filename, 'rb') as f:
        blob = pickle.load(f, encoding='latin1')
    data = np.asarray(blob['data'])
    labels = np.asarray(blob['labels' if 'labels' in blob else 'fine_labels'])
    n = data.shape[0]
    return data.reshape(n, 3, 32, 32).transpose(0, 2, 3, 1).astype('float'), labels


def load_CIFAR10(root):
    """The five training batches concatenated, and the test batch: (Xtr, ytr, Xte, yte)."""
    parts = [load_CIFAR_batch(os.path.join(root, 'data_batch_%d' % b)) for b in range(1, 6)]
    Xtr = np.concatenate([x for x, _ in parts])
    ytr = np.concatenate([y for _, y in parts])
    Xte, yte = load_CIFAR_batch(os.path.join(root, 'test_batch'))
    return Xtr, ytr, Xte, yte


def get_CIFAR10_data(data_dir, num_training=49000, num_validation=1000, num_test=1000):
    """{'X_train', 'y_train', 'X_val', 'y_val', 'X_test', 'y_test'}: the first `num_training` training
    images, the next `num_validation` as validation, the first `num_test` test images; the mean TRAINING
    image is subtracted from all three; images are returned channels-first (n, 3, 32, 32), float64."""
    Xtr, ytr, Xte, yte = load_CIFAR10(data_dir)
    X_val, y_val = Xtr[num_training:num_training + num_validation], ytr[num_training:num_training + num_validation]
    X_train, y_train = Xtr[:num_training], ytr[:num_training]
    X_test, y_test = Xte[:num_test], yte[:num_test]
    mean_image = np.mean(X_train, axis=0)
    out = {}
    for name, X, y in (('train', X_train, y_train), ('val', X_val, y_val), ('test', X_test, y_test)):
        out['X_' + name] = np.ascontiguousarray((X - mean_image).transpose(0, 3, 1, 2))
        out['y_' + name] = y
    return out


def adding_problem_generator(N, seq_len=30, high=1):
    """The adding problem (Le, Jaitly, Hinton 2015): X (N, seq_len, 2) = [values ~ U[0, high], a mask with
    exactly two ones], Y (N, 1) = the sum of the two marked values."""
    X_num = np.random.uniform(low=0, high=high, size=(N, seq_len, 1))
    X_mask = np.zeros((N, seq_len, 1))
    Y = np.ones((N, 1))
    for i in range(N):
        positions = np.random.choice(seq_len, size=2, replace=False)
        X_mask[i, positions] = 1
        Y[i, 0] = np.sum(X_num[i, positions])
    return np.append(X_num, X_mask, axis=2), Y


def gaussian_cluster_generator(num_samples=10000, num_features=500, num_classes=5):
    """`num_classes` Gaussian blobs (sigma 0.1 around uniform random centres): X (n, features), one-hot
    Y (n, classes), shuffled together."""
    mu = np.random.rand(num_classes, num_features)
    per = num_samples // num_classes
    n = per * num_classes
    x = np.zeros((n, num_features))
    y = np.zeros((n, num_classes))
    for i in range(num_classes):
        x[i * per:(i + 1) * per] = np.random.normal(mu[i], 0.1, (per, num_features))
        y[i * per:(i + 1) * per, i] = 1
    order = np.random.permutation(n)
    return x[order], y[order]
