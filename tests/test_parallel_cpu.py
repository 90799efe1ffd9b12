"""Data-parallel host logic on CPU: world size 2, `gloo` backend.

Each rank runs the front end on its NumPy variant (`only_numpy`), trains on ITS shard of the
global batch and exchanges gradients through `minpy_b200.parallel.allreduce_mean`, the call
`Solver._step` makes between backward and the update rules.  Checked: (1) every rank starts
from rank 0's parameters (`broadcast_params`), (2) the averaged shard gradients equal the
full-batch gradient (every loss in nn/layers.py is a mean over the batch), so two
data-parallel steps give the same parameters as two single-process steps on the whole batch,
(3) scalar reductions used by the bench (`max_over_ranks`, `allreduce_scalar_sum`).
"""
import os
import socket
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    s = socket.socket()
    s.bind(('127.0.0.1', 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _worker(rank, world, port, out_dir):
    os.environ.update(RANK=str(rank), LOCAL_RANK=str(rank), WORLD_SIZE=str(world), MASTER_ADDR='127.0.0.1',
                      MASTER_PORT=str(port))
    sys.path.insert(0, ROOT)
    import minpy
    minpy.set_global_policy('only_numpy')
    from minpy.nn.io import NDArrayIter
    from minpy.nn.solver import Solver
    from minpy_b200 import parallel, workloads
    assert parallel.init(backend='gloo') == world
    B, D, H = 8, 24, 16                         # per-rank batch
    X, Y = workloads.cifar_like(world * B * 2, seed=3)
    X = X.reshape(X.shape[0], -1)[:, :D].astype(np.float64)
    model = workloads.MLP(in_dim=D, hidden=H, classes=10)
    it = NDArrayIter(X, Y, batch_size=B, shuffle=False, num_parts=world, part_index=rank)
    solver = Solver(model, it, it, update_rule='sgd_momentum', optim_config={'learning_rate': 0.1, 'momentum': 0.9},
                    init_rule='gaussian', init_config={'stdvar': 0.1}, verbose=False)
    np.random.seed(100 + rank)                  # different draws per rank: broadcast must reconcile them
    solver.init()
    start = dict((k, np.array(v.asnumpy())) for k, v in model.params.items())
    for batch in it:
        solver._step(batch)  # pylint: disable=protected-access
    final = dict((k, np.array(v.asnumpy())) for k, v in model.params.items())
    # shuffled iterators must shard ONE permutation (rank 0's), and check_accuracy must give
    # every rank the same number (each rank scores its own shard)
    np.random.seed(1000 + rank)                 # per-rank RNG streams, as in separate processes
    ids = np.arange(world * B * 2)
    sit = NDArrayIter(ids.reshape(-1, 1).astype(np.float64), ids, batch_size=B, shuffle=True, num_parts=world,
                      part_index=rank)
    shard_ids = np.asarray(sit.label[0][1]).astype(np.int64)
    acc = solver.check_accuracy(it)
    mx = parallel.max_over_ranks(float(rank + 1))
    sm = parallel.allreduce_scalar_sum(float(rank + 1))
    np.savez(os.path.join(out_dir, 'rank%d.npz' % rank), mx=mx, sm=sm, nbatches=len(solver.loss_history),
             shard_ids=shard_ids, acc=acc,
             **dict([('start_' + k, v) for k, v in start.items()] + [('final_' + k, v) for k, v in final.items()]))
    parallel.barrier()
    parallel.shutdown()


@pytest.mark.timeout(300)
def test_two_rank_data_parallel_matches_full_batch(tmp_path):
    import torch.multiprocessing as mp
    world, port = 2, _free_port()
    mp.spawn(_worker, args=(world, port, str(tmp_path)), nprocs=world, join=True)
    r0 = np.load(str(tmp_path / 'rank0.npz'))
    r1 = np.load(str(tmp_path / 'rank1.npz'))
    names = [k[6:] for k in r0.files if k.startswith('start_')]
    assert names
    for k in names:                                                 # (1) same start, same end on every rank
        assert np.array_equal(r0['start_' + k], r1['start_' + k]), k
        assert np.allclose(r0['final_' + k], r1['final_' + k], rtol=0, atol=1e-12), k
    assert float(r0['mx']) == 2.0 and float(r0['sm']) == 3.0        # (3)
    assert int(r0['nbatches']) == 2
    both = np.concatenate([r0['shard_ids'], r1['shard_ids']])
    assert sorted(both.tolist()) == list(range(len(both)))           # the shards partition the data set
    assert not np.array_equal(np.sort(r0['shard_ids']), r0['shard_ids'])   # and it really was shuffled
    assert float(r0['acc']) == float(r1['acc'])                      # one accuracy for the whole job

    # (2) single process, whole global batch per step, from the same start
    sys.path.insert(0, ROOT)
    import minpy
    minpy.set_global_policy('only_numpy')
    import minpy.numpy as mnp
    from minpy.nn.io import NDArrayIter
    from minpy.nn.solver import Solver
    from minpy_b200 import workloads
    try:
        B, D, H = 8, 24, 16
        X, Y = workloads.cifar_like(world * B * 2, seed=3)
        X = X.reshape(X.shape[0], -1)[:, :D].astype(np.float64)
        # rank r of W owns rows r, r + W, r + 2W, ... (nn/io.py): the union of the ranks' i-th batches is
        # exactly rows [i*W*B, (i+1)*W*B) of the data set, i.e. the i-th batch of size W*B in natural order
        model = workloads.MLP(in_dim=D, hidden=H, classes=10)
        it = NDArrayIter(X, Y, batch_size=world * B, shuffle=False)
        solver = Solver(model, it, it, update_rule='sgd_momentum', optim_config={'learning_rate': 0.1, 'momentum': 0.9},
                        init_rule='gaussian', init_config={'stdvar': 0.1}, verbose=False)
        solver.init()
        for k in names:
            model.params[k] = mnp.array(r0['start_' + k])
        for batch in it:
            solver._step(batch)  # pylint: disable=protected-access
        for k in names:
            assert np.allclose(model.params[k].asnumpy(), r0['final_' + k], rtol=1e-10, atol=1e-12), k
    finally:
        minpy.set_global_policy('prefer_cuda')


def test_peer_memory_window_gating(monkeypatch):
    """_init_p2p_window: '0' never uses the window, and outside 2..8 ranks it is never used -- decided before
    the library is touched ('auto' uses it at every world size 2..8 since the push kernel, profiles/r2n4)."""
    from minpy_b200 import parallel

    class Boom(object):
        def __getattr__(self, name):
            raise AssertionError('the library must not be called: %s' % name)

    import minpy_b200.ops as ops
    monkeypatch.setattr(ops, 'L', lambda: Boom())
    monkeypatch.setattr(parallel, '_dist', lambda: Boom())
    for world, mode in ((4, '0'), (9, 'auto'), (2, '0'), (9, '1'), (1, '1'), (1, 'auto')):
        if mode == 'auto':
            monkeypatch.delenv('MINPY_P2P_ALLREDUCE', raising=False)
        else:
            monkeypatch.setenv('MINPY_P2P_ALLREDUCE', mode)
        parallel._state['p2p'] = 'stale'                         # pylint: disable=protected-access
        parallel._init_p2p_window(0, world)                      # pylint: disable=protected-access
        assert parallel._state['p2p'] is None and not parallel.p2p_active()   # pylint: disable=protected-access
