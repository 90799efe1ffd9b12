"""`Solver`: the training loop (reference: minpy/nn/solver.py:12-307, after cs231n).

One `_step` = forward + tape backward over every parameter (`core.grad_and_loss`) + one update
rule call per parameter.  New here: when a data-parallel process group is initialised
(`minpy_b200.parallel`), the tape-produced gradients are averaged across ranks with one NCCL
all-reduce over a flat buffer between the backward pass and the updates -- the place the
reference leaves open between solver.py:184 and :190.
"""
from __future__ import division
from __future__ import absolute_import
from __future__ import print_function

import time

import minpy.numpy as np
from minpy import core
from minpy.nn import optim, init

# pylint: disable=no-member, too-many-instance-attributes


def _begin_loss_readback(loss):
    """Start the device->host copy of a step's loss right away, on the library's readback stream
    (CudaArray.begin_readback): reading entry i of `loss_history` later waits for step i only, so a
    training loop can print / log losses one step behind without draining the GPU."""
    raw = getattr(loss, '_data', None)
    if isinstance(raw, dict):
        for value in raw.values():
            if hasattr(value, 'begin_readback'):
                value.begin_readback()


class Solver(object):
    """Trains a `ModelBase` with mini-batch gradient descent.

    Solver(model, train_dataiter, test_dataiter, task_type='classification', init_rule='xavier',
           init_config={}, update_rule='sgd', optim_config={}, lr_decay=1.0, num_epochs=10,
           train_acc_num_samples=1000, print_every=10, verbose=True)

    After `init()` and `train()`, `model.params` holds the parameters with the best validation
    score; `loss_history`, `train_acc_history`, `val_acc_history` hold the curves.
    """

    _DEFAULTS = (('task_type', 'classification'), ('init_rule', 'xavier'), ('init_config', {}),
                 ('update_rule', 'sgd'), ('optim_config', {}), ('lr_decay', 1.0), ('num_epochs', 10),
                 ('train_acc_num_samples', 1000), ('print_every', 10), ('verbose', True))

    def __init__(self, model, train_dataiter, test_dataiter, **kwargs):
        self.model = model
        self.train_dataiter = train_dataiter
        self.test_dataiter = test_dataiter
        for key, default in self._DEFAULTS:
            value = kwargs.pop(key, default)
            setattr(self, key, dict(value) if isinstance(default, dict) else value)
        # not in the reference: replay the step as a CUDA graph (minpy_b200/step_graph.py)
        import os
        self.step_graph = bool(kwargs.pop('step_graph', os.environ.get('MINPY_STEP_GRAPH', '0') == '1'))
        self._step_graph = None
        if kwargs:
            raise ValueError('Unrecognized arguments %s' % ', '.join('"%s"' % k for k in kwargs))
        if not hasattr(optim, self.update_rule):
            raise ValueError('Invalid update_rule "%s"' % self.update_rule)
        self.update_rule = getattr(optim, self.update_rule)
        self._reset()

    def _reset(self):
        """Clear histories and give every parameter its own optimizer / initializer config."""
        self.epoch = 0
        self.best_val_acc = 0
        self.best_params = {}
        self.loss_history = []
        self.train_acc_history = []
        self.val_acc_history = []
        self._reset_data_iterators()
        self.optim_configs = dict((p, dict(self.optim_config)) for p in self.model.param_configs)
        self.init_rules = {}
        self.init_configs = {}
        for p, pconfig in self.model.param_configs.items():
            if 'init_rule' in pconfig:
                rule, config = pconfig['init_rule'], pconfig.get('init_config', {})
            else:
                rule, config = self.init_rule, dict(self.init_config)
            if not hasattr(init, rule):
                raise ValueError('Invalid init_rule "%s"' % rule)
            self.init_rules[p] = getattr(init, rule)
            self.init_configs[p] = config

    def _reset_data_iterators(self):
        if self.train_dataiter:
            self.train_dataiter.reset()
        if self.test_dataiter:
            self.test_dataiter.reset()

    def init(self):
        """Create the parameters from their init rules (and copy the auxiliary ones)."""
        for name, config in self.model.param_configs.items():
            self.model.params[name] = self.init_rules[name](config['shape'], self.init_configs[name])
        for name, value in self.model.aux_param_configs.items():
            self.model.aux_params[name] = value
        self._sync_initial_params()

    def _sync_initial_params(self):
        """Data parallel: every rank starts from rank 0's parameters."""
        from minpy_b200 import parallel
        if parallel.world_size() > 1:
            parallel.broadcast_params(self.model.params)

    def _step(self, batch):
        """One gradient update on `batch` (eager, or a CUDA-graph replay of the same launch
        sequence when the solver was created with step_graph=True / MINPY_STEP_GRAPH=1)."""
        if self.step_graph and self._step_graph is None:
            from minpy_b200.step_graph import StepGraph
            self._step_graph = StepGraph(self)
        if self._step_graph is not None:
            return self._step_graph.step(batch)
        return self._step_eager(batch)

    def _step_eager(self, batch):
        names, grads, loss = self._step_grads(batch)
        from minpy_b200 import parallel
        if parallel.world_size() > 1:
            grads = parallel.allreduce_mean(grads)
        # the reference fetches the loss here (solver.py:187), a device->host sync every step;
        # the entry stays a lazy device scalar instead: float(entry) / entry.asnumpy() fetch it
        self.loss_history.append(loss)
        _begin_loss_readback(loss)
        self._step_update(names, grads)

    def _step_grads(self, batch):
        """Forward + tape backward: (parameter names, gradients, loss)."""
        def loss_func(*_params):
            # the parameters are read from self.model.params by the model; passing them as
            # arguments is what puts them on the tape
            predict = self.model.forward_batch(batch, mode='train')
            return self.model.loss_batch(batch, predict)

        names = list(self.model.params.keys())
        values = [self.model.params[k] for k in names]
        grads, loss = core.grad_and_loss(loss_func, argnum=range(len(values)))(*values)
        if len(values) == 1:
            grads = [grads]
        return names, grads, loss

    def _step_update(self, names, grads):
        """The update rules (reference solver.py:190-195)."""
        # same arithmetic as the per-parameter loop below; when every tensor is on the device the
        # whole loop is one multi-tensor kernel launch
        fast = optim.multi_update(self.update_rule, [self.model.params[k] for k in names], grads,
                                  [self.optim_configs[k] for k in names])
        self._update_was_multi_tensor = fast is not None
        if fast is not None:
            for name, next_w, next_config in zip(names, fast[0], fast[1]):
                self.model.params[name] = next_w
                self.optim_configs[name] = next_config
            return
        for name, dw in zip(names, grads):
            next_w, next_config = self.update_rule(self.model.params[name], dw, self.optim_configs[name])
            self.model.params[name] = next_w
            self.optim_configs[name] = next_config

    def _snapshot_params(self):
        """The parameters to keep as `best_params`.  The reference stores references (solver.py:295-301):
        updates rebind parameters, they never write in place.  A step graph DOES write in place (its
        replays update the buffers the captured kernels read), so then device copies are taken."""
        sg = self._step_graph
        if sg is None or not sg.writes_in_place():
            return dict(self.model.params)
        from minpy import array as marray
        out = {}
        for name, value in self.model.params.items():
            raw = value._get_latest_data() if isinstance(value, marray.Array) else value  # pylint: disable=protected-access
            out[name] = marray.wrap(raw.copy()) if hasattr(raw, 'copy') else value
        return out

    # -- checkpoint / resume (the reference saves parameters only, model.py:141-163) --
    def save_checkpoint(self, prefix):
        """`<prefix>.params` (ModelBase.save: parameters + auxiliary states, device arrays are read
        back) and `<prefix>.solver` (.npz: every optimizer state tensor and scalar per parameter --
        velocity / cache / m, v, t, the decayed learning rate --, epoch, best score, the curves)."""
        import numpy
        self.model.save(prefix)
        blob = {'epoch': numpy.asarray(self.epoch), 'best_val_acc': numpy.asarray(float(self.best_val_acc)),
                'train_acc_history': numpy.asarray(self.train_acc_history, dtype=numpy.float64),
                'val_acc_history': numpy.asarray(self.val_acc_history, dtype=numpy.float64),
                'loss_history': numpy.asarray([float(l) for l in self.loss_history], dtype=numpy.float64)}
        for name, config in self.optim_configs.items():
            for key, value in config.items():
                host = value.asnumpy() if hasattr(value, 'asnumpy') else value
                blob['optim/%s/%s' % (name, key)] = numpy.asarray(host)
        with open('%s.solver' % prefix, 'wb') as f:
            numpy.savez(f, **blob)

    def load_checkpoint(self, prefix):
        """Inverse of `save_checkpoint` into an initialised solver (`init()` first): parameters are
        written into the existing arrays, optimizer state is re-created in the bound array namespace
        (so on the device under a CUDA policy), and a captured step graph is dropped -- its replays
        read the old buffers -- so the next step captures afresh."""
        import numpy
        self.model.load(prefix)
        with numpy.load('%s.solver' % prefix) as z:
            blob = dict((k, z[k]) for k in z.files)
        self.epoch = int(blob['epoch'])
        self.best_val_acc = float(blob['best_val_acc'])
        self.train_acc_history = [float(v) for v in blob['train_acc_history']]
        self.val_acc_history = [float(v) for v in blob['val_acc_history']]
        self.loss_history = [float(v) for v in blob['loss_history']]
        for name in self.optim_configs:
            config = dict(self.optim_config)
            lead = 'optim/%s/' % name
            for key, value in blob.items():
                if not key.startswith(lead):
                    continue
                if value.ndim == 0:
                    item = value.item()
                    config[key[len(lead):]] = int(item) if value.dtype.kind in 'iu' else float(item)
                else:
                    config[key[len(lead):]] = np.array(value)
            self.optim_configs[name] = config
        self._step_graph = None

    def check_accuracy(self, dataiter, num_samples=None):
        """Classification: fraction of correct arg-max predictions; regression: mean loss.
        With `num_samples`, a random sub-iterator of that many rows is scored instead."""
        check = dataiter
        if num_samples is not None and dataiter.num_data > num_samples:
            check = dataiter.getsubiter(num_samples)
        else:
            check.reset()
        if self.task_type == 'classification':
            correct, seen = 0, 0
            for batch in check:
                predict = self.model.forward_batch(batch, mode='test')
                correct = correct + np.sum(np.argmax(predict, axis=1) == batch.label[0])
                seen += check.batch_size      # padded rows are counted too, as in the reference (:232)
            correct = correct.asnumpy() if hasattr(correct, 'asnumpy') else correct
            # data parallel: every rank scored its own shard; all ranks must see the SAME accuracy, or
            # `val_acc > best_val_acc` picks different epochs on different ranks and the weights diverge
            from minpy_b200 import parallel
            correct = parallel.allreduce_scalar_sum(float(correct))
            seen = parallel.allreduce_scalar_sum(seen)
            return float(correct) / seen
        if self.task_type == 'regression':
            total, batches = 0, 0
            for batch in check:
                predict = self.model.forward_batch(batch, mode='test')
                total = total + self.model.loss(predict, batch.label[0])
                batches += 1
            total = total.asnumpy() if hasattr(total, 'asnumpy') else total
            from minpy_b200 import parallel
            total = parallel.allreduce_scalar_sum(float(total))
            batches = parallel.allreduce_scalar_sum(batches)
            return float(total) / batches
        raise ValueError('Task type is either classification or regression.')

    def train(self):
        """Run `num_epochs` epochs; score on (a sample of) the training set and on the test set
        after each; decay the learning rate; keep the best parameters."""
        num_iterations = self.train_dataiter.getnumiterations() * self.num_epochs
        t = 0
        for epoch in range(self.num_epochs):
            start = time.time()
            self.epoch = epoch + 1
            for batch in self.train_dataiter:
                self._step(batch)
                if self.verbose and t % self.print_every == 0:
                    print('(Iteration %d / %d) loss: %f' % (t + 1, num_iterations, self.loss_history[-1]))
                t += 1
            train_acc = self.check_accuracy(self.train_dataiter, num_samples=self.train_acc_num_samples)
            val_acc = self.check_accuracy(self.test_dataiter)
            self.train_acc_history.append(train_acc)
            self.val_acc_history.append(val_acc)
            self._reset_data_iterators()
            if self.verbose:
                if self.task_type not in ('classification', 'regression'):
                    raise ValueError('task_type is not supported.')
                what = 'acc' if self.task_type == 'classification' else 'loss'
                print('(Epoch {} / {}) train {}: {}, val {}: {}, time: {}.'.format(
                    self.epoch, self.num_epochs, what, train_acc, what, val_acc, time.time() - start))
            if val_acc > self.best_val_acc:
                self.best_val_acc = val_acc
                self.best_params = self._snapshot_params()
            for config in self.optim_configs.values():
                config['learning_rate'] *= self.lr_decay
        self.model.params = self.best_params
