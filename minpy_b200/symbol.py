"""A small symbolic-graph façade with the MXNet 0.9 `mx.sym` surface that MinPy scripts touch.

The reference hands convolutional sub-graphs to MXNet: `examples/nn/cnn.py:32-44` builds
`mx.sym.Variable / Convolution / Activation / Pooling / Flatten` and wraps the result in
`minpy.core.Function` (minpy/core.py:98-218), which needs exactly `list_arguments`,
`list_outputs`, `list_auxiliary_states`, `infer_shape` and `simple_bind` from the symbol.
MXNet itself is an un-vendored dependency, so this module restates that surface (names,
argument order, auto-naming, shape rules) and `minpy_b200.symbol_exec` lowers the graph onto
libminpy_cuda.so.  The top-level `mxnet` package of this repository re-exports it so that
`import mxnet as mx; mx.sym.Convolution(...)` in an unmodified script resolves here.

Conventions restated from MXNet 0.9 (python/mxnet/symbol.py, src/operator/*-inl.h):
  * an operator called without `name=` is auto-named `<optype-lowercase><k>`;
  * a missing weight / bias / gamma / beta / label input becomes a Variable
    `<name>_weight`, `<name>_bias`, ...; BatchNorm owns the auxiliary states
    `<name>_moving_mean`, `<name>_moving_var`;
  * `list_arguments()` is the depth-first, input-order list of Variables; outputs are
    `<name>_output`;
  * Convolution weight is (num_filter, C, kh, kw); FullyConnected weight is
    (num_hidden, in) -- the transpose of `nn.layers.affine`'s (in, out).
"""
from __future__ import annotations

import collections

_counters = collections.defaultdict(int)


def _auto_name(op, name):
    if name is not None:
        return name
    k = _counters[op]
    _counters[op] += 1
    return '%s%d' % (op.lower(), k)


def _pair(v, default=None):
    if v is None:
        return default
    if isinstance(v, int):
        return (v, v)
    v = tuple(int(x) for x in v)
    return v if len(v) == 2 else (v[0], v[0])


class Node(object):
    """One operator application (or a Variable when op == 'null')."""
    __slots__ = ('op', 'name', 'attrs', 'inputs', 'aux')

    def __init__(self, op, name, attrs=None, inputs=(), aux=()):
        self.op = op
        self.name = name
        self.attrs = dict(attrs or {})
        self.inputs = list(inputs)      # Symbols' head nodes, in operator input order
        self.aux = list(aux)            # names of auxiliary states owned by this node


class Symbol(object):
    """Handle on the output of a node (single-output operators only)."""

    def __init__(self, node):
        self._node = node

    @property
    def name(self):
        return self._node.name

    # -- graph walks --------------------------------------------------------------------------
    def _topo(self):
        order, seen = [], set()

        def visit(node):
            if id(node) in seen:
                return
            seen.add(id(node))
            for inp in node.inputs:
                visit(inp)
            order.append(node)
        visit(self._node)
        return order

    def list_arguments(self):
        return [n.name for n in self._topo() if n.op == 'null']

    def list_outputs(self):
        return [self._node.name if self._node.op == 'null' else self._node.name + '_output']

    def list_auxiliary_states(self):
        out = []
        for n in self._topo():
            out.extend(n.aux)
        return out

    def get_internals(self):
        return [Symbol(n) for n in self._topo()]

    # -- shape inference ----------------------------------------------------------------------
    def infer_shape(self, **known):
        """(arg_shapes, out_shapes, aux_shapes) given the shapes of some arguments, in the
        order of list_arguments() / list_outputs() / list_auxiliary_states()."""
        shapes = {}
        arg_shape, aux_shape = dict((k, tuple(v)) for k, v in known.items()), {}
        for node in self._topo():
            if node.op == 'null':
                if node.name in arg_shape:
                    shapes[id(node)] = arg_shape[node.name]
                continue
            ins = node.inputs
            data = shapes.get(id(ins[0]))
            if data is None:
                raise ValueError('cannot infer shapes: input "%s" of "%s" is unknown' % (ins[0].name, node.name))
            out, params, auxs = _INFER[node.op](node, data, [shapes.get(id(i)) for i in ins])
            for inp, shp in zip(ins[1:], params):
                if inp.op == 'null' and shp is not None:
                    have = arg_shape.get(inp.name)
                    if have is not None and tuple(have) != tuple(shp):
                        raise ValueError('shape mismatch for "%s": given %s, inferred %s' % (inp.name, have, shp))
                    arg_shape[inp.name] = tuple(shp)
                    shapes[id(inp)] = tuple(shp)
            for name, shp in zip(node.aux, auxs):
                aux_shape[name] = tuple(shp)
            shapes[id(node)] = tuple(out)
        args = self.list_arguments()
        missing = [a for a in args if a not in arg_shape]
        if missing:
            raise ValueError('cannot infer shapes of arguments %s' % missing)
        return ([arg_shape[a] for a in args], [shapes[id(self._node)]],
                [aux_shape[a] for a in self.list_auxiliary_states()])

    # -- binding ------------------------------------------------------------------------------
    def simple_bind(self, ctx=None, grad_req='write', **shapes):
        """Executor with allocated argument / gradient / auxiliary arrays (core.py:148-150)."""
        from . import symbol_exec
        return symbol_exec.Executor(self, shapes, grad_req)

    # -- arithmetic between symbols (residual connections) ------------------------------------
    def __add__(self, other):
        if not isinstance(other, Symbol):
            raise TypeError('only Symbol + Symbol is supported')
        return Symbol(Node('_Plus', _auto_name('_Plus', None), {}, [self._node, other._node]))

    def __repr__(self):
        return '<Symbol %s>' % self._node.name


def _input(sym, owner, role):
    """Node of an operator input; a missing one becomes the Variable `<owner>_<role>`."""
    if sym is None:
        return Node('null', '%s_%s' % (owner, role))
    if not isinstance(sym, Symbol):
        raise TypeError('%s of %s must be a Symbol' % (role, owner))
    return sym._node  # pylint: disable=protected-access


def Variable(name, **_ignored):
    return Symbol(Node('null', name))


def Convolution(data=None, weight=None, bias=None, kernel=None, num_filter=None, stride=None, pad=None,
                no_bias=False, name=None, **_ignored):
    name = _auto_name('Convolution', name)
    if kernel is None or num_filter is None:
        raise ValueError('Convolution needs kernel= and num_filter=')
    attrs = {'kernel': _pair(kernel), 'num_filter': int(num_filter), 'stride': _pair(stride, (1, 1)),
             'pad': _pair(pad, (0, 0)), 'no_bias': bool(no_bias)}
    ins = [_input(data, name, 'data'), _input(weight, name, 'weight')]
    if not no_bias:
        ins.append(_input(bias, name, 'bias'))
    return Symbol(Node('Convolution', name, attrs, ins))


def Activation(data=None, act_type='relu', name=None, **_ignored):
    name = _auto_name('Activation', name)
    if act_type not in ('relu', 'sigmoid', 'tanh'):
        raise ValueError('unsupported act_type %r' % (act_type,))
    return Symbol(Node('Activation', name, {'act_type': act_type}, [_input(data, name, 'data')]))


def Pooling(data=None, kernel=None, pool_type='max', stride=None, pad=None, global_pool=False, name=None, **_ignored):
    name = _auto_name('Pooling', name)
    if pool_type not in ('max', 'avg'):
        raise ValueError('unsupported pool_type %r' % (pool_type,))
    attrs = {'kernel': _pair(kernel, (1, 1)), 'pool_type': pool_type, 'stride': _pair(stride, (1, 1)),
             'pad': _pair(pad, (0, 0)), 'global_pool': bool(global_pool)}
    return Symbol(Node('Pooling', name, attrs, [_input(data, name, 'data')]))


def Flatten(data=None, name=None, **_ignored):
    name = _auto_name('Flatten', name)
    return Symbol(Node('Flatten', name, {}, [_input(data, name, 'data')]))


def FullyConnected(data=None, weight=None, bias=None, num_hidden=None, no_bias=False, name=None, **_ignored):
    name = _auto_name('FullyConnected', name)
    if num_hidden is None:
        raise ValueError('FullyConnected needs num_hidden=')
    ins = [_input(data, name, 'data'), _input(weight, name, 'weight')]
    if not no_bias:
        ins.append(_input(bias, name, 'bias'))
    return Symbol(Node('FullyConnected', name, {'num_hidden': int(num_hidden), 'no_bias': bool(no_bias)}, ins))


def SoftmaxOutput(data=None, label=None, grad_scale=1.0, normalization='null', name=None, **_ignored):
    name = _auto_name('SoftmaxOutput', name)
    if normalization not in ('null', 'batch', 'valid'):
        raise ValueError('unsupported normalization %r' % (normalization,))
    attrs = {'grad_scale': float(grad_scale), 'normalization': normalization}
    return Symbol(Node('SoftmaxOutput', name, attrs, [_input(data, name, 'data'), _input(label, name, 'label')]))


def BatchNorm(data=None, gamma=None, beta=None, eps=1e-3, momentum=0.9, fix_gamma=True, use_global_stats=False,
              name=None, **_ignored):
    name = _auto_name('BatchNorm', name)
    attrs = {'eps': float(eps), 'momentum': float(momentum), 'fix_gamma': bool(fix_gamma),
             'use_global_stats': bool(use_global_stats)}
    ins = [_input(data, name, 'data'), _input(gamma, name, 'gamma'), _input(beta, name, 'beta')]
    return Symbol(Node('BatchNorm', name, attrs, ins, aux=[name + '_moving_mean', name + '_moving_var']))


# ------------------------------------------------------------------------- shape rules ------
def _conv_out(size, k, s, p):
    return (size + 2 * p - k) // s + 1


def _infer_conv(node, data, _ins):
    a = node.attrs
    if len(data) != 4:
        raise ValueError('Convolution expects NCHW input, got %s' % (data,))
    n, c, h, w = data
    kh, kw = a['kernel']
    out = (n, a['num_filter'], _conv_out(h, kh, a['stride'][0], a['pad'][0]), _conv_out(w, kw, a['stride'][1], a['pad'][1]))
    if out[2] <= 0 or out[3] <= 0:
        raise ValueError('Convolution "%s": kernel %s larger than padded input %s' % (node.name, a['kernel'], data))
    params = [(a['num_filter'], c, kh, kw)] + ([] if a['no_bias'] else [(a['num_filter'],)])
    return out, params, []


def _infer_pool(node, data, _ins):
    a = node.attrs
    n, c, h, w = data
    if a['global_pool']:
        return (n, c, 1, 1), [], []
    kh, kw = a['kernel']
    return (n, c, _conv_out(h, kh, a['stride'][0], a['pad'][0]), _conv_out(w, kw, a['stride'][1], a['pad'][1])), [], []


def _infer_fc(node, data, _ins):
    a = node.attrs
    n = data[0]
    d = 1
    for s in data[1:]:
        d *= s
    params = [(a['num_hidden'], d)] + ([] if a['no_bias'] else [(a['num_hidden'],)])
    return (n, a['num_hidden']), params, []


def _infer_flatten(_node, data, _ins):
    d = 1
    for s in data[1:]:
        d *= s
    return (data[0], d), [], []


def _infer_bn(_node, data, _ins):
    c = data[1]
    return tuple(data), [(c,), (c,)], [(c,), (c,)]


def _infer_plus(node, data, ins):
    if ins[1] is not None and tuple(ins[1]) != tuple(data):
        raise ValueError('"%s": operand shapes differ: %s vs %s' % (node.name, data, ins[1]))
    return tuple(data), [tuple(data)], []


_INFER = {
    'Convolution': _infer_conv,
    'Activation': lambda node, data, ins: (tuple(data), [], []),
    'Pooling': _infer_pool,
    'Flatten': _infer_flatten,
    'FullyConnected': _infer_fc,
    'SoftmaxOutput': lambda node, data, ins: (tuple(data), [(data[0],)], []),
    'BatchNorm': _infer_bn,
    '_Plus': _infer_plus,
}
