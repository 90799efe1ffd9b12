"""Lowering of a `minpy_b200.symbol` graph onto libminpy_cuda.so.

Replaces what `core.Function` obtained from an MXNet executor (minpy/core.py:148-190:
`simple_bind`, `arg_arrays`, `forward(is_train)`, `outputs`, `backward(out_grads)`,
`grad_arrays`, `aux_arrays`).  A `Plan` is built once per (symbol, input shapes): it fixes the
topological order, fuses `Convolution -> Activation(relu)` into the convolution epilogue and
knows which inputs need gradients.  `Plan.forward` returns a `Run` holding the activations
the backward pass needs; `Run.backward(out_grad)` returns one gradient per argument.

Operator arithmetic follows MXNet 0.9 (the reference's un-vendored dependency; SURVEY.md 8c
restates it): NCHW cross-correlation, 'valid' max pooling with the gradient routed to the
arg-max element, relu gradient `out > 0`, FullyConnected `x.W^T + b`, SoftmaxOutput whose
backward is `(p - onehot(label)) * grad_scale [/ N if normalization == 'batch']` and ignores
the incoming gradient (the same quirk as numpy_core.py:170-175).
"""
from __future__ import annotations

import numpy as np

from . import ops
from .cuda_array import CudaArray, _F32, _I32
from .lazy import concrete

_GT0 = 'MC_VJP_GT0_A'
import os as _os
_POOL_FUSE = _os.environ.get('MINPY_CONV_POOL_FUSE', '1') != '0'


class _Pooled(object):
    """Value of a Convolution (+ fused relu) node whose 2x2 max pooling ran in the convolution's epilogue:
    the un-pooled activation does not exist; the Pooling node downstream takes `y` / `idx` as its own."""
    __slots__ = ('y', 'idx', 'like')

    def __init__(self, y, idx, like):
        self.y, self.idx, self.like = y, idx, like

# test hook: when a list, `Plan.forward` appends the (NCHW, host) output of every Convolution node
# (post-relu when the activation is fused) -- the full-size parity tests read the device's relu /
# arg-max decisions from it.  Never set on the product path.
trace = None


def _sq(pair, what):
    if pair[0] != pair[1]:
        raise NotImplementedError('%s must be square, got %s' % (what, pair))
    return int(pair[0])


def _nchw(v):
    """Dense NCHW CudaArray of a value that may be a channels-last activation."""
    return ops.to_nchw(v) if isinstance(v, ops.Nhwc) else v


def _tc_mode():
    return ops._gemm_precision['mode'] == 1  # pylint: disable=protected-access


class Run(object):
    """State of one forward pass: outputs, updated auxiliary states and the backward closure."""

    def __init__(self, plan, values, saved, new_aux, train):
        self.plan, self._values, self._saved, self.new_aux, self._train = plan, values, saved, new_aux, train
        self.outputs = [_nchw(values[id(plan.head)])]

    def backward(self, out_grad=None):
        return self.plan._backward(self, out_grad)  # pylint: disable=protected-access


class Plan(object):
    """Execution plan of one symbol for fixed input shapes."""

    def __init__(self, symbol, input_shapes):
        self.symbol = symbol
        self.head = symbol._node  # pylint: disable=protected-access
        self.order = symbol._topo()  # pylint: disable=protected-access
        self.arg_names = symbol.list_arguments()
        self.aux_names = symbol.list_auxiliary_states()
        self.input_names = set(input_shapes or ())
        consumers = {}
        for node in self.order:
            for inp in node.inputs:
                consumers[id(inp)] = consumers.get(id(inp), 0) + 1
        # Convolution whose only consumer is a relu Activation: relu runs in the conv epilogue
        self.fused_relu = {}
        for node in self.order:
            if node.op == 'Activation' and node.attrs['act_type'] == 'relu':
                src = node.inputs[0]
                if src.op == 'Convolution' and consumers.get(id(src), 0) == 1 and src is not self.head:
                    self.fused_relu[id(src)] = node
        # Pooling whose input is such a fused relu (and its only consumer): the relu mask is
        # applied by the pooling backward kernel, not by a separate pass over the conv output
        self.pool_masks_relu = {}
        for node in self.order:
            if node.op == 'Pooling':
                src = node.inputs[0]
                if src.op == 'Activation' and consumers.get(id(src), 0) == 1 and \
                        any(act is src for act in self.fused_relu.values()):
                    self.pool_masks_relu[id(node)] = src.inputs[0]
        # Convolution -> relu -> max pooling 2x2 / 2 (every link with a single consumer): the pooling runs in the
        # convolution's epilogue when the shapes allow (decided per forward call)
        self.fused_pool = {}
        for node in self.order:
            if id(node) in self.pool_masks_relu:
                a = node.attrs
                if a['pool_type'] == 'max' and tuple(a['kernel']) == (2, 2) and tuple(a['stride']) == (2, 2) and \
                        tuple(a['pad']) == (0, 0) and not a['global_pool']:
                    self.fused_pool[id(self.pool_masks_relu[id(node)])] = node
        # which nodes need a gradient: anything downstream of a non-input argument
        self.needs_grad = {}
        for node in self.order:
            if node.op == 'null':
                self.needs_grad[id(node)] = True
            else:
                self.needs_grad[id(node)] = any(self.needs_grad[id(i)] for i in node.inputs)

    # -------------------------------------------------------------------------- forward -----
    def forward(self, args, aux, train=True, data_needs_grad=None):
        """args / aux: raw arrays aligned with list_arguments() / list_auxiliary_states()
        (None for absent ones).  `data_needs_grad`: names of arguments whose gradient the
        caller will use (default: all) -- lets the first convolution skip its dgrad."""
        lib = ops.L()
        values, saved = {}, {}
        by_name = dict(zip(self.arg_names, args))
        aux_by_name = dict(zip(self.aux_names, aux))
        new_aux = dict((k, None) for k in self.aux_names)
        skip_relu = set(id(n) for n in self.fused_relu.values())
        for node in self.order:
            key = id(node)
            if node.op == 'null':
                v = by_name.get(node.name)
                if v is not None and not isinstance(v, CudaArray):
                    v = ops.as_cuda(np.asarray(v))
                if isinstance(v, CudaArray) and v.dtype != _F32 and not node.name.endswith('_label'):
                    v = v.astype(_F32)
                values[key] = v
                continue
            x = values[id(node.inputs[0])]
            if x is None:
                raise ValueError('argument "%s" of the symbol was not supplied' % node.inputs[0].name)
            a = node.attrs
            if node.op == 'Convolution':
                w = values[id(node.inputs[1])]
                b = values[id(node.inputs[2])] if not a['no_bias'] else None
                relu = key in self.fused_relu
                stride, pad = _sq(a['stride'], 'stride'), _sq(a['pad'], 'pad')
                if stride == 1 and _tc_mode() and w.shape[1] <= 256:
                    # tensor-core path: channels-last in, channels-last out (no conversion between
                    # consecutive conv / relu / pool nodes)
                    if not isinstance(x, ops.Nhwc):
                        x = ops.to_nhwc(x, pad if ops.conv_merged_ok(w.shape[1], w.shape[3]) else 0)
                    if _POOL_FUSE and trace is None and key in self.fused_pool and ops.conv_pool_fusable(x, w, pad):
                        values[key] = _Pooled(*ops.conv2d_nhwc_fwd_pool(x, w, b, pad, relu=True, want_idx=train))
                    else:
                        values[key] = ops.conv2d_nhwc_fwd(x, w, b, pad, relu=relu)
                else:
                    x = _nchw(x)
                    values[key] = ops.conv2d_fwd(x, w, b, stride, pad, relu=relu)
                saved[key] = (x, w)
                if trace is not None:       # (the pooled epilogue is off while tracing: the test wants the un-pooled output)
                    trace.append(_nchw(values[key]).asnumpy())
            elif node.op == 'Activation':
                if key in skip_relu:
                    values[key] = x                     # already applied by the producing convolution
                else:
                    op = {'relu': lib.MC_RELU, 'sigmoid': lib.MC_SIGMOID, 'tanh': lib.MC_TANH}[a['act_type']]
                    if isinstance(x, ops.Nhwc):     # elementwise: layout-agnostic (pad channels are never read as data)
                        values[key] = ops.Nhwc(ops.unary(op, x.data), x.N, x.C, x.H, x.W, x.Cp, x.wpad)
                    else:
                        values[key] = ops.unary(op, x)
            elif node.op == 'Pooling':
                if a['pad'] != (0, 0):
                    raise NotImplementedError('padded pooling is not lowered')
                k, stride = a['kernel'], _sq(a['stride'], 'stride')
                if a['global_pool']:
                    k, stride = tuple(_nchw(x).shape[2:]), 1
                if isinstance(x, _Pooled):              # already pooled by the producing convolution
                    values[key] = x.y
                    saved[key] = (x.idx, x.like)
                elif a['pool_type'] == 'avg':
                    x = _nchw(x)
                    values[key] = ops.avgpool2d_fwd(x, k, stride)
                    saved[key] = ('avg', x.shape, k, stride)
                elif isinstance(x, ops.Nhwc) and k[0] == k[1] == stride and not x.wpad:
                    y, idx = ops.maxpool2d_nhwc_fwd(x, stride, want_idx=train)
                    values[key] = y
                    saved[key] = (idx, x)
                else:
                    x = _nchw(x)
                    y, arg = ops.maxpool2d_fwd(x, k, stride, want_argmax=train)
                    values[key] = y
                    saved[key] = (arg, x.shape)
            elif node.op == 'Flatten':
                x = _nchw(x)                    # Flatten is defined on the NCHW element order
                values[key] = x.reshape((x.shape[0], x.size // max(1, x.shape[0])))
            elif node.op == 'FullyConnected':
                x = _nchw(x)
                w = values[id(node.inputs[1])]
                b = values[id(node.inputs[2])] if not a['no_bias'] else None
                x2 = x.reshape((x.shape[0], x.size // max(1, x.shape[0])))
                values[key] = ops.gemm(x2, w, trans_b=True, bias=b)
                saved[key] = (x2, w)
            elif node.op == 'SoftmaxOutput':
                values[key] = ops.softmax_rows(x)
                saved[key] = values[id(node.inputs[1])]
            elif node.op == 'BatchNorm':
                x = _nchw(x)
                spatial = None
                if x.ndim == 4:
                    # spatial batchnorm: statistics per channel over (N, H, W) = the 2-D kernels on the
                    # (N*H*W, C) channels-last matrix
                    spatial = x.shape
                    x = ops.transpose(x, (0, 2, 3, 1)).reshape((x.shape[0] * x.shape[2] * x.shape[3], x.shape[1]))
                elif x.ndim != 2:
                    raise NotImplementedError('BatchNorm expects (N, C) or (N, C, H, W) data')
                gamma, beta = values[id(node.inputs[1])], values[id(node.inputs[2])]
                if a['fix_gamma']:
                    gamma = ops.ones(gamma.shape)
                rm = aux_by_name.get(node.aux[0])
                rv = aux_by_name.get(node.aux[1])
                rm = ops.zeros((x.shape[1],)) if rm is None else ops.as_f32(rm)
                rv = ops.ones((x.shape[1],)) if rv is None else ops.as_f32(rv)
                if train and not a['use_global_stats']:
                    N, D = x.shape
                    out = CudaArray.empty((N, D))
                    nrm, nrv = CudaArray.empty((D,)), CudaArray.empty((D,))
                    sm, si = CudaArray.empty((D,)), CudaArray.empty((D,))
                    lib.mc_batchnorm_train_fwd(x.ptr, gamma.ptr, beta.ptr, rm.ptr, rv.ptr, out.ptr, nrm.ptr, nrv.ptr,
                                               sm.ptr, si.ptr, N, D, a['eps'], a['momentum'])
                    new_aux[node.aux[0]], new_aux[node.aux[1]] = nrm, nrv
                    saved[key] = (x, gamma, sm, si, spatial)
                else:
                    N, D = x.shape
                    out = CudaArray.empty((N, D))
                    lib.mc_batchnorm_test_fwd(x.ptr, gamma.ptr, beta.ptr, rm.ptr, rv.ptr, out.ptr, N, D, a['eps'])
                    saved[key] = None
                if spatial is not None:
                    n, c, h, w = spatial
                    out = ops.transpose(out.reshape((n, h, w, c)), (0, 3, 1, 2))
                values[key] = out
            elif node.op == '_Plus':
                values[key] = ops.binary(lib.MC_ADD, _nchw(x), _nchw(values[id(node.inputs[1])]))
            else:
                raise NotImplementedError('operator %s is not lowered' % node.op)
        run = Run(self, values, saved, [new_aux[k] for k in self.aux_names], train)
        run.data_needs_grad = data_needs_grad
        return run

    # ------------------------------------------------------------------------- backward -----
    def _backward(self, run, out_grad):
        lib = ops.L()
        values, saved = run._values, run._saved  # pylint: disable=protected-access
        grads = {}
        head = self.head
        if head.op != 'SoftmaxOutput':
            if out_grad is None:
                raise ValueError('backward of a symbol without a loss head needs out_grads')
            if isinstance(out_grad, (tuple, list)):
                out_grad = out_grad[0]
            grads[id(head)] = concrete(out_grad, values[id(head)].shape)
            if isinstance(values[id(head)], ops.Nhwc):
                grads[id(head)] = ops.to_nhwc(grads[id(head)])
        wanted = run.data_needs_grad
        relu_masked = set()        # convolutions whose relu mask a pooling backward already applied

        def want(node):
            if node.op != 'null':
                return self.needs_grad[id(node)]
            return wanted is None or node.name in wanted

        def add(node, g):
            """Accumulate g into node's gradient, in the layout of the node's forward value."""
            if g is None or not want(node):
                return
            like = values.get(id(node))
            if isinstance(like, _Pooled):
                like = like.like                      # layout / shape of the un-pooled activation
            if isinstance(like, ops.Nhwc) and not isinstance(g, ops.Nhwc):
                g = ops.to_nhwc(g.reshape(like.shape))
            elif isinstance(g, ops.Nhwc) and not isinstance(like, ops.Nhwc):
                g = ops.to_nchw(g)
            have = grads.get(id(node))
            if have is None:
                grads[id(node)] = g
            elif isinstance(g, ops.Nhwc):
                grads[id(node)] = ops.Nhwc(ops.binary(lib.MC_ADD, have.data, g.data), g.N, g.C, g.H, g.W, g.Cp)
            else:
                grads[id(node)] = ops.binary(lib.MC_ADD, have, g)

        def chain1(g, opname, y):
            """g * f'(y) elementwise in one kernel, on either layout."""
            step = [(lib.enums[opname], y.data if isinstance(y, ops.Nhwc) else y, 0, None, 0, 0.0, 1)]
            if isinstance(y, ops.Nhwc):
                return ops.Nhwc(ops.run_chain(g.data, 1.0, step, y.data.shape), y.N, y.C, y.H, y.W, y.Cp)
            return ops.run_chain(g, 1.0, step, y.shape)

        fused_act = dict((id(act), cid) for cid, act in self.fused_relu.items())
        for node in reversed(self.order):
            key = id(node)
            if node.op == 'null':
                continue
            a = node.attrs
            if node.op == 'SoftmaxOutput':
                p, label = values[key], saved[key]
                rows, cols = p.shape
                dx = CudaArray.empty(p.shape)
                if label is None:
                    raise ValueError('SoftmaxOutput backward needs the label argument')
                if label.ndim == 1:
                    lab = label if label.dtype == _I32 else label.astype(_I32)
                    lib.mc_softmax_ce_bwd(p.ptr, lab.ptr, None, dx.ptr, rows, cols)
                else:
                    lib.mc_softmax_ce_bwd(p.ptr, None, ops.as_f32(label).ptr, dx.ptr, rows, cols)
                # the kernel yields (p - y) / rows; MXNet's default normalization is 'null'
                scale = a['grad_scale'] * (1.0 if a['normalization'] in ('batch', 'valid') else float(rows))
                if scale != 1.0:
                    dx = ops.binary(lib.MC_MUL, dx, scale)
                add(node.inputs[0], dx)
                continue
            g = grads.pop(key, None)
            if g is None:
                continue
            x_node = node.inputs[0]
            if node.op == 'Activation':
                if key in fused_act:
                    add(x_node, g)            # the convolution applies the relu mask to its own output
                elif want(x_node):
                    op = {'relu': _GT0, 'sigmoid': 'MC_VJP_SIGMOID_ANS', 'tanh': 'MC_VJP_TANH_ANS'}[a['act_type']]
                    add(x_node, chain1(g, op, values[key]))
            elif node.op == 'Convolution':
                x, w = saved[key]
                if key in self.fused_relu and key not in relu_masked:
                    g = chain1(g, _GT0, values[key])
                need_dx = want(x_node)
                if isinstance(x, ops.Nhwc):
                    dx, dw, db = ops.conv2d_nhwc_bwd(x, w, g, _sq(a['pad'], 'pad'), need_dx=need_dx,
                                                     need_db=not a['no_bias'])
                else:
                    dx, dw, db = ops.conv2d_bwd(x, w, g, _sq(a['stride'], 'stride'), _sq(a['pad'], 'pad'),
                                                need_dx=need_dx, need_db=not a['no_bias'])
                add(x_node, dx)
                add(node.inputs[1], dw)
                if not a['no_bias']:
                    add(node.inputs[2], db)
            elif node.op == 'Pooling':
                if want(x_node) and saved[key][0] == 'avg':
                    _, xshape, k, stride = saved[key]
                    add(x_node, ops.avgpool2d_bwd(_nchw(g), xshape, k, stride))
                elif want(x_node):
                    arg, xin = saved[key]
                    if isinstance(xin, ops.Nhwc):
                        conv = self.pool_masks_relu.get(key)
                        relu_y = values[key] if conv is not None else None
                        add(x_node, ops.maxpool2d_nhwc_bwd(g, arg, xin, _sq(a['stride'], 'stride'), relu_y=relu_y))
                        if conv is not None:
                            relu_masked.add(id(conv))
                    else:
                        add(x_node, ops.maxpool2d_bwd(g, arg, xin))
            elif node.op == 'Flatten':
                add(x_node, g.reshape(values[id(x_node)].shape))      # logical NCHW shape; add() re-lays it out
            elif node.op == 'FullyConnected':
                x2, w = saved[key]
                if want(x_node):
                    add(x_node, ops.gemm(g, w).reshape(values[id(x_node)].shape))          # g . W
                add(node.inputs[1], ops.gemm(g, x2, trans_a=True))                          # g^T . x
                if not a['no_bias']:
                    add(node.inputs[2], ops.reduce(lib.MC_RED_SUM, g, 0))
            elif node.op == 'BatchNorm':
                if saved[key] is None:
                    raise NotImplementedError('backward through inference-mode BatchNorm')
                x, gamma, sm, si, spatial = saved[key]
                N, D = x.shape
                g = _nchw(g)
                if spatial is not None:
                    g = ops.transpose(g, (0, 2, 3, 1)).reshape((N, D))
                dx, dgamma, dbeta = CudaArray.empty((N, D)), CudaArray.empty((D,)), CudaArray.empty((D,))
                lib.mc_batchnorm_train_bwd(x.ptr, g.ptr, gamma.ptr, sm.ptr, si.ptr, dx.ptr, dgamma.ptr, dbeta.ptr, N, D)
                if spatial is not None:
                    n, c, h, w = spatial
                    dx = ops.transpose(dx.reshape((n, h, w, c)), (0, 3, 1, 2))
                add(x_node, dx)
                add(node.inputs[1], ops.zeros((D,)) if a['fix_gamma'] else dgamma)
                add(node.inputs[2], dbeta)
            elif node.op == '_Plus':
                add(x_node, g)
                add(node.inputs[1], g)
        by_name = dict((n.name, grads.get(id(n))) for n in self.order if n.op == 'null')
        return [by_name[name] if by_name[name] is not None else 0.0 for name in self.arg_names]


class Executor(object):
    """`symbol.simple_bind(...)`: the MXNet executor surface core.Function used (core.py:148-190)."""

    def __init__(self, symbol, shapes, grad_req='write'):
        self._symbol = symbol
        self._plan = Plan(symbol, shapes)
        arg_shapes, _, aux_shapes = symbol.infer_shape(**shapes)
        self.arg_arrays = [ops.zeros(s) for s in arg_shapes]
        self.grad_arrays = [None] * len(arg_shapes) if grad_req == 'null' else [ops.zeros(s) for s in arg_shapes]
        self.aux_arrays = [ops.zeros(s) for s in aux_shapes]
        self.arg_dict = dict(zip(symbol.list_arguments(), self.arg_arrays))
        self.aux_dict = dict(zip(symbol.list_auxiliary_states(), self.aux_arrays))
        self.outputs = []
        self._run = None

    def forward(self, is_train=False, **kwargs):
        for k, v in kwargs.items():
            ops.as_cuda(v).copyto(self.arg_dict[k])
        self._run = self._plan.forward(self.arg_arrays, self.aux_arrays, train=is_train)
        for dst, new in zip(self.aux_arrays, self._run.new_aux):
            if new is not None:
                new.copyto(dst)
        self.outputs = list(self._run.outputs)
        return self.outputs

    def backward(self, out_grads=None):
        grads = self._run.backward(out_grads)
        for i, g in enumerate(grads):
            if self.grad_arrays[i] is None:
                continue
            if isinstance(g, CudaArray):
                g.copyto(self.grad_arrays[i])
            else:
                ops.L().mc_memset(self.grad_arrays[i].ptr, 0, self.grad_arrays[i].nbytes)
