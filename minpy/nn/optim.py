"""First-order update rules: `rule(w, dw, config) -> (next_w, config)`.

Formulas and config keys are those of the reference's minpy/nn/optim.py:33-141 (itself after
cs231n); state lives in the per-parameter `config` dict.  Under a CUDA policy each rule is one
fused kernel per parameter tensor instead of 4-12 elementwise passes.
"""
import minpy.numpy as np
from minpy.array_variants import ArrayType

# pylint: disable=no-member


def _cuda(name):
    prim = getattr(np, name, None)
    return prim if getattr(prim, 'type', None) == ArrayType.CUDA else None


def sgd(w, dw, config=None):
    """w - learning_rate * dw."""
    if config is None:
        config = {}
    config.setdefault('learning_rate', 1e-2)
    fused = _cuda('fused_sgd')
    if fused is not None:
        return fused(w, dw, config['learning_rate']), config
    w -= config['learning_rate'] * dw
    return w, config


def sgd_momentum(w, dw, config=None):
    """v = momentum * v - learning_rate * dw; w + v.  State: `velocity`."""
    if config is None:
        config = {}
    config.setdefault('learning_rate', 1e-2)
    config.setdefault('momentum', 0.9)
    v = config.get('velocity')
    if v is None:
        v = np.zeros_like(w)
    fused = _cuda('fused_sgd_momentum')
    if fused is not None:
        next_w, v = fused(w, dw, v, config['learning_rate'], config['momentum'])
    else:
        v = v * config['momentum'] - dw * config['learning_rate']
        next_w = w + v
    config['velocity'] = v
    return next_w, config


def rmsprop(x, dx, config=None):
    """cache = decay * cache + (1 - decay) * dx^2; x - lr * dx / (sqrt(cache) + eps)."""
    if config is None:
        config = {}
    config.setdefault('learning_rate', 1e-2)
    config.setdefault('decay_rate', 0.99)
    config.setdefault('epsilon', 1e-8)
    if 'cache' not in config:
        config['cache'] = np.zeros_like(x)
    cache = config['cache']
    fused = _cuda('fused_rmsprop')
    if fused is not None:
        next_x, cache = fused(x, dx, cache, config['learning_rate'], config['decay_rate'], config['epsilon'])
    else:
        cache = cache * config['decay_rate'] + dx ** 2 * (1 - config['decay_rate'])
        next_x = x - config['learning_rate'] * dx / (np.sqrt(cache) + config['epsilon'])
    config['cache'] = cache
    return next_x, config


def adam(x, dx, config=None):
    """Adam with bias-corrected moments; epsilon is added outside the square root."""
    if config is None:
        config = {}
    config.setdefault('learning_rate', 1e-3)
    config.setdefault('beta1', 0.9)
    config.setdefault('beta2', 0.999)
    config.setdefault('epsilon', 1e-8)
    if 'm' not in config:
        config['m'] = np.zeros_like(x)
    if 'v' not in config:
        config['v'] = np.zeros_like(x)
    config.setdefault('t', 0)
    m, v, t = config['m'], config['v'], config['t'] + 1
    fused = _cuda('fused_adam')
    if fused is not None:
        next_x, m, v = fused(x, dx, m, v, config['learning_rate'], config['beta1'], config['beta2'],
                             config['epsilon'], t)
    else:
        m = config['beta1'] * m + (1 - config['beta1']) * dx
        v = config['beta2'] * v + (1 - config['beta2']) * (dx ** 2)
        m_hat = m / (1 - config['beta1'] ** t)
        v_hat = v / (1 - config['beta2'] ** t)
        next_x = x - config['learning_rate'] * m_hat / (np.sqrt(v_hat) + config['epsilon'])
    config['m'], config['v'], config['t'] = m, v, t
    return next_x, config


# ------------------------------------------------------------------ multi-tensor fast path --
_RULE_SPEC = {
    # rule name: (code, defaults in order, state keys, hyper keys -> the 4 floats of mc_multi_update)
    'sgd': (0, (('learning_rate', 1e-2),), (), ('learning_rate',)),
    'sgd_momentum': (1, (('learning_rate', 1e-2), ('momentum', 0.9)), ('velocity',), ('learning_rate', 'momentum')),
    'rmsprop': (2, (('learning_rate', 1e-2), ('decay_rate', 0.99), ('epsilon', 1e-8)), ('cache',),
                ('learning_rate', 'decay_rate', 'epsilon')),
    'adam': (3, (('learning_rate', 1e-3), ('beta1', 0.9), ('beta2', 0.999), ('epsilon', 1e-8)), ('m', 'v'),
             ('learning_rate', 'beta1', 'beta2', 'epsilon')),
}


def multi_update(rule, params, grads, configs):
    """Apply update rule `rule` (one of this module's functions) to every (param, grad, config)
    triple: same results and same config contents as calling the rule once per parameter (the
    loop of the reference's nn/solver.py:190-195), but ONE kernel launch per group of tensors that
    share hyper-parameters when everything lives on the device.  Returns (next_params, configs),
    or None when the fast path does not apply (host arrays, scalar gradients, unknown rule) --
    the caller then runs the per-parameter loop."""
    spec = _RULE_SPEC.get(getattr(rule, '__name__', None))
    if spec is None or _cuda('fused_sgd') is None or rule is not globals().get(rule.__name__):
        return None
    import ctypes
    from minpy import array as marray
    from minpy_b200 import ops
    from minpy_b200.cuda_array import CudaArray, _F32
    code, defaults, state_keys, hyper_keys = spec

    def raw(v):
        if isinstance(v, marray.Array) and v.on_device():
            v = v.get_data(ArrayType.CUDA)
        return v if isinstance(v, CudaArray) and v.dtype == _F32 else None

    ws, gs, states = [], [], []
    for w, g, cfg in zip(params, grads, configs):
        rw, rg = raw(w), raw(g)
        if rw is None or rg is None or rw.size != rg.size:
            return None
        st = []
        for key in state_keys:
            cur = cfg.get(key)
            if cur is None:
                st.append(None)
            else:
                rs = raw(cur)
                if rs is None or rs.size != rw.size:
                    return None
                st.append(rs)
        ws.append(rw)
        gs.append(rg)
        states.append(st)
    for cfg in configs:
        for key, val in defaults:
            cfg.setdefault(key, val)
        if code == 3:
            cfg.setdefault('t', 0)
    # group by identical hyper-parameters (normally one group: Solver copies one optim_config)
    groups = {}
    for i, cfg in enumerate(configs):
        sig = tuple(float(cfg[k]) for k in hyper_keys) + ((int(cfg['t']) + 1,) if code == 3 else ())
        groups.setdefault(sig, []).append(i)
    lib = ops.L()
    from minpy_b200.cuda_array import _Buffer
    capturing = _Buffer._active_pool is not None  # pylint: disable=protected-access
    outs = [None] * len(ws)
    new_states = [None] * len(ws)
    for sig, idxs in groups.items():
        k = len(idxs)
        vp = ctypes.c_void_p * k
        if capturing:
            # inside a step-graph capture the replays overwrite parameters and optimizer state IN PLACE anyway
            # (minpy_b200/step_graph.py copies new -> old at the end of the graph): let the kernel do it -- every
            # element is read and written by the same thread -- and the graph carries no copy node per tensor
            # (mlp_c3: 36 device-to-device copies per step)
            w_out = [ws[i] for i in idxs]
            s_out = [[states[i][q] if states[i][q] is not None else CudaArray.empty(ws[i].shape) for i in idxs]
                     for q in range(len(state_keys))]
        else:
            w_out = [CudaArray.empty(ws[i].shape) for i in idxs]
            s_out = [[CudaArray.empty(ws[i].shape) for i in idxs] for _ in state_keys]
        hyper = (ctypes.c_float * 4)(*(list(sig[:len(hyper_keys)]) + [0.0] * (4 - len(hyper_keys))))

        def arr(items):
            return vp(*[x.ptr if x is not None else None for x in items])
        s1 = arr([states[i][0] for i in idxs]) if len(state_keys) >= 1 else None
        s2 = arr([states[i][1] for i in idxs]) if len(state_keys) >= 2 else None
        lib.mc_multi_update(code, k, arr([ws[i] for i in idxs]), arr([gs[i] for i in idxs]), s1, s2, arr(w_out),
                            arr(s_out[0]) if len(state_keys) >= 1 else None,
                            arr(s_out[1]) if len(state_keys) >= 2 else None,
                            (ctypes.c_int64 * k)(*[ws[i].size for i in idxs]), hyper, sig[-1] if code == 3 else 0)
        for j, i in enumerate(idxs):
            outs[i] = w_out[j]
            new_states[i] = [s_out[q][j] for q in range(len(state_keys))]
    next_params = []
    for i, cfg in enumerate(configs):
        for q, key in enumerate(state_keys):
            cfg[key] = marray.wrap(new_states[i][q])
        if code == 3:
            cfg['t'] = int(cfg['t']) + 1
        next_params.append(marray.wrap(outs[i]))
    return next_params, configs
