"""a7 host logic without a GPU (recorder in place of the library, as tests/test_fusion_logic.py):
`with gpu(i):` selects device i for the primitives inside the scope (`mc_init(i)`), results are tagged
with that device, an operand living on another device is copied over transiently (`mc_memcpy_peer`), and
leaving the scope returns to the home device.  Reference: minpy/context.py:92-153, primitive.py:236-240."""
import numpy as np
import pytest

from test_fusion_logic import _Lib


@pytest.fixture
def fake(monkeypatch):
    import minpy
    from minpy_b200 import ops, cuda_array, _lib
    from minpy_b200.cuda_array import _Buffer, _state
    lib = _Lib(_lib.parse_enums())
    counter = [4096]

    def fake_init(self, nbytes):
        self.nbytes = (nbytes + 511) & ~511 if nbytes else 512
        self.device, self.pool = _state['device'], None
        self.ptr = counter[0]
        counter[0] += self.nbytes

    monkeypatch.setattr(_Buffer, '__init__', fake_init)
    monkeypatch.setattr(_Buffer, '__del__', lambda self: None)
    monkeypatch.setattr(ops, 'L', lambda: lib)
    monkeypatch.setattr(cuda_array, 'get_lib', lambda: lib)
    monkeypatch.setitem(_state, 'device', None)
    monkeypatch.setitem(_state, 'home', None)
    monkeypatch.delenv('LOCAL_RANK', raising=False)
    minpy.set_global_policy('only_cuda')
    yield lib
    minpy.set_global_policy('prefer_cuda')


def test_gpu_scope_places_primitives_and_migrates_operands(fake):
    import minpy.numpy as mnp
    from minpy.array_variants import ArrayType
    from minpy.context import gpu, cpu
    from minpy_b200.cuda_array import _state
    x = mnp.array(np.ones((4, 4), np.float32))
    assert fake.names()[0] == 'mc_init' and fake.calls[0][1] == (0,)          # home device: LOCAL_RANK / 0
    assert _state['home'] == 0 and x.get_data(ArrayType.CUDA).device == 0
    del fake.calls[:]
    with gpu(1):
        y = x * 2.0
        assert ('mc_init', (1,)) in fake.calls and _state['device'] == 1
        peer = [c for c in fake.calls if c[0] == 'mc_memcpy_peer']
        assert len(peer) == 1 and peer[0][1][1] == 1 and peer[0][1][3] == 0     # dst device 1 <- src device 0
        assert y.get_data(ArrayType.CUDA).device == 1
        del fake.calls[:]
        z = y + y                                                               # both operands already on device 1
        assert 'mc_memcpy_peer' not in fake.names() and 'mc_init' not in fake.names()
        with cpu(0):                                                            # "no placement request": home device
            w = x + x
            assert w.get_data(ArrayType.CUDA).device == 0
        v = z - y
        assert v.get_data(ArrayType.CUDA).device == 1
    del fake.calls[:]
    u = x + x
    assert u.get_data(ArrayType.CUDA).device == 0 and _state['device'] == 0
    assert x.get_data(ArrayType.CUDA).device == 0                               # operands never move
    # a gradient closure runs where its forward ran
    from minpy import core
    with gpu(1):
        w1 = mnp.array(np.ones((4, 4), np.float32))
        f = core.grad(lambda t: mnp.sum(t * y))
    del fake.calls[:]
    with gpu(1):
        g = f(w1)
    assert g.get_data(ArrayType.CUDA).device == 1


def test_lazy_values_are_not_evaluated_by_the_device_check(fake):
    import minpy.numpy as mnp
    from minpy.nn import layers
    from minpy.array_variants import ArrayType
    x, w, b = mnp.array(np.ones((8, 4), np.float32)), mnp.array(np.ones((4, 4), np.float32)), mnp.array(np.zeros(4, np.float32))
    del fake.calls[:]
    a = layers.affine(x, w, b)
    assert 'mc_gemm' not in fake.names()
    h = layers.relu(a)                          # argument conversion (device check included) must not force `a`
    gemms = [c for c in fake.calls if c[0] == 'mc_gemm']
    assert len(gemms) == 1 and gemms[0][1][12] == 2
    assert a.get_data(ArrayType.CUDA).pending() and h.shape == (8, 4)
