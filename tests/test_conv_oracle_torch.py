"""Third-party agreement for the oracle's convolution / pooling restatement.

The reference's convolution arithmetic lives in MXNet (un-vendored, minpy/__init__.py:39), so no
golden vectors can exist for it ("parity unpinned").  What CAN be done is to hold the
oracle (oracle/minpy_oracle.py conv2d_* / maxpool2d_* / relu_mx) to an implementation that shares
no line with it: torch.nn.functional.conv2d / max_pool2d on the CPU in float64, differentiated by
torch.autograd.  This is agreement with an independent implementation of the same operator
definitions, NOT parity with the reference; DESIGN.md section 2 says so.

Shapes: the list tests/test_conv_gpu.py runs the CUDA kernels on, plus strided cases.  Also pinned:
the two semantic choices the oracle asserts -- `Activation('relu')` backward is g * (out > 0) (zero
gradient AT zero, unlike layers.relu) and max-pool ties route the gradient to the FIRST maximal
element of the window in row-major order.
"""
import numpy as np
import pytest

torch = pytest.importorskip('torch')
import torch.nn.functional as F  # noqa: E402

from oracle import minpy_oracle as orc  # noqa: E402

# (N, C, H, W, F, K, pad, stride)
SHAPES = [
    (4, 3, 32, 32, 32, 3, 1, 1), (4, 32, 16, 16, 64, 3, 1, 1), (4, 64, 8, 8, 128, 3, 1, 1),
    (3, 3, 32, 32, 32, 7, 0, 1), (2, 16, 32, 32, 48, 5, 2, 1), (2, 8, 64, 64, 16, 3, 1, 1),
    (5, 40, 16, 16, 24, 3, 1, 1), (2, 4, 12, 12, 8, 3, 1, 1), (1, 3, 8, 8, 4, 3, 1, 1),
    (2, 3, 20, 20, 8, 3, 0, 1), (3, 130, 8, 8, 260, 1, 0, 1), (2, 3, 9, 9, 5, 3, 1, 1),
    (2, 6, 10, 14, 7, 3, 2, 1), (2, 5, 15, 15, 6, 3, 1, 2), (2, 4, 16, 16, 8, 1, 0, 2),
]


def _t(a, grad=True):
    return torch.tensor(a, dtype=torch.float64, requires_grad=grad)


@pytest.mark.parametrize('shape', SHAPES)
def test_conv2d_forward_and_all_gradients_agree_with_torch(shape):
    N, C, H, W, Fo, K, pad, stride = shape
    rng = np.random.RandomState(sum(shape))
    x, w, b = rng.randn(N, C, H, W), rng.randn(Fo, C, K, K) / np.sqrt(C * K * K), rng.randn(Fo)
    y = orc.conv2d_fwd(x, w, b, stride, pad)
    tx, tw, tb = _t(x), _t(w), _t(b)
    ty = F.conv2d(tx, tw, tb, stride=stride, padding=pad)
    assert y.shape == tuple(ty.shape)
    np.testing.assert_allclose(y, ty.detach().numpy(), rtol=1e-11, atol=1e-11)
    dy = rng.randn(*y.shape)
    ty.backward(_t(dy, False))
    dx, dw, db = orc.conv2d_bwd(x, w, dy, stride, pad)
    np.testing.assert_allclose(dx, tx.grad.numpy(), rtol=1e-10, atol=1e-10)
    np.testing.assert_allclose(dw, tw.grad.numpy(), rtol=1e-10, atol=1e-10)
    np.testing.assert_allclose(db, tb.grad.numpy(), rtol=1e-10, atol=1e-10)


@pytest.mark.parametrize('shape', [(6, 5, 8, 8, 2, 2), (3, 4, 13, 13, 2, 2), (2, 3, 9, 9, 3, 2), (2, 2, 6, 6, 2, 1),
                                   (4, 32, 16, 16, 2, 2), (2, 6, 12, 12, 3, 3)])
def test_maxpool_value_argmax_and_gradient_agree_with_torch(shape):
    N, C, H, W, k, s = shape
    rng = np.random.RandomState(3)
    x = rng.randn(N, C, H, W)
    y, arg = orc.maxpool2d_fwd(x, k, s)
    tx = _t(x)
    ty, tidx = F.max_pool2d(tx, k, s, return_indices=True)
    assert np.array_equal(y, ty.detach().numpy())
    assert np.array_equal(arg, tidx.numpy())                 # flat h * W + w index, bit-exact
    dy = rng.randn(*y.shape)
    ty.backward(_t(dy, False))
    assert np.array_equal(orc.maxpool2d_bwd(dy, arg, x.shape), tx.grad.numpy())


def test_maxpool_ties_route_to_the_first_maximum():
    """A relu'd tensor is full of exact ties (zeros); the oracle sends the gradient to the first
    maximal element of the window in row-major order -- the choice torch's CPU kernel makes too."""
    rng = np.random.RandomState(5)
    x = np.maximum(rng.randn(3, 4, 8, 8), 0)
    x[0, 0, :2, :2] = 7.0                                    # a window whose four entries are all equal
    y, arg = orc.maxpool2d_fwd(x, 2, 2)
    assert arg[0, 0, 0, 0] == 0                              # first element of the first window
    tx = _t(x)
    ty, tidx = F.max_pool2d(tx, 2, 2, return_indices=True)
    assert np.array_equal(arg, tidx.numpy())
    dy = rng.randn(*y.shape)
    ty.backward(_t(dy, False))
    assert np.array_equal(orc.maxpool2d_bwd(dy, arg, x.shape), tx.grad.numpy())


def test_activation_relu_backward_is_zero_at_zero():
    """MXNet `Activation('relu')` backward = g * (out > 0): no gradient AT x == 0 (layers.relu passes
    it).  torch's relu makes the same choice."""
    x = np.array([[-1.0, 0.0, 2.0], [0.0, -0.0, 3.0]])

    def f(v):
        return orc.sum(orc.relu_mx(v) * np.array([[1.0, 2.0, 3.0], [4.0, 5.0, 6.0]]))
    g, _ = orc.grad_and_loss(f)(x)
    tx = _t(x)
    (F.relu(tx) * torch.tensor([[1.0, 2.0, 3.0], [4.0, 5.0, 6.0]], dtype=torch.float64)).sum().backward()
    assert np.array_equal(np.asarray(orc.raw(g)), tx.grad.numpy())
    assert np.asarray(orc.raw(g))[0, 1] == 0.0 and np.asarray(orc.raw(g))[1, 0] == 0.0
    g2, _ = orc.grad_and_loss(lambda v: orc.sum(orc.relu(v)))(x)           # layers.relu: passes g at 0
    assert np.asarray(orc.raw(g2))[0, 1] == 1.0


def test_conv_relu_pool_trunk_gradients_agree_with_torch():
    """The C4 trunk (3 x conv3x3-relu-maxpool + flatten + affine + softmax loss) differentiated by the
    oracle's tape against torch.autograd, every parameter gradient."""
    rng = np.random.RandomState(9)
    N = 6
    x = rng.randn(N, 3, 32, 32)
    y = rng.randint(0, 10, N)
    filters = (8, 12, 16)
    ws, bs, c = [], [], 3
    for f in filters:
        ws.append(rng.randn(f, c, 3, 3) * 0.2)
        bs.append(rng.randn(f) * 0.1)
        c = f
    w1, b1 = rng.randn(filters[-1] * 16, 10) * 0.1, rng.randn(10) * 0.1

    def loss(*ps):
        h = x
        for i in range(3):
            h = orc.maxpool2d(orc.relu_mx(orc.conv2d(h, ps[2 * i], ps[2 * i + 1], 1, 1)), 2, 2)
        return orc.softmax_loss(orc.affine(orc.flatten(h), ps[6], ps[7]), y)
    params = [v for pair in zip(ws, bs) for v in pair] + [w1, b1]
    grads, val = orc.grad_and_loss(loss, argnum=range(8))(*params)

    tp = [_t(p) for p in params]
    h = _t(x, False)
    for i in range(3):
        h = F.max_pool2d(F.relu(F.conv2d(h, tp[2 * i], tp[2 * i + 1], padding=1)), 2, 2)
    tl = F.cross_entropy(h.flatten(1) @ tp[6] + tp[7], torch.tensor(y, dtype=torch.long))
    tl.backward()
    assert abs(float(val) - tl.item()) <= 1e-12 * abs(tl.item())
    for g, t in zip(grads, tp):
        np.testing.assert_allclose(np.asarray(orc.raw(g)), t.grad.numpy(), rtol=1e-9, atol=1e-12)
