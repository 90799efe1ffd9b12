"""Phase timing from inside a kernel (library built by tools/probe.sh with -DMC_PROBE).

    MINPY_CUDA_LIB=$PWD/tools/_probe/libminpy_cuda_probe.so python tools/probe.py head 1024 512 10
    ... python tools/probe.py bnf 1024 512        (mc_bn_dropout_fwd,       batchnorm + dropout forward)
    ... python tools/probe.py bnb 1024 512        (mc_bn_dropout_relu_bwd,  dropout + batchnorm + relu backward)

Thread 0 of every block stores %globaltimer (ns) at the kernel's MC_STAMP points; the output is, per iteration,
the spread of block start times, the time at which the last block passed each stamp, and the per-phase median /
maximum over blocks.  Stamps cost a few hundred ns in total: compare phases, not absolute times.
"""
import ctypes
import sys

import numpy as np

from minpy_b200 import ops
from minpy_b200._lib import get_lib
from minpy_b200.cuda_array import CudaArray


def dev(a):
    return CudaArray.from_numpy(np.ascontiguousarray(a))


def case_head(lib, rows=1024, K=512, C=10):
    rng = np.random.RandomState(0)
    h, W = dev(rng.randn(rows, K).astype(np.float32)), dev((rng.randn(K, C) * 0.05).astype(np.float32))
    b, y = dev(np.zeros(C, np.float32)), dev(rng.randint(0, C, rows).astype(np.int32))
    logits, dl, dh, loss = ops.empty((rows, C)), ops.empty((rows, C)), ops.empty((rows, K)), ops.empty(())
    keep = (h, W, b, y, logits, dl, dh, loss)
    return (lambda: lib.mc_head_softmax_loss(h.ptr, W.ptr, b.ptr, y.ptr, logits.ptr, dl.ptr, dh.ptr, loss.ptr, rows, K, C),
            (rows + 7) // 8, keep)


def case_bnf(lib, N=1024, D=512):
    rng = np.random.RandomState(0)
    x = dev(rng.randn(N, D).astype(np.float32))
    g, b, rm, rv = dev(np.ones(D, np.float32)), dev(np.zeros(D, np.float32)), dev(np.zeros(D, np.float32)), dev(np.ones(D, np.float32))
    out, nm, nv, sm, si = ops.empty((N, D)), ops.empty((D,)), ops.empty((D,)), ops.empty((D,)), ops.empty((D,))
    keep = (x, g, b, rm, rv, out, nm, nv, sm, si)
    return (lambda: lib.mc_bn_dropout_fwd(x.ptr, g.ptr, b.ptr, rm.ptr, rv.ptr, out.ptr, nm.ptr, nv.ptr, sm.ptr, si.ptr,
                                          N, D, 1e-5, 0.9, 0.5, 1234, 0), D // 8, keep)


def case_bnb(lib, N=1024, D=512):
    rng = np.random.RandomState(0)
    x, dy = dev(rng.randn(N, D).astype(np.float32)), dev(rng.randn(N, D).astype(np.float32))
    g, sm, si = dev(np.ones(D, np.float32)), dev(np.zeros(D, np.float32)), dev(np.ones(D, np.float32))
    dx, dg, db, dbias = ops.empty((N, D)), ops.empty((D,)), ops.empty((D,)), ops.empty((D,))
    keep = (x, dy, g, sm, si, dx, dg, db, dbias)
    return (lambda: lib.mc_bn_dropout_relu_bwd(x.ptr, dy.ptr, g.ptr, sm.ptr, si.ptr, dx.ptr, dg.ptr, db.ptr, dbias.ptr,
                                               N, D, 1, 0.5, 1234, 0, 1), D // 8, keep)


def case_gemm(lib, M=1024, N=512, K=512, ta=0, tb=0, bias=1, relu=0):
    """Stamps: 0 entry, 1 barriers + TMEM ready, 2 after the PDL wait, 3 producer issued its last TMA load,
    4 accumulator complete, 5 tile stored, 6 TMEM released."""
    rng = np.random.RandomState(0)
    A = dev(rng.randn(*((K, M) if ta else (M, K))).astype(np.float32))
    B = dev(rng.randn(*((N, K) if tb else (K, N))).astype(np.float32))
    bv = dev(rng.randn(N).astype(np.float32))
    C = ops.empty((M, N))
    keep = (A, B, bv, C)
    return (lambda: lib.mc_gemm(ta, tb, M, N, K, A.ptr, M if ta else K, B.ptr, K if tb else N, C.ptr, N,
                                bv.ptr if bias else None, relu, 0, 1), 1024, keep)


def case_wgh(lib, layer=1, B=512):
    """conv_wgrad_halo_kernel on C4 layer 0 / 1 / 2.  Stamps: 0 entry, 1 set-up done, 2 after the PDL wait,
    3 accumulators complete, 4 partials stored, 5 TMEM released."""
    C, F, HW = ((3, 32, 32), (32, 64, 16), (64, 128, 8))[layer]
    rng = np.random.RandomState(0)
    wpad = 1 if layer == 0 else 0
    Cp = (C + 3) // 4 * 4
    x = dev(rng.randn(B, HW, HW + 2 * wpad, Cp).astype(np.float32))
    dy = dev(rng.randn(B, HW, HW, F).astype(np.float32))
    dw, db = ops.empty((F, C, 3, 3)), ops.empty((F,))
    keep = (x, dy, dw, db)
    return (lambda: lib.mc_conv2d_nhwc_wgrad(x.ptr, dy.ptr, dw.ptr, db.ptr, B, C, HW, HW, Cp, wpad, F, F, 3, 3, 1), 1024, keep)


def case_lstmp(lib, T=100, N=128, H=1024):
    """lstm_persist_fwd_kernel, second-to-last time step.  Stamps: 0 step begins, 1 producers' flags seen, 2 TMA loads
    issued, 3 accumulator complete, 4 partial parked, 5 cluster barrier passed, 6 partials summed, 7 gate arithmetic done
    and stores issued, 8 block barrier, 9 flag published, 10 second cluster barrier passed."""
    rng = np.random.RandomState(0)
    A = dev(rng.randn(T, N, 4 * H).astype(np.float32))
    Wh = dev((rng.randn(H, 4 * H) / np.sqrt(H)).astype(np.float32))
    hb, cb = dev(np.zeros((T + 1, N, H), np.float32)), dev(np.zeros((T + 1, N, H), np.float32))
    gt = dev(np.zeros((T, N, 4 * H), np.float32))
    keep = (A, Wh, hb, cb, gt)
    return (lambda: lib.mc_lstm_seq_fwd(A.ptr, Wh.ptr, hb.ptr, cb.ptr, gt.ptr, T, N, H, 1), H // 8, keep)


def case_lstmb(lib, T=100, N=128, H=1024):
    """lstm_persist_bwd_kernel, second-to-last iteration.  Stamps: 0 iteration begins, 1 first producers' flags seen,
    2 TMA loads issued, 3 accumulator complete, 4 partial parked, 5 cluster barrier passed, 6 partials summed,
    7 gate backward done and dA stores issued, 8 flag published, 9 second cluster barrier passed."""
    rng = np.random.RandomState(0)
    gt = dev(rng.rand(T, N, 4 * H).astype(np.float32))
    cb = dev((rng.randn(T + 1, N, H) * 0.5).astype(np.float32))
    Wh = dev((rng.randn(H, 4 * H) / np.sqrt(H)).astype(np.float32))
    dH = dev((rng.randn(T + 1, N, H) * 0.1).astype(np.float32))
    dA = dev(np.zeros((T, N, 4 * H), np.float32))
    keep = (gt, cb, Wh, dH, dA)
    return (lambda: lib.mc_lstm_seq_bwd(gt.ptr, cb.ptr, Wh.ptr, dH.ptr, dA.ptr, T, N, H, 1), H // 8, keep)


CASES = {'lstms': case_lstmb, 'lstmb': case_lstmb, 'lstmp': case_lstmp, 'wgh': case_wgh, 'gemm': case_gemm, 'head': case_head, 'bnf': case_bnf, 'bnb': case_bnb}


def main():
    tag = sys.argv[1]
    lib = get_lib()
    call, blocks, _keep = CASES[tag](lib, *(int(v) for v in sys.argv[2:]))
    blocks = min(1024, blocks)
    read = getattr(lib.cdll, 'mc_probe_read_' + tag)
    read.argtypes = [ctypes.c_void_p]
    buf = np.zeros(1024 * 16, np.uint64)
    for it in range(6):
        call()
        lib.mc_sync()
        assert read(buf.ctypes.data) == 0
        t = buf.reshape(1024, 16)[:blocks].astype(np.int64)
        if it == 0:
            first = t.copy()
            continue
        live = (t != first).any(axis=1) if tag in ('gemm', 'wgh') else np.ones(len(t), bool)   # blocks this launch has
        t = t[live]
        used = [i for i in range(16) if t[:, i].max() > 0]
        t0 = t[:, 0].min()
        if it < 3:
            continue
        print('%s it %d: %d blocks, entry spread %d ns; last block past stamp i at (ns): %s'
              % (tag, it, len(t), t[:, 0].max() - t0, [int(t[:, i].max() - t0) for i in used]))
        full = [i for i in used if (t[:, i] >= t0).all()]
        ph = np.diff(t[:, full], axis=1)
        print('   phases between stamps %s: median %s  max %s'
              % (full, np.median(ph, axis=0).astype(int).tolist(), ph.max(axis=0).tolist()))


if __name__ == '__main__':
    main()
