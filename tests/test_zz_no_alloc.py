"""mpcdl_rollout_static_scratch (include/mpcdl.h): the entry point that allocates nothing -- the window-8 tcgen05 GRU / LSTM
rollout with its per-CTA sequences in the library's static device array -- against mpcdl_rollout with a caller scratch.

Both are called through the C ABI (ctypes, plain device pointers); the kernel is the same, so the results must be bit-equal
(the bound the repo's kernel-vs-kernel tests use, e.g. test_two_group_mlp_kernel_bitwise_equal).  Two calls in a row on
different inputs: the array keeps the first call's contents and the second call must not see them.  Collected last
(file name), so a failure here hides no other test under `pytest -x`; it passes alone in a fresh process and after the others."""
import ctypes as C

import pytest
import torch

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _args(m, norm, blob, init, U, states, costs, cref, noise, eps, carry, precision):
    from mpcdl import _lib
    hidden, layers, window = m._desc()
    a = _lib.RolloutArgs()
    a.model = _lib.ModelDesc(m.KIND, hidden, layers, window)
    a.norm, a.init, a.U = norm.data_ptr(), init.data_ptr(), U.data_ptr()
    a.states, a.costs, a.cost_ref = states.data_ptr(), costs.data_ptr(), cref.data_ptr()
    for i in range(6):
        a.cost_qw[i] = 1.0 + 0.1 * i
    a.cost_rw[0], a.cost_rw[1] = 0.1, 0.2
    if noise is not None:
        a.noise_wblob = noise.packed(torch.device(DEV)).data_ptr()
        a.noise_history, a.noise_size, a.noise_layers, a.noise_act = noise.cfg()
        a.eps = eps.data_ptr()
    if carry is not None:
        a.stateful, a.carry = 1, carry.data_ptr()
    a.N, a.H = U.shape[0], U.shape[1]
    a.precision, a.wblob_tc = _lib.PRECISION[precision], blob.data_ptr()
    return a


@pytest.mark.parametrize("kind,precision,stateful,use_noise", [("gru", "tf32x3", False, True), ("lstm", "tf32", True, False)])
def test_static_scratch_entry_point_equals_caller_scratch(kind, precision, stateful, use_noise):
    import mpcdl
    from mpcdl import _lib, pack, synth
    lib = _lib.lib
    torch.manual_seed(7)
    layers, h = 2, 6
    m = (mpcdl.GRU_Model if kind == "gru" else mpcdl.LSTM_Model)(8, 64, 6, layers, False).eval().to(DEV)
    m._window = 8                                                # what Module.rollout derives from init [N, 8, 8]
    noise = None
    if use_noise:
        noise = mpcdl.FeedforwardNoiseModel(6, 2, 3, n_layers=2, size=32, activation="tanh", output_activation="identity").to(DEV)
    rows = synth.load_rows()
    blob = m._blob_tc(torch.device(DEV))
    norm = pack.identity_norm(8).to(DEV)
    stream = torch.cuda.current_stream().cuda_stream
    for call, n in enumerate((8192, 3000)):                      # the maximum, then fewer tiles over stale contents
        init, U = synth.make_inputs(rows, n, h, window=8, seed=11 + call)
        init, U = init.to(DEV).contiguous(), U.to(DEV).contiguous()
        cref = synth.make_cost(rows, h)[0].to(DEV).contiguous()
        eps = synth.make_eps(n, h, seed=5 + call).to(DEV).contiguous() if use_noise else None
        outs = []
        for entry in ("caller", "static"):
            states = torch.full((n, h, 6), float("nan"), device=DEV)
            costs = torch.full((n,), float("nan"), device=DEV)
            carry = torch.zeros(layers * (2 if kind == "lstm" else 1), n, 64, device=DEV) if stateful else None
            a = _args(m, norm, blob, init, U, states, costs, cref, noise, eps, carry, precision)
            if entry == "caller":
                nbytes = lib.mpcdl_rollout_scratch_bytes(C.byref(a))
                assert nbytes == min((n + 127) // 128, 148) * 8 * 64 * 128 * 4
                scratch = torch.empty(nbytes, dtype=torch.uint8, device=DEV)
                a.scratch, a.scratch_bytes = scratch.data_ptr(), nbytes
                _lib.check(lib.mpcdl_rollout(C.byref(a), stream))
            else:
                mem = torch.cuda.memory_allocated()
                _lib.check(lib.mpcdl_rollout_static_scratch(C.byref(a), stream))     # scratch fields left NULL / 0
                assert torch.cuda.memory_allocated() == mem
            torch.cuda.synchronize()
            outs.append((states, costs, carry))
        (s0, c0, h0), (s1, c1, h1) = outs
        assert torch.isfinite(s0).all() and torch.isfinite(c0).all()
        assert torch.equal(s0, s1) and torch.equal(c0, c1)
        if stateful:
            assert h0.abs().max() > 0 and torch.equal(h0, h1)


def test_static_scratch_entry_point_rejects_what_it_cannot_hold():
    import mpcdl
    from mpcdl import _lib, pack
    lib = _lib.lib
    m = mpcdl.GRU_Model(8, 64, 6, 1, False).eval().to(DEV)
    m._window = 8
    buf = torch.zeros(16, device=DEV)
    blob, norm = m._blob_tc(torch.device(DEV)), pack.identity_norm(8).to(DEV)

    def call(n, window=8, precision="tf32"):
        a = _lib.RolloutArgs()
        a.model = _lib.ModelDesc(_lib.GRU, 64, 1, window)
        a.norm, a.init, a.U, a.states = norm.data_ptr(), buf.data_ptr(), buf.data_ptr(), buf.data_ptr()
        a.wblob, a.wblob_tc = blob.data_ptr(), blob.data_ptr()
        a.N, a.H, a.precision = n, 1, _lib.PRECISION[precision]
        return lib.mpcdl_rollout_static_scratch(C.byref(a), None), lib.mpcdl_last_error().decode()

    rc, msg = call(8193)
    assert rc == -2 and "N=8193 exceeds the static scratch (8192 trajectories" in msg          # MPCDL_ERR_UNSUPPORTED
    rc, msg = call(100, window=4)
    assert rc == -2 and "window=8" in msg
    rc, msg = call(100, precision="fp32")
    assert rc == -2 and "precision=0" in msg
    assert lib.mpcdl_rollout_static_scratch(None, None) == -1
    torch.cuda.synchronize()
