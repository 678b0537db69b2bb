"""CPU: host-side logic of the drop-in boundary -- dataset/statistics mirror, weight packers,
checkpoint-name schema, state_dict compatibility, error behaviour without a device."""
import numpy as np
import pytest
import torch

from conftest import FWD, RNN, load_ckpt
from oracle import rollout_ref as R


def test_robotdata_matches_oracle_bitwise(bags, gold_stats):
    import mpcdl
    ds = mpcdl.RobotData(bags)
    f, l = R.mlp_dataset(bags)
    fn, ln = R.apply_zscore(f, l, R.zscore_stats(f, l))
    assert torch.equal(ds.featrs, fn) and torch.equal(ds.labels, ln) and len(ds) == 4550
    for k in ("f_mean", "f_std", "l_mean", "l_std"):
        assert torch.equal(ds.stats[k], gold_stats[0][k])
    dr = mpcdl.RobotData(bags, 4)
    f, l = R.rnn_dataset(bags, 4)
    fn, ln = R.apply_zscore(f, l, R.zscore_stats(f, l))
    assert torch.equal(dr.featrs, fn) and torch.equal(dr.labels, ln) and len(dr) == 4546
    assert tuple(dr.stats["f_mean"].shape) == (4, 8)
    x, y = dr[3]
    assert x.shape == (4, 8) and y.shape == (6,)


def test_robotdata_ragged_and_short_bags():
    import mpcdl
    rows = mpcdl.synth.load_rows()[:40].tolist()
    bags = [rows[:10], rows[10:13], rows[13:14], rows[14:40]]       # a 1-row bag contributes nothing
    ds = mpcdl.RobotData(bags)
    f, l = R.mlp_dataset(bags)
    assert len(ds) == len(f) == 9 + 2 + 0 + 25
    dr = mpcdl.RobotData(bags, 4)
    f, l = R.rnn_dataset(bags, 4)                                   # bags shorter than size+2 rows are skipped
    assert len(dr) == len(f) == 5 + 0 + 0 + 21
    fn, ln = R.apply_zscore(f, l, R.zscore_stats(f, l))
    assert torch.equal(dr.featrs, fn) and torch.equal(dr.labels, ln)


def _mlp_from_blob(blob, hidden, layers, x):
    """numpy interpreter of the packed MLP blob: per layer [bias(width)] [k][width]."""
    o, widths = 0, [(8, hidden)] + [(hidden, hidden)] * layers + [(hidden, 8)]
    for i, (k, w) in enumerate(widths):
        b = blob[o:o + w]; o += w
        W = blob[o:o + k * w].reshape(k, w); o += k * w
        x = x @ W + b
        if i < len(widths) - 1:
            x = np.maximum(x, 0)
    assert o == blob.size
    return x[:, :6]


@pytest.mark.parametrize("name", FWD)
def test_pack_mlp_roundtrip(name, bags):
    from mpcdl import pack
    sd = load_ckpt(name)
    h, L = int(name.split("_")[1]), int(name.split("_")[2])
    blob = pack.pack_mlp(sd, h, L).numpy().astype(np.float64)
    m, _ = R.build_from_name(name, sd)
    x = torch.randn(64, 8, generator=torch.Generator().manual_seed(0))
    with torch.no_grad():
        ref = m(x).numpy()
    got = _mlp_from_blob(blob, h, L, x.numpy().astype(np.float64))
    assert np.abs(got - ref).max() / np.abs(ref).max() < 1e-5


def test_pack_mlp_skips_dropout_entries():
    import mpcdl
    from mpcdl import pack
    torch.manual_seed(0)
    m = mpcdl.Basic_Model(8, 16, 6, 2, True)                       # Linear indices 0,1,3,5 (basic_model.py:53-54)
    keys = sorted({int(k.split(".")[1]) for k in m.state_dict()})
    assert keys == [0, 1, 3, 5]
    blob = pack.pack_mlp(m.state_dict(), 16, 2).numpy().astype(np.float64)
    ref = R.RefBasic(8, 16, 6, 2, True)
    ref.load_state_dict(m.state_dict())
    x = torch.randn(32, 8)
    with torch.no_grad():
        want = ref.eval()(x).numpy()
    got = _mlp_from_blob(blob, 16, 2, x.numpy().astype(np.float64))
    assert np.abs(got - want).max() < 1e-5


def _sig(x):
    return 1 / (1 + np.exp(-x))


def _rnn_from_blob(blob, kind, h, L, X, RJ=4):
    """numpy interpreter of the packed GRU/LSTM blob (gate-interleaved columns [jg][G][RJ],
    bias block [jg][4][RJ]); X [B,T,8] -> [B,6], zero initial state."""
    G = 3 if kind == "GRU" else 4
    njg = h // RJ
    o, seq = 0, [X[:, t] for t in range(X.shape[1])]

    def cols(M):      # [..., jg, G, RJ] -> per gate [..., h]
        M = M.reshape(M.shape[:-1] + (njg, G, RJ))
        return [M[..., g, :].reshape(M.shape[:-3] + (h,)) for g in range(G)]

    for l in range(L):
        k = 8 if l == 0 else h
        bias = blob[o:o + 4 * h].reshape(njg, 4, RJ); o += 4 * h
        b = [bias[:, g, :].reshape(h) for g in range(4)]
        Wx = cols(blob[o:o + k * G * h].reshape(k, G * h)); o += k * G * h
        Wh = cols(blob[o:o + h * G * h].reshape(h, G * h)); o += h * G * h
        hs, c, out = np.zeros((X.shape[0], h)), np.zeros((X.shape[0], h)), []
        for x in seq:
            if kind == "GRU":      # bias blocks: r (ih+hh), z (ih+hh), n_x (ih), n_h (hh)
                r = _sig(x @ Wx[0] + hs @ Wh[0] + b[0]); z = _sig(x @ Wx[1] + hs @ Wh[1] + b[1])
                n = np.tanh(x @ Wx[2] + b[2] + r * (hs @ Wh[2] + b[3]))
                hs = (1 - z) * n + z * hs
            else:
                i, f, g, og = (x @ Wx[q] + hs @ Wh[q] + b[q] for q in range(4))
                c = _sig(f) * c + _sig(i) * np.tanh(g)
                hs = _sig(og) * np.tanh(c)
            out.append(hs)
        seq = out
    bfc = blob[o:o + 8]; o += 8
    Wfc = blob[o:o + h * 8].reshape(h, 8); o += h * 8
    assert o == blob.size
    return (seq[-1] @ Wfc + bfc)[:, :6]


@pytest.mark.parametrize("name", [RNN[0], RNN[4]])
def test_pack_rnn_roundtrip(name, bags):
    from mpcdl import _lib, pack
    sd = load_ckpt(name)
    kind = name.split("_")[0]
    blob = pack.pack_rnn(sd, _lib.GRU if kind == "GRU" else _lib.LSTM, 64, 4, 4).numpy().astype(np.float64)
    m, _ = R.build_from_name(name, sd)
    X = torch.randn(32, 4, 8, generator=torch.Generator().manual_seed(1))
    with torch.no_grad():
        ref = m(X, m.zero_state(32))[0].numpy()
    got = _rnn_from_blob(blob, kind, 64, 4, X.numpy().astype(np.float64))
    assert np.abs(got - ref).max() / np.abs(ref).max() < 1e-5


def test_pack_noise_layout():
    from mpcdl import pack
    sd = load_ckpt("noise_seed0_h3_s32_l2")
    blob = pack.pack_noise(sd, 3, 32, 2).numpy()
    assert blob.size == (24 + 1) * 32 + (32 + 1) * 32 + (32 + 1) * 12
    W0 = blob[:24 * 32].reshape(24, 32)
    assert np.array_equal(W0, sd["0.weight"].numpy().T) and np.array_equal(blob[24 * 32:25 * 32], sd["0.bias"].numpy())


@pytest.mark.parametrize("name", [FWD[-1], RNN[0], RNN[5]])
def test_dropin_modules_load_reference_state_dicts(name):
    import mpcdl
    sd = load_ckpt(name)
    m, meta = mpcdl.from_checkpoint_name(name + ".pt", sd)
    ref, rmeta = R.build_from_name(name, sd)
    assert list(m.state_dict().keys()) == list(ref.state_dict().keys()) == list(sd.keys())
    assert meta["hidden"] == 64 and meta["layers"] == 4
    if meta["kind"] != "mlp":
        assert (m.hidden, m.layers, m.state_f) == (64, 4, name.endswith("True"))
        m.init_hidden(5)
        hs = m.h_state if meta["kind"] == "gru" else m.h_state[0]
        assert tuple(hs.shape) == (4, 5, 64)


def test_no_cpu_fallback():
    """CPU tensors must fail loudly: the op is registered for the CUDA dispatch key only."""
    import mpcdl
    m, _ = mpcdl.from_checkpoint_name("Basic_16_1_32", load_ckpt("Basic_16_1_32"))
    with pytest.raises((RuntimeError, NotImplementedError)):
        m(torch.zeros(4, 8))
    with pytest.raises((RuntimeError, NotImplementedError)):
        m.rollout(torch.zeros(4, 6), torch.zeros(4, 3, 2))
    with pytest.raises(RuntimeError):
        mpcdl.Basic_Model(7, 16, 6, 1, False)
    with pytest.raises(RuntimeError):
        mpcdl.pack.pack_mlp(load_ckpt("Basic_16_1_32"), 64, 1)


def test_synth_inputs_are_deterministic_and_bounded():
    from mpcdl import synth
    rows = synth.load_rows()
    a = synth.make_inputs(rows, 100, 30, window=4, seed=9)
    b = synth.make_inputs(rows, 100, 30, window=4, seed=9)
    assert torch.equal(a[0], b[0]) and torch.equal(a[1], b[1])
    assert a[0].shape == (100, 4, 8) and a[1].shape == (100, 30, 2)
    u = torch.as_tensor(rows[:, 6:8], dtype=torch.float32)
    assert (a[1] >= u.amin(0)).all() and (a[1] <= u.amax(0)).all()


def test_pack_rnn_tensor_core_layout():
    """mpcdl_pack_rnn_tc against a numpy restatement: tf32 hi/lo split (round to nearest, ties away),
    canonical K-major UMMA core-matrix order, GRU weight_ih rows packed [n; r; z]."""
    import numpy as np
    import torch
    from conftest import load_ckpt
    from mpcdl import _lib, pack

    def rna(x):
        u = x.astype(np.float32).view(np.uint32)
        return ((u + np.uint32(0x1000)) & np.uint32(0xFFFFE000)).view(np.float32)

    for name, kind, G in (("GRU_64_4_4_16_False", _lib.GRU, 3), ("LSTM_64_4_4_16_False", _lib.LSTM, 4)):
        sd = load_ckpt(name)
        blob = pack.pack_rnn(sd, kind, 64, 4, 4, tensor_core=True).numpy()
        pre = "gru" if kind == _lib.GRU else "lstm"
        L, h, N = 4, 64, G * 64
        off = L * h * 4
        b_ih, b_hh = sd[f"{pre}.bias_ih_l1"].numpy(), sd[f"{pre}.bias_hh_l1"].numpy()
        bias = blob[h * 4:2 * h * 4].reshape(h, 4)                        # layer 1
        if kind == _lib.GRU:
            want = np.stack([b_ih[:h] + b_hh[:h], b_ih[h:2 * h] + b_hh[h:2 * h], b_ih[2 * h:], b_hh[2 * h:]], 1)
        else:
            want = (b_ih + b_hh).reshape(4, h).T
        assert np.array_equal(bias, want)
        fc = blob[off:off + h * 8 + 8]
        assert np.array_equal(fc[:h * 8].reshape(h, 8)[:, :6], sd["full_c.weight"].numpy().T)
        assert np.array_equal(fc[h * 8:h * 8 + 6], sd["full_c.bias"].numpy())
        pos = (off + h * 8 + 8 + 31) // 32 * 32
        for l in range(L):
            for part, key in enumerate(("weight_ih", "weight_hh")):
                W = sd[f"{pre}.{key}_l{l}"].numpy()
                K = W.shape[1]
                if kind == _lib.GRU and part == 0:
                    W = np.concatenate([W[2 * h:], W[:h], W[h:2 * h]])       # [n; r; z]
                kc = min(K, 16)
                hi = rna(W)
                lo = rna(W - hi)
                for c in range(K // kc):
                    chunk = blob[pos:pos + 2 * kc * N]
                    for half, ref in ((0, hi), (1, lo)):
                        blk = chunk[half * kc * N:(half + 1) * kc * N].reshape(kc // 4, N // 8, 8, 4)   # [k/4][n/8][n%8][k%4]
                        got = blk.transpose(1, 2, 0, 3).reshape(N, kc)
                        assert np.array_equal(got, ref[:, c * kc:(c + 1) * kc]), (name, l, part, c, half)
                    pos += 2 * kc * N
        assert pos == blob.size


def test_pack_mlp_tensor_core_layout():
    """mpcdl_pack_mlp_tc against a numpy restatement: biases [L+1][64], FP32 output head W^T [64][8] + bias[8],
    then per layer the tf32 hi block and lo block in canonical K-major UMMA order [k/4][n/8][n%8][k%4]."""
    import numpy as np
    from conftest import load_ckpt
    from mpcdl import pack

    def rna(x):
        u = x.astype(np.float32).view(np.uint32)
        return ((u + np.uint32(0x1000)) & np.uint32(0xFFFFE000)).view(np.float32)

    for name, L in (("Basic_64_1_32", 1), ("Basic_64_2_64", 2), ("Basic_64_4_32", 4)):
        sd = load_ckpt(name)
        blob = pack.pack_mlp(sd, 64, L, tensor_core=True).numpy()
        h = 64
        for l in range(L + 1):
            assert np.array_equal(blob[l * h:(l + 1) * h], sd[f"layers.{l}.bias"].numpy())
        off = (L + 1) * h
        assert np.array_equal(blob[off:off + h * 8].reshape(h, 8)[:, :6], sd[f"layers.{L + 1}.weight"].numpy().T)
        assert np.array_equal(blob[off + h * 8:off + h * 8 + 6], sd[f"layers.{L + 1}.bias"].numpy())
        pos = (off + h * 8 + 8 + 31) // 32 * 32
        for l in range(L + 1):
            W = sd[f"layers.{l}.weight"].numpy()
            K = W.shape[1]
            hi = rna(W)
            lo = rna(W - hi)
            for ref in (hi, lo):
                blk = blob[pos:pos + K * h].reshape(K // 4, h // 8, 8, 4)
                assert np.array_equal(blk.transpose(1, 2, 0, 3).reshape(h, K), ref), (name, l)
                pos += K * h
        assert pos == blob.size
    with pytest.raises(RuntimeError):                # hidden 16: FP32 kernels only
        pack.pack_mlp(load_ckpt("Basic_16_1_32"), 16, 1, tensor_core=True)


def test_pack_lstm256_bf16_layout():
    """mpcdl_pack_rnn_tc for the hidden-256 LSTM (rollout_lstm_tc.cu, BASELINE config 4) against a numpy restatement:
    BF16 round-to-nearest-even, per layer / 64-unit chunk the x slots then the h slots (64 k each), every slot
    [CTA rank][k/8][n/8][8 rows][8 k] -- each CTA of the pair streams the canonical K-major UMMA block of ITS 128 of the
    chunk's 256 gate rows -- with row n = gate n/64, unit 64*chunk + n%64; layer 0 holds [W | W] (K = 16) against the
    [x_hi | x_lo] operand."""
    import numpy as np
    import torch
    from mpcdl import _lib, pack

    def bf16(x):
        u = np.ascontiguousarray(x, dtype=np.float32).view(np.uint32)
        return ((u + np.uint32(0x7FFF) + ((u >> np.uint32(16)) & np.uint32(1))) >> np.uint32(16)).astype(np.uint16)

    torch.manual_seed(0)
    L, h = 2, 256
    ref = torch.nn.ModuleDict(dict(lstm=torch.nn.LSTM(8, h, L, batch_first=True), full_c=torch.nn.Linear(h, 6)))
    sd = {k: v.detach() for k, v in ref.state_dict().items()}
    blob = pack.pack_rnn(sd, _lib.LSTM, h, L, 4, tensor_core=True).numpy()
    assert torch.equal(torch.from_numpy(bf16(np.float32([1.0, 1.00390625, 1.01171875, -3.1415927]))).to(torch.int32),
                       torch.tensor([1.0, 1.00390625, 1.01171875, -3.1415927]).to(torch.bfloat16).view(torch.int16).to(torch.int32) & 0xFFFF)
    for l in range(L):
        b = (sd[f"lstm.bias_ih_l{l}"] + sd[f"lstm.bias_hh_l{l}"]).numpy().reshape(4, 4, 64)        # [gate][chunk][j]
        assert np.array_equal(blob[l * 1024:(l + 1) * 1024].reshape(4, 4, 64), b.transpose(1, 0, 2))
    off = L * 1024
    assert np.array_equal(blob[off:off + h * 8].reshape(h, 8)[:, :6], sd["full_c.weight"].numpy().T)
    assert np.array_equal(blob[off + h * 8:off + h * 8 + 6], sd["full_c.bias"].numpy())
    pos = (off + h * 8 + 8 + 31) // 32 * 32
    w16 = blob[pos:].view(np.uint16)
    p = 0

    def slot(W, c, ks):                      # -> the [256 rows][len(ks)] block of chunk c in slot order
        rows = np.array([(n // 64) * h + c * 64 + n % 64 for n in range(256)])
        blk = bf16(W[rows][:, ks])                                                                  # [n][kk]
        return np.concatenate([blk[128 * r:128 * r + 128].reshape(16, 8, len(ks) // 8, 8).transpose(2, 0, 1, 3).reshape(-1)
                               for r in range(2)])                                                  # [rank][kk/8][n/8][n%8][kk%8]

    for l in range(L):
        wi, wh = sd[f"lstm.weight_ih_l{l}"].numpy(), sd[f"lstm.weight_hh_l{l}"].numpy()
        for c in range(4):
            if l == 0:
                want = slot(wi, c, np.arange(16) % 8)
                assert np.array_equal(w16[p:p + want.size], want), (l, c, "x0")
                p += want.size
            else:
                for s in range(4):
                    want = slot(wi, c, np.arange(64 * s, 64 * s + 64))
                    assert np.array_equal(w16[p:p + want.size], want), (l, c, "x", s)
                    p += want.size
            for s in range(4):
                want = slot(wh, c, np.arange(64 * s, 64 * s + 64))
                assert np.array_equal(w16[p:p + want.size], want), (l, c, "h", s)
                p += want.size
    assert p == w16.size


def test_cost_and_norm_caches_follow_content():
    """Round-1 advisor finding: the cost / statistics caches keyed on id() / data_ptr, so a helper that rebuilt
    `qw = [w] * 6` on every call (the id is reused after the list is freed) got the FIRST call's weights forever.
    Non-tensor inputs are no longer cached; tensor inputs are cached by object identity + version with the cache
    holding the tensors, so a freed-and-reallocated tensor at the same address cannot alias an entry."""
    import mpcdl
    m = mpcdl.Basic_Model(8, 16, 6, 1, False)
    cpu = torch.device("cpu")
    ref = torch.zeros(3, 6)
    for w in (1.0, 2.0, 3.0, 4.0):
        qw, rw = [w] * 6, [w / 2] * 2                      # fresh temporaries on every call
        assert m._cost((ref, qw, rw), cpu)[1] == [w] * 6 + [w / 2] * 2
    for w in (1.0, 2.0, 3.0):                              # tensors that die between calls (storage may be recycled)
        got = m._cost((ref, torch.full((6,), w), torch.full((2,), w)), cpu)[1]
        assert got == [w] * 8
    qw, rw = torch.ones(6), torch.ones(2)                  # in-place edits move the version counter
    assert m._cost((ref, qw, rw), cpu)[1] == [1.0] * 8
    qw.mul_(5.0)
    assert m._cost((ref, qw, rw), cpu)[1] == [5.0] * 6 + [1.0] * 2
    st = {k: torch.ones(8 if k[0] == "f" else 6) for k in ("f_mean", "f_std", "l_mean", "l_std")}
    a = m._norm(st, cpu)
    assert m._norm(st, cpu) is a                           # same tensors, unchanged -> cached
    st["f_std"].add_(1.0)
    b = m._norm(st, cpu)
    assert b is not a and float(b[8]) == 2.0
    for v in (1.0, 2.0):                                   # numpy statistics are converted on every call
        stn = {k: np.full(8 if k[0] == "f" else 6, v, dtype=np.float32) for k in st}
        assert float(m._norm(stn, cpu)[0]) == v
