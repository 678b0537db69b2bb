"""Drop-in nn.Modules with the reference's constructors, attributes and state_dict keys whose
forward / rollout run on the fused sm_100a kernels.

    Basic_Model(inputs, hidden, output, layers, do_drop)        model_scripts/basic_model.py:43-61
    GRU_Model / LSTM_Model(inputs, hidden, output, layers, state_f)   model_scripts/deep_models.py:44-92
    FeedforwardNoiseModel(state_size, action_size, history, **params) model_scripts/noise_model.py:19-52

The torch sub-modules (``layers``, ``gru``/``lstm``, ``full_c``, ``model``) are kept purely as parameter
containers so ``load_state_dict(torch.load(file))`` of the shipped checkpoints works unchanged
(eval_models.py:27-28, 54-55); rollout() is one launch of the library and no torch compute kernel (forward() of a recurrent
model adds a zero fill of its dummy controls and a copy of the carried h_state).  Inputs must be CUDA
tensors: there is no CPU path (the dispatcher raises for CPU tensors).
"""
from __future__ import annotations

import torch
import torch.nn as nn

from . import _lib, ops, pack

_ACT_MODULES = {"relu": nn.ReLU, "tanh": nn.Tanh, "leaky_relu": nn.LeakyReLU, "sigmoid": nn.Sigmoid,
                "selu": nn.SELU, "softplus": nn.Softplus, "identity": nn.Identity}


class _Packed:
    """Device-resident packed blob, rebuilt when any parameter changes (version counter / device)."""

    def __init__(self):
        self._key, self._blob = None, None

    def get(self, params, device, build):
        key = (str(device),) + tuple((p.data_ptr(), p._version) for p in params)
        if key != self._key:
            self._blob = build().to(device)
            self._key = key
        return self._blob


class _RolloutMixin:
    KIND = None

    def _params(self):
        """The module's Parameter objects, enumerated once (nn.Module.parameters() walks the module tree: ~40 us per
        call, a third of the host time of a rollout).  In-place updates are still seen through the version counters in
        _Packed; .to() / load_state_dict() drop the list.  After REPLACING a Parameter object by hand, call
        ``model._plist = None``."""
        pl = self.__dict__.get("_plist")
        if pl is None:
            pl = list(self.parameters())
            self.__dict__["_plist"] = pl
        return pl

    def _apply(self, fn, *args, **kwargs):
        self.__dict__["_plist"] = None
        return super()._apply(fn, *args, **kwargs)

    def load_state_dict(self, *args, **kwargs):
        self.__dict__["_plist"] = None
        return super().load_state_dict(*args, **kwargs)

    def _desc(self):
        raise NotImplementedError

    def _blob(self, device):
        raise NotImplementedError

    @staticmethod
    def _same(held, vals):
        """True if `vals` are the very tensor objects the cache holds, unchanged (version counters).  The cache keeps
        the source tensors alive, so neither an id nor a data pointer can be recycled under it; anything that is not a
        torch tensor (lists, numpy arrays, python floats) is never cached -- it is converted on every call.
        Blind spot, as for ``_Packed``: writes through ``.data`` do not move the version counter."""
        return len(held) == len(vals) and all(a is b and a._version == v for (a, v), b in zip(held, vals))

    def _norm(self, stats, device):
        """Device copy of the statistics vector (28 / 76 floats).  Cached only for torch tensors (see ``_same``)."""
        if stats is None:
            key = ("identity", str(device), self._desc()[2])
            if getattr(self, "_norm_key", None) != key:
                self._norm_id = pack.identity_norm(self._desc()[2]).to(device)
                self._norm_key = key
            return self._norm_id
        vals = [stats[k] for k in ("f_mean", "f_std", "l_mean", "l_std")]
        cacheable = all(isinstance(v, torch.Tensor) for v in vals)
        c = self.__dict__.get("_norm_cache")
        if cacheable and c is not None and c[0] == device and self._same(c[1], vals):
            return c[2]
        buf = pack.pack_norm({k: torch.as_tensor(v).detach().cpu() for k, v in stats.items()}).to(device)
        self.__dict__["_norm_cache"] = (device, [(v, v._version) for v in vals], buf) if cacheable else None
        return buf

    def _cost(self, cost, device):
        """(ref [H,6], qw [6], rw [2]) -> (device ref, 8 python floats).  Weights given as lists / numpy / CPU tensors
        are read on every call (8 floats); only torch tensors are cached, by object identity + version, with the cache
        holding them (weights that live on the GPU are read back -- a device sync -- once, not on every call)."""
        c = self.__dict__.get("_cost_cache")
        cacheable = all(isinstance(v, torch.Tensor) for v in cost)
        if cacheable and c is not None and c[0] == device and self._same(c[1], cost):
            return c[2]
        cref = torch.as_tensor(cost[0]).to(device, torch.float32).contiguous()
        cw = [float(v) for v in torch.as_tensor(cost[1]).detach().cpu().reshape(-1)] + \
             [float(v) for v in torch.as_tensor(cost[2]).detach().cpu().reshape(-1)]
        if len(cw) != 8:
            raise RuntimeError(f"mpcdl: cost weights must be qw [6] and rw [2], got {len(cw)} values")
        self.__dict__["_cost_cache"] = (device, [(v, v._version) for v in cost], (cref, cw)) if cacheable else None
        return cref, cw

    @torch.no_grad()
    def rollout(self, init, U, stats=None, *, noise=None, eps=None, seed=0, traj_offset=0, cost=None,
                return_states=True, stateful=None, carry=None, tune=(), precision=None, sample=None, shared_init=False,
                out=None):
        """Closed-loop rollout of N trajectories over H steps in one kernel launch.

        init  s0 [N,6] (MLP) or window [N,T,8] (GRU/LSTM), raw units; U [N,H,2] controls.
        stats dict(f_mean,f_std,l_mean,l_std) from ``mpcdl.data.zscore_stats`` (None = identity).
        noise FeedforwardNoiseModel (residual injection); eps [N,H,6] draws or None -> Philox(seed).
        cost  (ref [H,6], qw [6], rw [2]) -> per-trajectory quadratic tracking cost.
        precision "fp32" (FFMA kernels, the parity path) | "tf32x3" | "tf32" (tcgen05 tensor-core kernels,
              GRU/LSTM hidden 64, window 4 or 8) | "bf16" (tcgen05 kernel of the LSTM with hidden 256: BF16
              operands, FP32 accumulation and cell state); anything else raises -- there is no silent fallback.
        sample  MPPI call: dict(n=N, sigma=(s_a, s_steer), lo=(..), hi=(..), seed=int) -> U is the NOMINAL sequence
              [H,2] and trajectory n applies clamp(U + sigma * z_n) with z_n drawn in-kernel (Philox keyed by
              seed, traj_offset + n, t); no [N,H,2] tensor exists.  shared_init: init is ONE state / window.
        out   (states [N,H,6] | None, costs [N] | None): caller-owned result buffers to write into instead of fresh ones.
        Returns (states [N,H,6] or None, costs [N] or None[, carry if stateful])."""
        hidden, layers, window = self._desc()
        dev = U.device
        stateful = bool(getattr(self, "state_f", False)) if stateful is None else bool(stateful)
        stateful = stateful and self.KIND != _lib.MLP
        ctrl, ctrl_seed, n_samples = [], 0, 0
        if sample is not None:
            ctrl = [float(v) for k in ("sigma", "lo", "hi") for v in sample[k]]
            ctrl_seed, n_samples = int(sample.get("seed", 0)), int(sample["n"])
        if stateful and carry is None:
            planes = layers * (2 if self.KIND == _lib.LSTM else 1)
            carry = torch.zeros(planes, n_samples or U.shape[0], hidden, device=dev)
        nblob, ncfg = None, []
        if noise is not None:
            nblob, ncfg = noise.packed(dev), noise.cfg()
        cref, cw = None, []
        if cost is not None:
            cref, cw = self._cost(cost, dev)
        precision = getattr(self, "precision", "fp32") if precision is None else precision
        if precision not in _lib.PRECISION:
            raise RuntimeError(f"mpcdl: unknown precision {precision!r} (fp32 | tf32x3 | tf32 | bf16)")
        prec = _lib.PRECISION[precision]
        # the blob of the precision in use only (the other one is never packed or uploaded)
        states, costs = ops.OPS[self.KIND](self._blob(dev) if prec == 0 else None, self._norm(stats, dev), init, U, hidden, layers, window,
                                           stateful, carry if stateful else None, nblob, ncfg, eps, int(seed),
                                           int(traj_offset), cref, cw, bool(return_states), cost is not None,
                                           list(tune), prec, self._blob_tc(dev) if prec else None,
                                           ctrl, ctrl_seed, n_samples, bool(shared_init),
                                           out[0] if out is not None else None, out[1] if out is not None else None)
        out = (states if return_states else None, costs if cost is not None else None)
        return out + (carry,) if stateful else out


class Basic_Model(_RolloutMixin, nn.Module):
    KIND = _lib.MLP

    def __init__(self, inputs, hidden, output, layers, do_drop):
        super().__init__()
        if inputs != 8 or output != 6:
            raise RuntimeError("mpcdl: the fused kernels implement the BARC shape inputs=8, output=6")
        self.layers = nn.ModuleList()
        self.layers.append(nn.Linear(inputs, hidden))
        for _ in range(layers):
            self.layers.append(nn.Linear(hidden, hidden))
            if do_drop:
                self.layers.append(nn.Dropout(0.1))       # shifts the Linear indices (basic_model.py:53-54)
        self.layers.append(nn.Linear(hidden, output))
        self._hidden, self._nlayers = hidden, layers
        self.precision = "fp32"        # "fp32" | "tf32x3" | "tf32" (tcgen05, hidden 64)
        self._packed = _Packed()

    def _desc(self):
        return self._hidden, self._nlayers, 0

    def _blob(self, device):
        return self._packed.get(self._params(), device,
                                lambda: pack.pack_mlp(self.state_dict(), self._hidden, self._nlayers))

    def _blob_tc(self, device):
        if not hasattr(self, "_packed_tc"):
            self._packed_tc = _Packed()
        return self._packed_tc.get(self._params(), device,
                                   lambda: pack.pack_mlp(self.state_dict(), self._hidden, self._nlayers, tensor_core=True))

    def forward(self, x):
        """x [B,8] z-scored features -> [B,6] z-scored next state (one fused launch, H = 1).
        Dropout (do_drop) is the identity in eval mode, which is the only mode the rollout path has."""
        if self.training and any(isinstance(m, nn.Dropout) for m in self.layers):
            raise RuntimeError("mpcdl: training-mode dropout is outside the fused inference path; call .eval()")
        x = x.contiguous()
        states, _ = self.rollout(x[:, :6].contiguous(), x[:, None, 6:8].contiguous(), None)
        return states[:, 0, :]


class _RNNBase(_RolloutMixin, nn.Module):
    def __init__(self, inputs, hidden, output, layers, state_f):
        super().__init__()
        if inputs != 8 or output != 6:
            raise RuntimeError("mpcdl: the fused kernels implement the BARC shape inputs=8, output=6")
        self.hidden = hidden
        self.layers = layers
        self.state_f = state_f
        self.h_state = None
        self.precision = "fp32"        # default arithmetic of forward()/rollout(): "fp32" | "tf32x3" | "tf32" (tcgen05)
        self._packed = _Packed()
        self._window = None

    def _rnn(self):
        raise NotImplementedError

    def _desc(self):
        return self.hidden, self.layers, self._window

    def _blob(self, device):
        return self._packed.get(self._params(), device,
                                lambda: pack.pack_rnn(self.state_dict(), self.KIND, self.hidden, self.layers, self._window))

    def _blob_tc(self, device):
        if not hasattr(self, "_packed_tc"):
            self._packed_tc = _Packed()
        return self._packed_tc.get(self._params(), device,
                                   lambda: pack.pack_rnn(self.state_dict(), self.KIND, self.hidden, self.layers, self._window,
                                                         tensor_core=True))

    def rollout(self, init, U, stats=None, **kw):
        self._window = int(init.shape[-2])
        return super().rollout(init, U, stats, **kw)

    def forward(self, x):
        """x [B,T,8] z-scored window -> [B,6].  state_f: h_state carries across calls and has a fixed
        batch size (deep_models.py:61-65); it is allocated on x.device (the reference allocates on a
        module-global device)."""
        if not self.state_f or self.h_state is None:
            self.init_hidden(x.size(0), x.device)
        x = x.contiguous()
        B = x.size(0)
        carry = self._carry_in(B)
        U = torch.zeros(B, 1, 2, device=x.device)
        states, _, carry = self.rollout(x, U, None, stateful=True, carry=carry)
        self._carry_out(carry)
        return states[:, 0, :]


class GRU_Model(_RNNBase):
    KIND = _lib.GRU

    def __init__(self, inputs, hidden, output, layers, state_f):
        super().__init__(inputs, hidden, output, layers, state_f)
        self.gru = nn.GRU(inputs, hidden, layers, batch_first=True)
        self.full_c = nn.Linear(hidden, output)

    def init_hidden(self, batch_s, device=None):
        device = device if device is not None else self.full_c.weight.device
        self.h_state = torch.zeros(self.layers, batch_s, self.hidden, device=device)

    def _carry_in(self, B):
        if self.h_state.shape[1] != B:
            raise RuntimeError(f"Expected hidden size {(self.layers, B, self.hidden)}, got {list(self.h_state.shape)}")
        return self.h_state.detach().clone().contiguous()

    def _carry_out(self, carry):
        self.h_state = carry


class LSTM_Model(_RNNBase):
    KIND = _lib.LSTM

    def __init__(self, inputs, hidden, output, layers, state_f):
        super().__init__(inputs, hidden, output, layers, state_f)
        self.lstm = nn.LSTM(inputs, hidden, layers, batch_first=True)
        self.full_c = nn.Linear(hidden, output)

    def init_hidden(self, batch_s, device=None):
        device = device if device is not None else self.full_c.weight.device
        self.h_state = (torch.zeros(self.layers, batch_s, self.hidden, device=device),
                        torch.zeros(self.layers, batch_s, self.hidden, device=device))

    def _carry_in(self, B):
        if self.h_state[0].shape[1] != B:
            raise RuntimeError(f"Expected hidden[0] size {(self.layers, B, self.hidden)}, got {list(self.h_state[0].shape)}")
        return torch.cat([self.h_state[0].detach(), self.h_state[1].detach()], dim=0).contiguous()

    def _carry_out(self, carry):
        self.h_state = (carry[:self.layers], carry[self.layers:])


def build_mlp(input_size, output_size, n_layers, size, activation="tanh", output_activation="identity", **params):
    """Same module indices as utils/pytorch_utils.py:19-57 (Linear at even positions)."""
    seq, d = [], input_size
    for _ in range(n_layers):
        seq += [nn.Linear(d, size), _ACT_MODULES[activation]()]
        d = size
    seq += [nn.Linear(d, output_size), _ACT_MODULES[output_activation]()]
    return nn.Sequential(*seq)


class FeedforwardNoiseModel(nn.Module):
    """Residual-noise net; inside ``rollout(noise=...)`` it is evaluated by the fused kernel with the
    history buffering of CasadiDynamicCLBicycleNoise.get_prediction (noise_model.py:95-103)."""

    def __init__(self, state_size, action_size, history, **params):
        super().__init__()
        if state_size != 6 or action_size != 2:
            raise RuntimeError("mpcdl: the fused kernels implement state_size=6, action_size=2")
        self.state_size, self.action_size, self.history = state_size, action_size, history
        self.n_layers, self.size = params["n_layers"], params["size"]
        self.activation = params.get("activation", "tanh")
        if self.activation not in _lib.ACT or params.get("output_activation", "identity") != "identity":
            raise RuntimeError("mpcdl: fused noise net supports hidden activation tanh/relu/sigmoid/identity and "
                               "identity output activation")
        self.model = build_mlp(input_size=(state_size + action_size) * history, output_size=state_size * 2, **params)
        self._packed = _Packed()

    def cfg(self):
        return [self.history, self.size, self.n_layers, _lib.ACT[self.activation]]

    def _apply(self, fn, *args, **kwargs):
        self.__dict__["_plist"] = None
        return super()._apply(fn, *args, **kwargs)

    def packed(self, device):
        pl = self.__dict__.get("_plist")                 # enumerated once, like _RolloutMixin._params
        if pl is None:
            pl = self.__dict__["_plist"] = list(self.model.parameters())
        return self._packed.get(pl, device,
                                lambda: pack.pack_noise(self.model.state_dict(), self.history, self.size, self.n_layers))

    def forward(self, q, u, eps=None, return_params=False):
        """q [B,6*history] (q order, oldest first), u [B,2*history] -> Normal(mu, softplus(p)).rsample() [B,6]
        (noise_model.py:34-40) in one kernel launch.  The reference draws eps from torch's global RNG;
        here eps is either supplied ([B,6], the parity path) or drawn in-kernel by Philox keyed with
        ``torch.initial_seed()`` and a per-module call counter (reproducible under torch.manual_seed)."""
        q, u = q.contiguous(), u.contiguous()
        self._calls = getattr(self, "_calls", 0) + (eps is None)
        out, mu, scale = torch.ops.mpcdl.noise_forward(self.packed(q.device), self.cfg(), q, u, eps,
                                                       int(torch.initial_seed()) & 0x7FFFFFFFFFFFFFFF, 0, self._calls, bool(return_params))
        return (out, mu, scale) if return_params else out

    def get_prediction(self, q, u, eps=None):
        """numpy in, numpy out (noise_model.py:42-45); the arrays go to the device the parameters live on."""
        import numpy as np
        dev = next(self.model.parameters()).device
        qt = torch.from_numpy(np.ascontiguousarray(q, dtype=np.float32)).to(dev)
        ut = torch.from_numpy(np.ascontiguousarray(u, dtype=np.float32)).to(dev)
        et = None if eps is None else torch.from_numpy(np.ascontiguousarray(eps, dtype=np.float32)).to(dev)
        return self(qt, ut, et).cpu().numpy()

    def export(self, path="../model_data", name="noise_model.pkl"):
        import os
        torch.save(self.model.state_dict(), os.path.join(path, name))

    def load(self, path="../model_data", name="noise_model.pkl"):
        import os
        self.model.load_state_dict(torch.load(os.path.join(path, name), map_location="cpu"))


Q_PERM = [4, 5, 3, 0, 1, 2]    # q = frame-order state[Q_PERM]   (noise_model.py:108, utils/data_utils.py:47-49)
Q_INV = [3, 4, 5, 2, 0, 1]     # state = q[Q_INV]


class CasadiDynamicCLBicycleNoise:
    """nominal step + learned residual noise (noise_model.py:54-111), same constructor and methods.

    The reference's nominal model is the CasADi bicycle of ``mpclab_common`` -- a private submodule that is not
    in the tree.  Here ``nominal_model`` is either one of the learned dynamics modules of this
    package (``Basic_Model`` / ``GRU_Model`` / ``LSTM_Model``: the substitution BASELINE.json's north_star makes)
    together with its z-score ``stats``, or any object with the reference's ``qu2state / step / state2q``
    protocol (then ``vehicle_state`` must be passed in, there is no VehicleState type here).
    ``get_prediction`` keeps the reference's batch-1, stateful-deque semantics; ``rollout`` is the fused,
    batched form of calling it in the open-loop feedback of run_dynamics_model (scripts/test_model.py:67-87)."""

    def __init__(self, nominal_model, noise_model, dynamics_uses_frenet, track=None, history=1, stats=None):
        from collections import deque
        if dynamics_uses_frenet:
            raise RuntimeError("mpcdl: Frenet-frame dynamics need mpclab_common's track (absent); use global coordinates")
        self.nominal_model, self.noise_model = nominal_model, noise_model
        self.dynamics_uses_frenet, self.track, self.history, self.stats = dynamics_uses_frenet, track, history, stats
        self.q_buffer, self.u_buffer = deque(maxlen=history), deque(maxlen=history)

    def _device(self):
        return next(self.noise_model.model.parameters()).device

    def get_nominal_prediction(self, q, u, vehicle_state=None):
        import numpy as np
        assert q.shape == (6,) and u.shape == (2,), f"Got unexpected shape: q={q.shape}, u={u.shape}"
        m = self.nominal_model
        if isinstance(m, Basic_Model):
            dev = self._device()
            s0 = torch.from_numpy(np.asarray(q, dtype=np.float32)[Q_INV][None].copy()).to(dev)
            U = torch.from_numpy(np.asarray(u, dtype=np.float32)[None, None].copy()).to(dev)
            return m.rollout(s0, U, self.stats)[0][0, 0].cpu().numpy()[Q_PERM]
        if hasattr(m, "qu2state"):
            if vehicle_state is None:
                raise RuntimeError("mpcdl: pass vehicle_state (mpclab_common's VehicleState is not available here)")
            m.qu2state(vehicle_state, q, u)
            m.step(vehicle_state)
            return m.state2q(vehicle_state)
        raise RuntimeError("mpcdl: nominal_model must be a Basic_Model or follow the qu2state/step/state2q protocol "
                           "(window models are rolled out with .rollout)")

    def _push(self, q, u):
        self.q_buffer.append(q)
        self.u_buffer.append(u)
        while len(self.q_buffer) < self.history:       # noise_model.py:97-99
            self.q_buffer.append(q)
            self.u_buffer.append(u)

    def get_prediction(self, q, u, vehicle_state=None, eps=None):
        import numpy as np
        assert len(q.shape) == 1 and len(u.shape) == 1
        self._push(q, u)
        nominal_next_q = self.get_nominal_prediction(q, u, vehicle_state)
        noise = self.noise_model.get_prediction(np.asarray(self.q_buffer).reshape(1, -1),
                                                np.asarray(self.u_buffer).reshape(1, -1), eps)
        return nominal_next_q + noise

    def rollout(self, q0, U, eps=None, seed=0, traj_offset=0, **kw):
        """q0 [N,6] (q order; or a frame-order window [N,T,8] for GRU/LSTM nominal models), U [N,H,2] ->
        predicted q [N,H,6]: nominal learned step + residual noise, the whole horizon in one launch."""
        m = self.nominal_model
        if not isinstance(m, _RolloutMixin):
            raise RuntimeError("mpcdl: the fused rollout needs a learned nominal model of this package")
        init = q0[:, Q_INV].contiguous() if q0.dim() == 2 else q0
        states = m.rollout(init, U, self.stats, noise=self.noise_model, eps=eps, seed=seed, traj_offset=traj_offset, **kw)[0]
        return states[:, :, Q_PERM]

    def step(self, state):
        raise NotImplementedError("mpcdl: VehicleState stepping belongs to mpclab_common (absent)")


class DynamicsNN(CasadiDynamicCLBicycleNoise):
    """The noise net as the whole dynamics (noise_model.py:113-128): next q = sample of the net."""

    def get_prediction(self, q, u, vehicle_state=None, eps=None):
        import numpy as np
        assert len(q.shape) == 1 and len(u.shape) == 1
        self._push(q, u)
        return self.noise_model.get_prediction(np.asarray(self.q_buffer).reshape(1, -1),
                                               np.asarray(self.u_buffer).reshape(1, -1), eps)

    def rollout(self, q0, U, eps=None, seed=0, traj_offset=0):
        """q0 [N,6], U [N,H,2] -> [N,H,6]; out[:,t] = get_prediction(out[:,t-1], U[:,t]) with fresh deques."""
        nm = self.noise_model
        if nm.history != self.history:
            raise RuntimeError(f"mpcdl: wrapper history {self.history} != noise model history {nm.history}")
        return torch.ops.mpcdl.dynamics_nn_rollout(nm.packed(U.device), nm.cfg(), q0.contiguous(), U.contiguous(), eps,
                                                   int(seed), int(traj_offset))


def from_checkpoint_name(name, sd=None, device=None):
    """File-name schema is the config (eval_models.py:19-25, 43-52):
    Basic_{h}_{L}_{batch} | {GRU,LSTM}_{h}_{L}_{ft}_{batch}_{state_f}."""
    p = name.replace(".pt", "").replace(".npz", "").split("_")
    kind, hidden, layers = p[0], int(p[1]), int(p[2])
    if kind == "Basic":
        m, meta = Basic_Model(8, hidden, 6, layers, False), dict(kind="mlp", window=0, state_f=False)
    else:
        state_f = p[-1] == "True"
        m = (GRU_Model if kind == "GRU" else LSTM_Model)(8, hidden, 6, layers, state_f)
        m._window = int(p[3])
        meta = dict(kind=kind.lower(), window=int(p[3]), state_f=state_f)
    meta.update(hidden=hidden, layers=layers)
    if sd is not None:
        m.load_state_dict({k: torch.as_tensor(v) for k, v in sd.items()})
    m.eval()
    if device is not None:
        m.to(device)
    return m, meta


def load_model(path, device=None):
    """Checkpoint file -> (module, meta): the file name is the config (eval_models.py:19-25, 43-52), the state_dict loads
    unchanged (``.pt`` via torch.load as eval_models.py:28 does, or the ``.npz`` export), and the z-score statistics
    persisted beside it by ``mpcdl.data.save_stats(mpcdl.data.stats_path(path), stats)`` -- the reference computes them
    at import and throws them away -- are attached as ``module.stats`` (None if there is no such file)."""
    import os

    import numpy as np

    from . import data
    if path.endswith(".npz"):
        sd = {k: torch.from_numpy(v) for k, v in np.load(path).items()}
    else:
        sd = torch.load(path, map_location="cpu")
    m, meta = from_checkpoint_name(os.path.basename(path), sd, device)
    sp = data.stats_path(path)
    m.stats = data.load_stats(sp) if os.path.exists(sp) else None
    return m, meta
