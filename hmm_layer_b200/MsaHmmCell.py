"""HmmCell / MsaHmmCell with the reference's interface (reference: hmm_layer/MsaHmmCell.py).

The cell owns an emitter (or list) and a transitioner plugin, exactly like the reference.  What differs:
  * `make_reverse_direction_offspring` gives the reverse cell its own reversed *view* of the transitioner, so
    building the pair no longer flips the forward cell to A^T (defect D1);
  * the chunked initial state never multiplies the identity in place (defect D2);
  * constants are created on the parameters' device (the reference is CPU-only)."""
import torch
import torch.nn as nn


class HmmCell(nn.Module):
    def __init__(self, num_states, dim, emitter, transitioner, use_step_counter=False, use_fake_step_counter=False, **kwargs):
        super().__init__(**kwargs)
        self.num_states = num_states
        self.num_models = len(num_states)
        self.max_num_states = max(num_states)
        self.dim = dim
        self.emitter = nn.ModuleList(emitter if isinstance(emitter, (list, tuple)) else [emitter])
        self.transitioner = transitioner
        self.epsilon = 1e-16
        self.reverse = False
        self.use_step_counter = use_step_counter
        self.use_fake_step_counter = use_fake_step_counter
        self.recurrent_init()

    def __setattr__(self, name, value):
        if isinstance(value, nn.Module):   # a plugin was swapped: derived tensors and the cached module list are stale
            self.__dict__.pop("_ri_modules", None)
            self.__dict__["_ri_key"] = None
        super().__setattr__(name, value)

    # ---- parameter -> matrices, once per call (reference :41-59) ----
    def _param_key(self):
        """Identity + version of every parameter / buffer the derived tensors depend on, including those of the base
        transitioner behind a reverse cell's view (it is deliberately not a registered sub-module) and the identity of
        every plugin (an emitter swapped inside the ModuleList changes the key).  Only consulted after
        `freeze_derived(True)`."""
        ids = (id(self.transitioner), *(id(e) for e in self.emitter))
        cached = self.__dict__.get("_ri_modules")
        if cached is None or cached[0] != ids:
            mods = list(self.modules())
            base = getattr(self.transitioner, "base", None)
            if isinstance(base, nn.Module):
                mods += [m for m in base.modules() if all(m is not x for x in mods)]
            cached = (ids, tuple(mods))
            object.__setattr__(self, "_ri_modules", cached)
        key = [ids]
        for m in cached[1]:
            for t in m._parameters.values():
                if t is not None:
                    key.append((t.data_ptr(), t._version))
            for t in m._buffers.values():
                if t is not None:
                    key.append((t.data_ptr(), t._version))
        return tuple(key)

    def freeze_derived(self, on=True):
        """Opt-in for inference loops: after freeze_derived(True), `recurrent_init` rebuilds A / B / pi only when the
        identity or the autograd version of a parameter changed (optimizer steps, load_state_dict, in-place ops and
        plugin swaps are all seen).  NOT seen: writes through `.data` (`p.data.copy_(...)`), which do not bump the
        version -- call `invalidate()` after those.  Default (off) = the reference's behaviour: rebuild on every call
        (MsaHmmCell.py:41-50)."""
        object.__setattr__(self, "_frozen", bool(on))
        object.__setattr__(self, "_ri_key", None)
        return self

    def invalidate(self):
        """Forget the derived tensors kept by `freeze_derived(True)`; the next call rebuilds them."""
        object.__setattr__(self, "_ri_key", None)

    def derived_fresh(self):
        """True when frozen derived tensors exist and still match the parameters (always False when not frozen)."""
        return bool(self.__dict__.get("_frozen", False)) and self.__dict__.get("_ri_key") is not None \
            and self.__dict__.get("_ri_key") == self._param_key()

    def recurrent_init(self, fresh=False):
        """Derived tensors of this call (reference MsaHmmCell.py:41-59): rebuilt on every call like the reference,
        unless `freeze_derived(True)` was requested and no parameter changed.  fresh=True rebuilds regardless (a
        differentiable call needs its own autograd graph from the parameters)."""
        if not self.reverse and self.use_step_counter:
            self.step_counter = torch.tensor(-1, dtype=torch.int32)
        frozen = self.__dict__.get("_frozen", False)
        key = self._param_key() if frozen else None
        if frozen and not fresh and self.__dict__.get("_ri_key") == key:
            return
        self.transitioner.recurrent_init()
        for em in self.emitter:
            em.recurrent_init()
        self.log_A_dense = self.transitioner.make_log_A()
        self.log_A_dense_t = self.log_A_dense.transpose(1, 2)
        self.init_dist = self.make_initial_distribution()
        object.__setattr__(self, "_ri_key", key)
        object.__setattr__(self, "_gene_params", None)

    def make_initial_distribution(self):
        return self.transitioner.make_initial_distribution()

    @property
    def device(self):
        return self.init_dist.device

    def emission_probs(self, inputs, end_hints=None, training=False):
        em_probs = None   # (iterating the ModuleList: slicing it builds a new nn.ModuleList on every call)
        for em in self.emitter:
            e = em(inputs, end_hints=end_hints, training=training)
            em_probs = e if em_probs is None else em_probs * e
        return em_probs

    # ---- one DP step (reference :73-106) ----
    def forward(self, emission_probs, states, training=None, init=False):
        from . import ops
        A = self.transitioner.A_transposed if self.reverse else self.transitioner.A
        out, scaled, loglik = ops.cell_step(emission_probs, states[0], states[1], A, self.num_models,
                                            reverse=self.reverse, init=init)
        if not self.reverse and self.use_step_counter:
            self.step_counter += 1
        return out, [scaled, loglik]

    # ---- initial states (reference :108-142) ----
    def get_initial_state(self, inputs=None, batch_size=None, parallel_factor=1):
        k, q, dev = self.num_models, self.max_num_states, self.device
        if parallel_factor == 1:
            if self.reverse:
                init = torch.ones((k * batch_size, q), dtype=torch.float32, device=dev)
            else:
                init = self.make_initial_distribution().detach().reshape(k, 1, q).expand(k, batch_size, q).reshape(-1, q).contiguous()
            return [init, torch.zeros((k * batch_size, 1), dtype=torch.float32, device=dev)]
        pf = parallel_factor
        b = batch_size // pf
        A = self.transitioner.A.detach()
        eye = torch.eye(q, dtype=torch.float32, device=dev).expand(k, b, pf, q, q)
        if self.reverse:
            fe = inputs[:, 0, :].reshape(k, b, pf, q).roll(-1, dims=2)
            rows = fe.unsqueeze(-1) * A.transpose(1, 2)[:, None, None]       # [i, j] = e_next(i) * A[j, i]
            boundary = pf - 1
        else:
            rows = A[:, None, None].expand(k, b, pf, q, q)
            boundary = 0
        is_boundary = (torch.arange(pf, device=dev) == boundary).view(1, 1, pf, 1, 1)
        init = torch.where(is_boundary, eye, rows).reshape(k * batch_size, q * q).contiguous()
        return [init, torch.zeros((k * batch_size, q), dtype=torch.float32, device=dev)]

    def get_aux_loss(self):
        return sum(em.get_aux_loss() for em in self.emitter)

    def get_prior_log_density(self):
        em_priors = [torch.sum(em.get_prior_log_density(), dim=1) for em in self.emitter]
        trans_priors = self.transitioner.get_prior_log_densities()
        return sum(em_priors) + sum(trans_priors.values())

    def make_reverse_direction_offspring(self):
        rev = HmmCell.__new__(HmmCell)
        nn.Module.__init__(rev)
        rev.num_states, rev.num_models, rev.max_num_states, rev.dim = self.num_states, self.num_models, self.max_num_states, self.dim
        rev.emitter = self.emitter
        rev.transitioner = self.transitioner.reversed_view() if hasattr(self.transitioner, "reversed_view") else self.transitioner
        rev.epsilon, rev.use_step_counter, rev.use_fake_step_counter = self.epsilon, False, False
        rev.reverse = True
        rev.recurrent_init()
        return rev

    def reverse_direction(self, reverse=True):
        """Reference :160-162.  Only flips this cell; a transitioner shared with another cell is swapped for a
        reversed view instead of being mutated."""
        self.reverse = reverse
        object.__setattr__(self, "_ri_key", None)   # derived tensors belong to the previous orientation
        if reverse and hasattr(self.transitioner, "reversed_view"):
            self.transitioner = self.transitioner.reversed_view()
        elif not reverse and hasattr(self.transitioner, "base"):
            self.transitioner = self.transitioner.base


class MsaHmmCell(HmmCell):
    """Profile-HMM cell, q = 2*length+3 (reference :164-182).  Defaults to ProfileHMMEmitter / ProfileHMMTransitioner
    like the reference.  The plugins learn the model lengths BEFORE the base constructor runs its first
    `recurrent_init` (the shipped order, :179-182, initialises from plugins that do not know their size yet),
    and the emitter is built for `dim` input symbols right away."""

    def __init__(self, length, dim=24, emitter=None, transitioner=None, **kwargs):
        from .Emitter import ProfileHMMEmitter
        from .Transitioner import ProfileHMMTransitioner
        self.length = [length] if not isinstance(length, list) else length
        k = len(self.length)
        if emitter is None:
            emitter = _default_emitter(k)
        if transitioner is None:
            transitioner = _default_transitioner(k)
        for em in (emitter if isinstance(emitter, (list, tuple)) else [emitter]):
            if hasattr(em, "set_lengths"):
                em.set_lengths(self.length)
            if hasattr(em, "build") and not getattr(em, "built", True):
                em.build((k, 1, 1, dim))
        if hasattr(transitioner, "set_lengths"):
            transitioner.set_lengths(self.length)
        if hasattr(transitioner, "build"):
            transitioner.build()
        super().__init__([2 * l + 3 for l in self.length], dim, emitter, transitioner, **kwargs)


def _default_emitter(k):
    from . import Initializers as I
    from .Emitter import ProfileHMMEmitter
    return ProfileHMMEmitter(emission_init=[I.make_default_emission_init() for _ in range(k)],
                             insertion_init=[I.make_default_insertion_init() for _ in range(k)])


def _default_transitioner(k):
    from . import Initializers as I
    from .Transitioner import ProfileHMMTransitioner
    return ProfileHMMTransitioner(transition_init=[I.make_default_transition_init() for _ in range(k)],
                                  flank_init=[I.make_default_flank_init() for _ in range(k)])
