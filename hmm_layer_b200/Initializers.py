"""Initialisers of the profile-HMM kernels (reference: hmm_layer/Initializers.py).  An initialiser here is anything
`materialize` understands: a callable `f(shape) -> tensor`, an array-like, or a scalar.  The defaults restate the
reference's default values (make_default_transition_init :367-422, EntryInitializer :184-214, ExitInitializer
:222-249, MatchTransitionInitializer :258-299, make_default_flank_init :363-364).

The reference's default EMISSION initialiser is a placeholder (a 15-class gene-model kernel used as amino-acid
background, :158-181); the default here is a uniform background over the alphabet, and real background
frequencies are passed in by the caller like in learnMSA."""
import math

import numpy as np
import torch


def materialize(init, shape):
    """float32 tensor of `shape` from a callable, an array-like (broadcast) or a scalar."""
    shape = tuple(shape)
    if callable(init) and not isinstance(init, torch.Tensor):
        v = init(shape)
    else:
        v = init
    v = torch.as_tensor(np.asarray(v.detach().cpu()) if isinstance(v, torch.Tensor) else np.asarray(v), dtype=torch.float32)
    return v.expand(shape).clone() if v.shape != shape else v.clone()


class ConstantInitializer:
    def __init__(self, value):
        self.value = value

    def __call__(self, shape, dtype=None, device=None):
        v = torch.as_tensor(np.asarray(self.value), dtype=torch.float32)
        return v.expand(tuple(shape)).clone()

    def __repr__(self):
        return f"Const(shape={np.shape(self.value)})"


class RandomNormalInitializer:
    def __init__(self, mean=0.0, stddev=0.05):
        self.mean, self.stddev = mean, stddev

    def __call__(self, shape, dtype=None, device=None):
        return self.mean + self.stddev * torch.randn(tuple(shape))

    def __repr__(self):
        return f"Norm({self.mean}, {self.stddev})"


class EntryInitializer:
    """First entry 0, the rest log(1 / (n - 1)): about half of the entry mass on the first match state."""

    def __call__(self, shape, dtype=None, device=None):
        n = shape[0]
        rest = math.log(1.0 / (n - 1)) if n > 1 else 0.0
        return torch.cat([torch.zeros(1), torch.full((n - 1,), rest)])

    def __repr__(self):
        return "DefaultEntry()"


class ExitInitializer:
    """log(0.5 / (n - 1)) everywhere."""

    def __call__(self, shape, dtype=None, device=None):
        n = shape[0]
        return torch.full(tuple(shape), math.log(0.5 / (n - 1)) if n > 1 else math.log(0.5))

    def __repr__(self):
        return "DefaultExit()"


class MatchTransitionInitializer:
    """Entry i of softmax(val + noise) scaled by the probability of not leaving through the end state."""

    def __init__(self, val, i, scale):
        self.val, self.i, self.scale = torch.tensor(val, dtype=torch.float32), i, scale

    def __call__(self, shape, dtype=None, device=None):
        n = shape[0]
        z = self.scale * torch.randn(n, 1)
        p_exit = 0.5 / (n - 1) if n > 1 else 0.5
        prob = torch.softmax(self.val[None, :] + z, dim=-1) * (1 - p_exit)
        return torch.log(prob[:, self.i])

    def __repr__(self):
        return f"DefaultMatchTransition({self.val.tolist()}, {self.i}, {self.scale})"


def make_default_flank_init():
    return ConstantInitializer(0.)


def make_default_transition_init(MM=1, MI=-1, MD=-1, II=-0.5, IM=0, DM=0, DD=-0.5, FC=0, FE=-1, R=-9, RF=0, T=0, scale=0.1):
    m = [MM, MI, MD]
    return {
        "begin_to_match": EntryInitializer(), "match_to_end": ExitInitializer(),
        "match_to_match": MatchTransitionInitializer(m, 0, scale), "match_to_insert": MatchTransitionInitializer(m, 1, scale),
        "insert_to_match": RandomNormalInitializer(IM, scale), "insert_to_insert": RandomNormalInitializer(II, scale),
        "match_to_delete": MatchTransitionInitializer(m, 2, scale),
        "delete_to_match": RandomNormalInitializer(DM, scale), "delete_to_delete": RandomNormalInitializer(DD, scale),
        "left_flank_loop": RandomNormalInitializer(FC, scale), "left_flank_exit": RandomNormalInitializer(FE, scale),
        "right_flank_loop": RandomNormalInitializer(FC, scale), "right_flank_exit": RandomNormalInitializer(FE, scale),
        "unannotated_segment_loop": RandomNormalInitializer(FC, scale), "unannotated_segment_exit": RandomNormalInitializer(FE, scale),
        "end_to_unannotated_segment": RandomNormalInitializer(R, scale), "end_to_right_flank": RandomNormalInitializer(RF, scale),
        "end_to_terminal": RandomNormalInitializer(T, scale),
    }


class EmissionInitializer:
    """Match-state emission kernel: the background log-distribution repeated for every match state."""

    def __init__(self, dist=None):
        self.dist = dist

    def __call__(self, shape, dtype=None, device=None):
        s = shape[-1]
        d = torch.zeros(s) if self.dist is None else torch.as_tensor(np.asarray(self.dist), dtype=torch.float32).reshape(-1)
        assert d.numel() == s, f"background distribution has {d.numel()} entries, the alphabet {s}"
        return d.expand(tuple(shape)).clone()


def make_default_emission_init():
    return EmissionInitializer()


def make_default_insertion_init():
    return EmissionInitializer()
