"""hmm_layer_b200 — B200 (sm_100a) kernels for the hmm_layer forward / forward-backward path behind the
reference's Python interface.  Module names mirror the reference package (`hmm_layer/`)."""
from .MsaHmmCell import HmmCell, MsaHmmCell
from .MsaHMMLayer import MsaHmmLayer
from .BaseRNN import BaseRNN
from .Bidirectional import Bidirectional
from .TotalProbabilityCell import TotalProbabilityCell
from .dense_plugins import DenseEmitter, DenseTransitioner
from .Emitter import ProfileHMMEmitter
from .Transitioner import ProfileHMMTransitioner
from .gene_pred_hmm_emitter import GenePredHMMEmitter, SimpleGenePredHMMEmitter
from .gene_pred_hmm_transitioner import (GenePredHMMTransitioner, GenePredMultiHMMTransitioner,
                                         SimpleGenePredHMMTransitioner)

__all__ = ["HmmCell", "MsaHmmCell", "MsaHmmLayer", "BaseRNN", "Bidirectional", "TotalProbabilityCell", "DenseEmitter",
           "DenseTransitioner", "ProfileHMMEmitter", "ProfileHMMTransitioner", "GenePredHMMEmitter", "SimpleGenePredHMMEmitter",
           "GenePredHMMTransitioner", "GenePredMultiHMMTransitioner", "SimpleGenePredHMMTransitioner"]
