"""Bidirectional wrapper with the reference's interface (reference: hmm_layer/Bidirectional.py:6-180): runs the
forward layer on the sequence and the backward layer on the flipped sequence, flips the backward output back and
merges.  The reference's own script builds the backward layer with reverse=True
(tests/parallel_rnn_forward.py:57-63), which in the reference flips the sequence twice and scans left-to-right
(defect D3).  That wiring is accepted here (one warning): such a layer flips the sequence itself, so this wrapper
does not flip it a second time and the scan runs right-to-left, as the script's author intended."""
import warnings

import torch
import torch.nn as nn


class Bidirectional(nn.Module):
    def __init__(self, layer, backward_layer, merge_mode="concat"):
        super().__init__()
        if not isinstance(layer, nn.Module):
            raise ValueError(f"Please initialize `Bidirectional` layer with a `nn.Module` instance. Received: {layer}")
        if backward_layer is not None and not isinstance(backward_layer, nn.Module):
            raise ValueError(f"`backward_layer` need to be a `nn.Module` instance. Received: {backward_layer}")
        if merge_mode not in ["sum", "concat", None]:
            raise ValueError(f"Invalid merge mode. Received: {merge_mode}. Merge mode should be one of "
                             '{"sum", "concat", None}')
        if backward_layer is not None and getattr(backward_layer, "reverse", False):
            warnings.warn("Bidirectional: backward_layer was built with reverse=True; it flips the sequence itself, so "
                          "Bidirectional will not flip it again (the reference flips twice, defect D3)", stacklevel=2)
        if backward_layer is not None:
            for attr in ("batch_first", "hidden_size"):   # reference :100-110
                fv, bv = getattr(layer, attr, None), getattr(backward_layer, attr, None)
                if fv != bv:
                    raise ValueError("Forward layer and backward layer are expected to have the same value for attribute "
                                     f'"{attr}", got "{fv}" for forward layer and "{bv}" for backward layer')
        self.forward_layer = layer
        self.backward_layer = backward_layer
        self.merge_mode = merge_mode
        self.return_sequences = bool(getattr(layer, "return_sequences", False))
        self.return_state = bool(getattr(layer, "return_state", False))

    def forward(self, sequences, initial_state=None, **kwargs):
        if initial_state is not None:
            half = len(initial_state) // 2
            forward_state, backward_state = list(initial_state[:half]), list(initial_state[half:])
        else:
            forward_state, backward_state = None, None
        axis = 1 if self.backward_layer.batch_first else 0
        fwd = self.forward_layer(sequences, forward_state)
        self_flipping = bool(getattr(self.backward_layer, "reverse", False))
        bwd = self.backward_layer(sequences if self_flipping else torch.flip(sequences, [axis]), backward_state)
        if self.return_state:
            forward_output, forward_states = fwd
            backward_output, backward_states = bwd
        else:
            forward_output, backward_output = fwd, bwd
        backward_output = torch.flip(backward_output, [axis])
        if self.merge_mode == "concat":
            output = torch.cat([forward_output, backward_output], dim=-1)
        elif self.merge_mode == "sum":
            output = forward_output + backward_output
        else:
            output = [forward_output, backward_output]
        if self.return_state:
            return (output, *forward_states, *backward_states)
        return output
