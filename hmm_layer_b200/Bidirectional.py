"""Bidirectional wrapper with the reference's interface (reference: hmm_layer/Bidirectional.py:6-180): runs the
forward layer on the sequence and the backward layer on the flipped sequence, flips the backward output back and
merges.  The backward layer must NOT itself be built with reverse=True (the reference's test script does that and
scans the middle segment left-to-right, defect D3)."""
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
            raise ValueError("backward_layer must not be built with reverse=True: Bidirectional already flips the sequence")
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
        bwd = self.backward_layer(torch.flip(sequences, [axis]), backward_state)
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
